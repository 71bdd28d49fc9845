// Library-level entry points: error reporting, version, device query.
#include <stdarg.h>
#include <stdlib.h>
#include <string.h>

#include <atomic>
#include <mutex>
#include <unordered_set>

#include "common.cuh"

namespace dli {

static thread_local char g_err[512] = {0};

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

void set_pdl(bool on);

int fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return code;
}

void prepare_kernel(const void*) {}  // hook for one-time per-kernel attributes (none needed today)

static std::atomic<bool> g_pdl{false};
bool pdl_enabled() { return g_pdl.load(std::memory_order_relaxed); }
void set_pdl(bool on) { g_pdl.store(on, std::memory_order_relaxed); }

int current_device_slot() {
  int dev = -1;
  const cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) {
    fail(DLI_ERR_CUDA, "cudaGetDevice: %s", cudaGetErrorString(e));
    return -1;
  }
  if (dev < 0 || dev >= kMaxDevices) {
    fail(DLI_ERR_CUDA, "device ordinal %d is outside the %d devices this library keeps per-device state for", dev,
         kMaxDevices);
    return -1;
  }
  return dev;
}

int sm_count_cached() {
  static std::atomic<int> cached[kMaxDevices];
  const int dev = current_device_slot();
  if (dev < 0) return -1;
  int n = cached[dev].load(std::memory_order_relaxed);
  if (n == 0) {
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return -1;
    cached[dev].store(n, std::memory_order_relaxed);
  }
  return n;
}

}  // namespace dli

extern "C" {

const char* dli_last_error(void) { return dli::g_err; }

int dli_version(void) { return 1004; }  // 1004: + dli_gemm_qkv_rope_append_bf16 (1003: + dli_gemv_argmax_bf16; per-device kernel set-up)

int dli_set_pdl(int enable) {
  const int before = dli::pdl_enabled() ? 1 : 0;
  dli::set_pdl(enable != 0);
  return before;
}

int dli_sm_count(void) {
  int n = dli::sm_count_cached();
  if (n <= 0) {
    cudaGetLastError();
    return dli::fail(DLI_ERR_CUDA, "dli_sm_count: no usable CUDA device");
  }
  return n;
}

}  // extern "C"
