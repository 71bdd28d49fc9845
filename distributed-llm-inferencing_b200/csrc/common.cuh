// Shared device/host helpers for the dli_b200 kernels (sm_100a only).
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include <mutex>

#include "../../include/dli.h"

namespace dli {

// ---- host-side error plumbing (abi.cu owns the storage) -------------------
void set_error(const char* fmt, ...);
int fail(int code, const char* fmt, ...);

#define DLI_CHECK_ARG(cond, ...)                                  \
  do {                                                            \
    if (!(cond)) return ::dli::fail(DLI_ERR_INVALID, __VA_ARGS__); \
  } while (0)

#define DLI_CUDA_LAUNCH_CHECK(what)                                             \
  do {                                                                          \
    cudaError_t e__ = cudaGetLastError();                                       \
    if (e__ != cudaSuccess)                                                     \
      return ::dli::fail(DLI_ERR_CUDA, "%s: %s", what, cudaGetErrorString(e__)); \
  } while (0)

#define DLI_CUDA_CALL(expr)                                                       \
  do {                                                                            \
    cudaError_t e__ = (expr);                                                     \
    if (e__ != cudaSuccess)                                                       \
      return ::dli::fail(DLI_ERR_CUDA, "%s: %s", #expr, cudaGetErrorString(e__)); \
  } while (0)

inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }
int sm_count_cached();
int current_device_slot();  // abi.cu: ordinal of the current device (0..63), or -1 after recording the error

// cudaFuncSetAttribute and occupancy results belong to a (kernel, DEVICE) pair, so a launcher keeps one
// of these per kernel instantiation and repeats its set-up the first time each device calls it (a host
// that drives several GPUs from one process would otherwise fail to launch on the second one).
constexpr int kMaxDevices = 64;
struct PerDeviceOnce {
  std::mutex mu;
  bool done[kMaxDevices] = {};
  int value[kMaxDevices] = {};      // optional payload (e.g. resident CTAs per SM)
  size_t bytes[kMaxDevices] = {};   // optional payload (e.g. dynamic shared memory the set-up covered)
};
#define DLI_ONCE_PER_DEVICE(state, dev_var, ...)                    \
  const int dev_var = ::dli::current_device_slot();                 \
  if (dev_var < 0) return DLI_ERR_CUDA;                             \
  {                                                                 \
    std::lock_guard<std::mutex> lock__((state).mu);                 \
    if (!(state).done[dev_var]) {                                   \
      __VA_ARGS__;                                                  \
      (state).done[dev_var] = true;                                 \
    }                                                               \
  }
bool pdl_enabled();  // abi.cu: programmatic dependent launch on/off (dli_set_pdl)

// Every kernel of the library is launched through this helper.  With PDL enabled the
// launch carries cudaLaunchAttributeProgrammaticStreamSerialization, so the kernel may
// start while its predecessor in the stream is still draining; each kernel therefore
// calls pdl_wait() before it touches anything a predecessor may have written.
void prepare_kernel(const void* fn);  // abi.cu: one-time per-kernel attributes

template <typename... KArgs, typename... Args>
inline cudaError_t launch_kernel(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream,
                                 Args... args) {
  prepare_kernel(reinterpret_cast<const void*>(kernel));
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

// Same, as thread-block clusters of `cluster_x` CTAs along x (distributed shared memory).
template <typename... KArgs, typename... Args>
inline cudaError_t launch_kernel_cluster(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem,
                                         cudaStream_t stream, unsigned cluster_x, Args... args) {
  prepare_kernel(reinterpret_cast<const void*>(kernel));
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = cluster_x;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 2 : 1;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

#define DLI_LAUNCH(what, ...)                                                        \
  do {                                                                               \
    cudaError_t e__ = ::dli::launch_kernel(__VA_ARGS__);                             \
    if (e__ != cudaSuccess)                                                          \
      return ::dli::fail(DLI_ERR_CUDA, "%s: %s", what, cudaGetErrorString(e__));    \
  } while (0)

// ---- device helpers ----------------------------------------------------------
typedef __nv_bfloat16 bf16;
typedef __nv_bfloat162 bf162;

struct __align__(16) bf16x8 {
  bf162 v[4];
};

// Programmatic dependent launch: block until the preceding kernel of the stream has
// completed and its writes are visible (no-op when launched without the attribute) ...
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
// ... and allow the next kernel of the stream to start launching.
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

// 2^x as ONE MUFU.EX2: exp2f() without -ftz wraps the instruction in a range test and two predicated multiplies
// so that results below 2^-126 come out as denormals (4 issue slots per call); softmax weights that small
// contribute nothing, so the flash-attention loops flush them to zero instead.
__device__ __forceinline__ float ex2_ftz(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// 128-bit streaming load that does not pollute L1 (weights / KV are read once).
__device__ __forceinline__ uint4 ld_stream_16B(const void* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}
__device__ __forceinline__ uint4 ld_16B(const void* p) {
  return *reinterpret_cast<const uint4*>(p);
}
__device__ __forceinline__ void st_16B(void* p, uint4 v) { *reinterpret_cast<uint4*>(p) = v; }

// bf16 pair packed in a u32 -> two floats (exact: bf16 is the top half of fp32).
__device__ __forceinline__ float bf_lo(uint32_t u) { return __uint_as_float(u << 16); }
__device__ __forceinline__ float bf_hi(uint32_t u) { return __uint_as_float(u & 0xffff0000u); }
__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi) {
  bf162 t = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&t);
}
__device__ __forceinline__ float round_bf16(float x) {
  return __bfloat162float(__float2bfloat16_rn(x));
}
__device__ __forceinline__ void unpack8(const uint4& u, float* f) {
  f[0] = bf_lo(u.x); f[1] = bf_hi(u.x); f[2] = bf_lo(u.y); f[3] = bf_hi(u.y);
  f[4] = bf_lo(u.z); f[5] = bf_hi(u.z); f[6] = bf_lo(u.w); f[7] = bf_hi(u.w);
}
__device__ __forceinline__ uint4 pack8(const float* f) {
  uint4 u;
  u.x = pack_bf16(f[0], f[1]); u.y = pack_bf16(f[2], f[3]);
  u.z = pack_bf16(f[4], f[5]); u.w = pack_bf16(f[6], f[7]);
  return u;
}

// Activations, evaluated the way torch evaluates them on a bf16 tensor:
// upcast to fp32, compute, round once to bf16 (the caller rounds).
__device__ __forceinline__ float silu_f(float x) { return x / (1.0f + expf(-x)); }
// gelu_new (tanh approximation), TF:activations.py:59-66
__device__ __forceinline__ float gelu_new_f(float x) {
  const float k = 0.7978845608028654f;  // sqrt(2/pi)
  return 0.5f * x * (1.0f + tanhf(k * (x + 0.044715f * x * x * x)));
}

}  // namespace dli
