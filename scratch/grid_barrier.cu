// Micro-benchmark: cost of a software grid barrier (cooperative launch) vs a PDL kernel boundary.
#include <cooperative_groups.h>
#include <cuda_runtime.h>
#include <cstdio>
namespace cg = cooperative_groups;

__device__ __forceinline__ unsigned ld_acquire(const unsigned* p) {
  unsigned v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

__global__ void barrier_loop(unsigned* ctr, int iters, float* sink) {
  const unsigned n = gridDim.x;
  float acc = 0.f;
  for (int it = 0; it < iters; ++it) {
    acc += 1.f;
    __syncthreads();
    if (threadIdx.x == 0) {
      __threadfence();
      atomicAdd(ctr, 1u);
      const unsigned target = (unsigned)(it + 1) * n;
      while (ld_acquire(ctr) < target) {}
    }
    __syncthreads();
  }
  if (acc < 0) sink[0] = acc;
}

__global__ void cg_barrier_loop(int iters, float* sink) {
  cg::grid_group g = cg::this_grid();
  float acc = 0.f;
  for (int it = 0; it < iters; ++it) { acc += 1.f; g.sync(); }
  if (acc < 0) sink[0] = acc;
}

__global__ void empty_kernel(float* sink) {
  asm volatile("griddepcontrol.wait;" ::: "memory");
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  if (sink == nullptr) return;
}

int main() {
  unsigned* ctr; float* sink;
  cudaMalloc(&ctr, 4); cudaMalloc(&sink, 4);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  const int iters = 2000;
  for (int cfg = 0; cfg < 3; ++cfg) {
    int ctas = cfg == 0 ? 148 : cfg == 1 ? 296 : 444, thr = 256;
    for (int rep = 0; rep < 2; ++rep) {
      cudaMemset(ctr, 0, 4);
      int it = iters; void* args[] = {&ctr, &it, &sink};
      cudaEventRecord(e0);
      cudaError_t e = cudaLaunchCooperativeKernel((void*)barrier_loop, dim3(ctas), dim3(thr), args, 0, 0);
      cudaEventRecord(e1); cudaEventSynchronize(e1);
      float ms; cudaEventElapsedTime(&ms, e0, e1);
      if (rep) printf("atomic barrier  ctas=%d: %.3f us/barrier (%s)\n", ctas, ms * 1e3 / iters, cudaGetErrorString(e));
    }
    for (int rep = 0; rep < 2; ++rep) {
      int it = iters; void* args[] = {&it, &sink};
      cudaEventRecord(e0);
      cudaError_t e = cudaLaunchCooperativeKernel((void*)cg_barrier_loop, dim3(ctas), dim3(thr), args, 0, 0);
      cudaEventRecord(e1); cudaEventSynchronize(e1);
      float ms; cudaEventElapsedTime(&ms, e0, e1);
      if (rep) printf("cg grid.sync    ctas=%d: %.3f us/barrier (%s)\n", ctas, ms * 1e3 / iters, cudaGetErrorString(e));
    }
  }
  // PDL chain of empty kernels, 444 CTAs x 256
  for (int pdl = 0; pdl < 2; ++pdl) {
    cudaLaunchConfig_t cfg = {}; cfg.gridDim = dim3(444); cfg.blockDim = dim3(256);
    cudaLaunchAttribute at; at.id = cudaLaunchAttributeProgrammaticStreamSerialization; at.val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = &at; cfg.numAttrs = pdl;
    for (int rep = 0; rep < 2; ++rep) {
      cudaEventRecord(e0);
      for (int i = 0; i < iters; ++i) cudaLaunchKernelEx(&cfg, empty_kernel, sink);
      cudaEventRecord(e1); cudaEventSynchronize(e1);
      float ms; cudaEventElapsedTime(&ms, e0, e1);
      if (rep) printf("empty kernel chain pdl=%d: %.3f us/launch (stream, not graph)\n", pdl, ms * 1e3 / iters);
    }
  }
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
