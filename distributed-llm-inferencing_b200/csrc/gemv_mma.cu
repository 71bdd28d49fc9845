// Weight-streaming projection for a FEW activation rows (3 <= M <= 16) that needs NO scratch: the
// loop structure of the CUDA-core GEMV (gemv.cu) with the multiply-adds moved to mma.sync, so the
// kernel stays an HBM stream where the scalar GEMV becomes FMA-issue bound (4.1-5.0 TB/s at 3-4
// rows against 6.0-7.3 at 1-2).  Measured 4.2-5.5 TB/s (8B shapes) and 5.0-6.3 TB/s (70B shapes)
// at 3-8 rows.  dli_linear_bf16 uses it when it has no split-K workspace and the matrix has too
// few 128-row tiles to fill the GPU with the swap-AB tcgen05 kernel (that kernel: 1.8 TB/s on
// 4096 x 14336).  Inside the decode step the engine passes a workspace and the split-K tcgen05
// stream is faster (profiles/r1_notes.md), so this kernel is not on that path.
//
//   y[m, n] = epi( sum_k x[m,k] * w[n,k] ),   w in the HF Linear layout [N, K]
//
// Orientation: x is the A operand (16 rows, rows >= M are zero registers), 8 weight rows are
// the B operand of an m16n8k16 MMA.  Both come straight from global memory into fragments: the
// contraction index of an MMA may be permuted freely as long as A and B agree, so lane
// (g = lane/4, t = lane%4) takes ONE 16-byte load of w[n0+g][k0+8t .. k0+8t+8) per 32-element
// chunk and uses elements 0-3 for one MMA and 4-7 for the next; x[g][same 8 elements] (and
// x[g+8] when M > 8) supplies the matching A registers from L1/L2 (x is at most 16 x K bf16).
// A task is 8 weight rows (gated: 4 gate rows + their 4 up rows); a CTA's 8 warps interleave
// 256-element groups of K, so one CTA iteration reads 8 contiguous 4 KB row segments, with 8
// loads of 16 B per thread in flight like the GEMV.  Partial tiles are combined through shared
// memory in warp order (deterministic).  Algorithmic bytes per launch: N*K*2.
#include <stdlib.h>

#include "mma_utils.cuh"

namespace dli {

constexpr int kMmaGemvThreads = 256;
constexpr int kMmaGemvWarps = 8;

__device__ __forceinline__ size_t gated_row_mma(int n, int is_up) {
  return (size_t)(n >> 7) * 256 + (n & 127) + (is_up ? 128 : 0);
}

template <bool kGated, bool kTwoTiles>
__global__ void __launch_bounds__(kMmaGemvThreads, 3)
gemv_mma_kernel(const bf16* __restrict__ x, const bf16* __restrict__ w, const bf16* __restrict__ bias,
                const bf16* residual, bf16* y, int M, int N, int K, int epilogue, int ntasks, int pf_l2) {
  // partial output tiles: [parity][warp][activation row 0..15][weight row 0..7]
  __shared__ float part[2][kMmaGemvWarps][16][8];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int g = lane >> 2, t = lane & 3;
  constexpr int kOutPerTask = kGated ? 4 : 8;
  const int n_out = N;  // gated: N = F output columns, w has 2F rows
  const uint4 zero = make_uint4(0u, 0u, 0u, 0u);

  auto weight_row = [&](int task, int r) -> const bf16* {
    size_t row;
    if (kGated) {
      const int n = min(task * 4 + (r & 3), N - 1);
      row = gated_row_mma(n, r >> 2);
    } else {
      row = (size_t)min(task * 8 + r, N - 1);
    }
    return w + row * (size_t)K;
  };

  if (pf_l2 > 0 && (int)blockIdx.x < ntasks && tid < 8) {
    // before the dependency wait: head of this CTA's first 8 weight rows into L2 (see gemv.cu)
    const uint32_t bytes = (uint32_t)min((size_t)K * 2, (size_t)pf_l2) & ~15u;
    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(weight_row(blockIdx.x, tid)), "r"(bytes) : "memory");
  }
  pdl_wait();
  pdl_trigger();

  const bf16* xa_row = x + (size_t)(g < M ? g : 0) * K;
  const bf16* xb_row = x + (size_t)(g + 8 < M ? g + 8 : 0) * K;
  const bool a_on = g < M, b_on = kTwoTiles && (g + 8 < M);
  const int ngroups = (K + 255) >> 8;

  int parity = 0;
  for (int task = blockIdx.x; task < ntasks; task += gridDim.x, parity ^= 1) {
    const bf16* wr = weight_row(task, g);
    float c[4] = {0.f, 0.f, 0.f, 0.f};
    for (int gi = warp; gi < ngroups; gi += kMmaGemvWarps) {
      const int kbase = (gi << 8) + 8 * t;
      uint4 wv[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const int k = kbase + 32 * i;
        wv[i] = (k + 8 <= K) ? ld_stream_16B(wr + k) : zero;
      }
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const int k = kbase + 32 * i;
        const bool in = k + 8 <= K;
        const uint4 xa = (a_on && in) ? ld_16B(xa_row + k) : zero;
        const uint4 xb = (b_on && in) ? ld_16B(xb_row + k) : zero;
        const uint32_t a_lo[4] = {xa.x, xb.x, xa.y, xb.y};
        const uint32_t a_hi[4] = {xa.z, xb.z, xa.w, xb.w};
        mma_16816(c, a_lo, wv[i].x, wv[i].y);
        mma_16816(c, a_hi, wv[i].z, wv[i].w);
      }
    }
    // c0,c1: activation row g, weight rows 2t,2t+1; c2,c3: activation row g+8
    float(*pw)[8] = part[parity][warp];
    pw[g][2 * t] = c[0];
    pw[g][2 * t + 1] = c[1];
    if (kTwoTiles) {
      pw[g + 8][2 * t] = c[2];
      pw[g + 8][2 * t + 1] = c[3];
    }
    __syncthreads();

    const int n0 = task * kOutPerTask;
    for (int id = tid; id < M * kOutPerTask; id += kMmaGemvThreads) {
      const int m = id / kOutPerTask, j = id - m * kOutPerTask;
      float s = 0.f, s2 = 0.f;
#pragma unroll
      for (int ww = 0; ww < kMmaGemvWarps; ++ww) {
        s += part[parity][ww][m][j];
        if (kGated) s2 += part[parity][ww][m][j + 4];
      }
      const int n = n0 + j;
      if (n < n_out) {
        float v;
        if (kGated) {
          const float gte = round_bf16(s), up = round_bf16(s2);
          v = round_bf16(silu_f(gte)) * up;
        } else {
          if (epilogue & DLI_EPI_BIAS) s += __bfloat162float(bias[n]);
          v = round_bf16(s);
          if (epilogue & DLI_EPI_GELU_NEW) v = round_bf16(gelu_new_f(v));
          if (epilogue & DLI_EPI_RESIDUAL) v += __bfloat162float(residual[(size_t)m * n_out + n]);
        }
        y[(size_t)m * n_out + n] = __float2bfloat16_rn(v);
      }
    }
    // the other parity buffer is only rewritten after the next __syncthreads()
  }
}

template <bool kGated, bool kTwoTiles>
static int launch_gemv_mma(const bf16* x, const bf16* w, const bf16* bias, const bf16* residual, bf16* y, int M, int N,
                           int K, int epilogue, cudaStream_t s) {
  constexpr int kOutPerTask = kGated ? 4 : 8;
  const int ntasks = (N + kOutPerTask - 1) / kOutPerTask;
  const int sms = sm_count_cached();
  if (sms <= 0) return fail(DLI_ERR_CUDA, "gemv_mma: no CUDA device");
  static PerDeviceOnce once;
  DLI_ONCE_PER_DEVICE(once, dev, {
    int o = 0;
    DLI_CUDA_CALL(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&o, gemv_mma_kernel<kGated, kTwoTiles>,
                                                                kMmaGemvThreads, 0));
    once.value[dev] = o > 0 ? o : 1;
  });
  const int occ = once.value[dev];
  const int grid = ntasks < sms * occ ? ntasks : sms * occ;
  static const int pf_l2 = getenv("DLI_GEMV_PF_L2") ? atoi(getenv("DLI_GEMV_PF_L2")) : 8192;
  DLI_LAUNCH("gemv_mma_kernel", gemv_mma_kernel<kGated, kTwoTiles>, dim3(grid), dim3(kMmaGemvThreads), (size_t)0, s, x,
             w, bias, residual, y, M, N, K, epilogue, ntasks, pf_l2);
  return DLI_OK;
}

}  // namespace dli

extern "C" {

int dli_gemv_mma_bf16(const void* x, const void* w, const void* bias, const void* residual, void* y, int M, int N,
                      int K, int epilogue, dli_stream_t stream) {
  using namespace dli;
  DLI_CHECK_ARG(x && w && y, "dli_gemv_mma_bf16: null pointer");
  DLI_CHECK_ARG(M >= 0 && M <= 16 && N > 0 && K > 0 && K % 8 == 0,
                "dli_gemv_mma_bf16: need M <= 16 and K %% 8 == 0 (got M=%d N=%d K=%d)", M, N, K);
  DLI_CHECK_ARG(aligned16(x) && aligned16(w), "dli_gemv_mma_bf16: x and w must be 16-byte aligned");
  DLI_CHECK_ARG(!(epilogue & DLI_EPI_BIAS) || bias, "dli_gemv_mma_bf16: DLI_EPI_BIAS without bias");
  DLI_CHECK_ARG(!(epilogue & DLI_EPI_RESIDUAL) || residual, "dli_gemv_mma_bf16: DLI_EPI_RESIDUAL without residual");
  if (M == 0) return DLI_OK;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (M <= 8)
    return launch_gemv_mma<false, false>((const bf16*)x, (const bf16*)w, (const bf16*)bias, (const bf16*)residual,
                                         (bf16*)y, M, N, K, epilogue, s);
  return launch_gemv_mma<false, true>((const bf16*)x, (const bf16*)w, (const bf16*)bias, (const bf16*)residual,
                                      (bf16*)y, M, N, K, epilogue, s);
}

int dli_gemv_mma_gated_silu_bf16(const void* x, const void* w_gu, void* y, int M, int F, int K, dli_stream_t stream) {
  using namespace dli;
  DLI_CHECK_ARG(x && w_gu && y, "dli_gemv_mma_gated_silu_bf16: null pointer");
  DLI_CHECK_ARG(M >= 0 && M <= 16 && F > 0 && F % 128 == 0 && K > 0 && K % 8 == 0,
                "dli_gemv_mma_gated_silu_bf16: need M <= 16, F %% 128 == 0, K %% 8 == 0 (got M=%d F=%d K=%d)", M, F, K);
  DLI_CHECK_ARG(aligned16(x) && aligned16(w_gu), "dli_gemv_mma_gated_silu_bf16: x and w_gu must be 16-byte aligned");
  if (M == 0) return DLI_OK;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (M <= 8)
    return launch_gemv_mma<true, false>((const bf16*)x, (const bf16*)w_gu, nullptr, nullptr, (bf16*)y, M, F, K, 0, s);
  return launch_gemv_mma<true, true>((const bf16*)x, (const bf16*)w_gu, nullptr, nullptr, (bf16*)y, M, F, K, 0, s);
}

}  // extern "C"
