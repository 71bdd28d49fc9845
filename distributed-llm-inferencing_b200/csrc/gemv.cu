// Weight-streaming GEMV for decode (M <= 4 rows of x per pass), CUDA cores.
//
// y[m, n] = epi( sum_k x[m,k] * w[n,k] ),  w in the HF Linear layout [N, K].
//
// HBM-bound by design: every weight byte is read exactly once per pass with
// 128-bit streaming loads (ld.global.nc.L1::no_allocate); algorithmic bytes per
// launch = N*K*2 (+ M*K*2 for x, + M*N*2 for y; both negligible).  A CTA works
// on kRows weight rows at a time; its 8 warps split K between them so that one
// CTA iteration reads kRows contiguous 4 KB row segments (256 lanes x 16 B),
// then partial sums are combined through shared memory (one warp per output).
// Tasks (row groups) are dealt round-robin to a grid sized as a multiple of the
// SM count, so every SM streams the same number of bytes.
#include "common.cuh"

namespace dli {

constexpr int kGemvThreads = 256;
constexpr int kGemvWarps = kGemvThreads / 32;
constexpr int kGemvRows = 4;  // weight rows per task

// Row index of logical row `n` of the gated [gate|up] matrix, interleaved in
// blocks of 128 rows (see dli.h: dli_gemv_gated_silu_bf16).
__device__ __forceinline__ size_t gated_row(int n, int is_up) {
  return (size_t)(n >> 7) * 256 + (n & 127) + (is_up ? 128 : 0);
}

template <int M, bool kGated>
__global__ void __launch_bounds__(kGemvThreads)
gemv_kernel(const bf16* __restrict__ x, const bf16* __restrict__ w,
            const bf16* __restrict__ bias, const bf16* residual, bf16* y, int N, int K,
            int epilogue, int ntasks) {
  // partial sums: [parity][value][thread]; value = row * M + m
  __shared__ float part[2][kGemvRows * M][kGemvThreads];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int kvec = K >> 3;
  const int iters = (kvec + kGemvThreads - 1) / kGemvThreads;
  // gated: a task is 2 output columns (2 gate rows + their 2 up rows)
  constexpr int kOutPerTask = kGated ? kGemvRows / 2 : kGemvRows;
  const uint4* xv = reinterpret_cast<const uint4*>(x);

  int parity = 0;
  for (int task = blockIdx.x; task < ntasks; task += gridDim.x, parity ^= 1) {
    const int n0 = task * kOutPerTask;
    const uint4* wrow[kGemvRows];
#pragma unroll
    for (int r = 0; r < kGemvRows; ++r) {
      size_t row;
      if (kGated) {
        const int n = min(n0 + (r % kOutPerTask), N - 1);
        row = gated_row(n, r / kOutPerTask);
      } else {
        row = (size_t)min(n0 + r, N - 1);
      }
      wrow[r] = reinterpret_cast<const uint4*>(w + row * (size_t)K);
    }
    float acc[kGemvRows][M];
#pragma unroll
    for (int r = 0; r < kGemvRows; ++r)
#pragma unroll
      for (int m = 0; m < M; ++m) acc[r][m] = 0.f;

    for (int j = 0; j < iters; j += 2) {
      const int i0 = tid + j * kGemvThreads, i1 = i0 + kGemvThreads;
      const bool p0 = i0 < kvec, p1 = i1 < kvec;
      uint4 wv0[kGemvRows], wv1[kGemvRows];
      const uint4 zero = make_uint4(0, 0, 0, 0);
#pragma unroll
      for (int r = 0; r < kGemvRows; ++r) {
        wv0[r] = p0 ? ld_stream_16B(wrow[r] + i0) : zero;
        wv1[r] = p1 ? ld_stream_16B(wrow[r] + i1) : zero;
      }
      uint4 xa[M], xb[M];
#pragma unroll
      for (int m = 0; m < M; ++m) {
        xa[m] = p0 ? ld_16B(xv + (size_t)m * kvec + i0) : zero;
        xb[m] = p1 ? ld_16B(xv + (size_t)m * kvec + i1) : zero;
      }
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        float xf[M][8];
#pragma unroll
        for (int m = 0; m < M; ++m) unpack8(h ? xb[m] : xa[m], xf[m]);
#pragma unroll
        for (int r = 0; r < kGemvRows; ++r) {
          float wf[8];
          unpack8(h ? wv1[r] : wv0[r], wf);
#pragma unroll
          for (int m = 0; m < M; ++m)
#pragma unroll
            for (int e = 0; e < 8; ++e) acc[r][m] = fmaf(wf[e], xf[m][e], acc[r][m]);
        }
      }
    }

#pragma unroll
    for (int r = 0; r < kGemvRows; ++r)
#pragma unroll
      for (int m = 0; m < M; ++m) part[parity][r * M + m][tid] = acc[r][m];
    __syncthreads();

    // one warp per output element (col, m): sum 256 partials in a fixed order
    for (int o = warp; o < kOutPerTask * M; o += kGemvWarps) {
      const int c = o / M, m = o - c * M;
      float s = 0.f, s2 = 0.f;
#pragma unroll
      for (int ww = 0; ww < kGemvWarps; ++ww) {
        s += part[parity][c * M + m][ww * 32 + lane];
        if (kGated) s2 += part[parity][(c + kOutPerTask) * M + m][ww * 32 + lane];
      }
      s = warp_sum(s);
      if (kGated) s2 = warp_sum(s2);
      const int n = n0 + c;
      if (lane == 0 && n < N) {
        float v;
        if (kGated) {
          const float g = round_bf16(s), u = round_bf16(s2);
          v = round_bf16(silu_f(g)) * u;
        } else {
          if (epilogue & DLI_EPI_BIAS) s += __bfloat162float(bias[n]);
          v = round_bf16(s);
          if (epilogue & DLI_EPI_GELU_NEW) v = round_bf16(gelu_new_f(v));
          if (epilogue & DLI_EPI_RESIDUAL) v += __bfloat162float(residual[(size_t)m * N + n]);
        }
        y[(size_t)m * N + n] = __float2bfloat16_rn(v);
      }
    }
    // the other parity buffer is only rewritten after the next __syncthreads()
  }
}

template <int M, bool kGated>
static int launch_gemv(const bf16* x, const bf16* w, const bf16* bias, const bf16* residual,
                       bf16* y, int N, int K, int epilogue, cudaStream_t s) {
  constexpr int kOutPerTask = kGated ? kGemvRows / 2 : kGemvRows;
  const int ntasks = (N + kOutPerTask - 1) / kOutPerTask;
  const int sms = sm_count_cached();
  if (sms <= 0) return fail(DLI_ERR_CUDA, "gemv: no CUDA device");
  static int occ = 0;  // resident CTAs per SM for this instantiation
  if (occ == 0) {
    int o = 0;
    DLI_CUDA_CALL(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&o, gemv_kernel<M, kGated>,
                                                                kGemvThreads, 0));
    occ = o > 0 ? o : 1;
  }
  const int grid = ntasks < sms * occ ? ntasks : sms * occ;
  gemv_kernel<M, kGated><<<grid, kGemvThreads, 0, s>>>(x, w, bias, residual, y, N, K, epilogue, ntasks);
  DLI_CUDA_LAUNCH_CHECK("gemv_kernel");
  return DLI_OK;
}

template <bool kGated>
static int gemv_dispatch(const void* x, const void* w, const void* bias, const void* residual,
                         void* y, int M, int N, int K, int epilogue, dli_stream_t stream) {
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  // M > 4 is served in passes of <= 4 rows (weights re-streamed once per pass).
  for (int m0 = 0; m0 < M; m0 += 4) {
    const int mm = (M - m0) < 4 ? (M - m0) : 4;
    const bf16* xp = (const bf16*)x + (size_t)m0 * K;
    const bf16* rp = residual ? (const bf16*)residual + (size_t)m0 * N : nullptr;
    bf16* yp = (bf16*)y + (size_t)m0 * N;
    int rc;
    switch (mm) {
      case 1: rc = launch_gemv<1, kGated>(xp, (const bf16*)w, (const bf16*)bias, rp, yp, N, K, epilogue, s); break;
      case 2: rc = launch_gemv<2, kGated>(xp, (const bf16*)w, (const bf16*)bias, rp, yp, N, K, epilogue, s); break;
      case 3: rc = launch_gemv<3, kGated>(xp, (const bf16*)w, (const bf16*)bias, rp, yp, N, K, epilogue, s); break;
      default: rc = launch_gemv<4, kGated>(xp, (const bf16*)w, (const bf16*)bias, rp, yp, N, K, epilogue, s); break;
    }
    if (rc != DLI_OK) return rc;
  }
  return DLI_OK;
}

}  // namespace dli

extern "C" {

int dli_gemv_bf16(const void* x, const void* w, const void* bias, const void* residual, void* y,
                  int M, int N, int K, int epilogue, dli_stream_t stream) {
  using namespace dli;
  DLI_CHECK_ARG(x && w && y, "dli_gemv_bf16: null pointer");
  DLI_CHECK_ARG(M >= 0 && N > 0 && K > 0 && K % 8 == 0,
                "dli_gemv_bf16: need K %% 8 == 0 (got M=%d N=%d K=%d)", M, N, K);
  DLI_CHECK_ARG(!(epilogue & DLI_EPI_BIAS) || bias, "dli_gemv_bf16: DLI_EPI_BIAS without bias");
  DLI_CHECK_ARG(!(epilogue & DLI_EPI_RESIDUAL) || residual,
                "dli_gemv_bf16: DLI_EPI_RESIDUAL without residual");
  DLI_CHECK_ARG(aligned16(x) && aligned16(w), "dli_gemv_bf16: x and w must be 16-byte aligned");
  if (M == 0) return DLI_OK;
  return gemv_dispatch<false>(x, w, bias, residual, y, M, N, K, epilogue, stream);
}

int dli_gemv_gated_silu_bf16(const void* x, const void* w_gu, void* y, int M, int F, int K,
                             dli_stream_t stream) {
  using namespace dli;
  DLI_CHECK_ARG(x && w_gu && y, "dli_gemv_gated_silu_bf16: null pointer");
  DLI_CHECK_ARG(M >= 0 && F > 0 && F % 128 == 0 && K > 0 && K % 8 == 0,
                "dli_gemv_gated_silu_bf16: need F %% 128 == 0, K %% 8 == 0 (got M=%d F=%d K=%d)", M, F, K);
  DLI_CHECK_ARG(aligned16(x) && aligned16(w_gu), "dli_gemv_gated_silu_bf16: pointers must be 16-byte aligned");
  if (M == 0) return DLI_OK;
  return gemv_dispatch<true>(x, w_gu, nullptr, nullptr, y, M, F, K, 0, stream);
}

}  // extern "C"
