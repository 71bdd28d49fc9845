// Weight-streaming GEMV for decode (M <= 4 rows of x per pass), CUDA cores.
//
// y[m, n] = epi( sum_k xn[m,k] * w[n,k] ),  w in the HF Linear layout [N, K],
// xn = x, or (kNorm) xn = ln_w * bf16(x * rsqrt(mean(x^2) + eps)): the RMSNorm of
// the residual stream computed in the kernel prologue with HF's rounding points,
// so the decode step needs no separate norm kernel.
//
// HBM-bound by design: every weight byte is read exactly once per pass with
// 128-bit streaming loads (ld.global.nc.L1::no_allocate); algorithmic bytes per
// launch = N*K*2 (+ M*K*2 for x, + M*N*2 for y; both negligible).  A CTA works
// on kRows weight rows at a time; its 8 warps split K between them so that one
// CTA iteration reads kRows contiguous 4 KB row segments (256 lanes x 16 B),
// then partial sums are combined through shared memory (one warp per output).
// Groups of kRows rows are dealt round-robin to a grid sized as a multiple of the SM count,
// so at any moment the grid streams one contiguous band of the matrix (two balanced
// dealings were measured and lost, see `deal` in the kernel; so did a software-pipelined
// 2-row variant for the small matrices, profiles/r2_notes.md).
//
// Programmatic dependent launch: a CTA of the next kernel may move in while this
// kernel drains; it waits (pdl_wait) before reading activations.  Prefetching the
// first task's weights ahead of that wait (registers, or cp.async into shared
// memory) was measured and gave nothing: the shared-memory variant costs L1
// capacity that the streaming loads need (profiles/r1_notes.md).
#include <stdlib.h>

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

template <int M, bool kGated, bool kNorm>
__global__ void __launch_bounds__(kGemvThreads, M <= 2 ? 3 : 2)
gemv_kernel(const bf16* __restrict__ x, const bf16* __restrict__ w, const bf16* __restrict__ bias,
            const bf16* residual, bf16* y, int N, int K, int epilogue,
            const bf16* __restrict__ ln_w, float eps, int pf_l2, float* __restrict__ cand, int deal) {
  // partial sums: [parity][value][thread]; value = row * M + m
  __shared__ float part[2][kGemvRows * M][kGemvThreads];
  __shared__ float red[kGemvWarps][M];
  extern __shared__ __align__(16) uint8_t xnorm_raw[];  // kNorm: normalised x, [M][K] bf16
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int kvec = K >> 3;
  const int iters = (kvec + kGemvThreads - 1) / kGemvThreads;
  // gated: a task is 2 output columns (2 gate rows + their 2 up rows)
  constexpr int kOutPerTask = kGated ? kGemvRows / 2 : kGemvRows;
  const uint4 zero = make_uint4(0, 0, 0, 0);

  // How the N outputs are dealt to the CTAs (`deal`):
  //  0  (default) whole groups of kOutPerTask outputs round-robin (group i -> CTA i % grid): at any moment the grid
  //     streams one contiguous band of the matrix, but the last wave is ragged — on the small matrices a
  //     third of the CTAs carry one group more than the rest (o/down: 1024 groups on 444 CTAs = 3 vs 2);
  //  1  one contiguous, equally sized share per CTA (N/grid outputs, +-1), walked in groups of <= kOutPerTask
  //     (measured: 444 far-apart streams cost the gate/up matrix 10 % of its bandwidth, 35.2 -> 39.0 us);
  //  2  the full waves round-robin as in 0, then the outputs of the ragged last wave split evenly over ALL
  //     CTAs in groups of 1..kOutPerTask — the access pattern of 0 with the balance of 1.
  // Measured on the Llama-3-8B batch-1 step (B200, profiles/r2_notes.md): 0 -> 2.985 ms, 2 -> 3.062 ms,
  // 1 -> 3.171 ms.  The ragged wave costs less than short groups do (fewer loads in flight per thread), so
  // 0 stays the default; 1 and 2 remain selectable (DLI_GEMV_DEAL) to reproduce the comparison.
  const int full_waves = deal == 2 ? (N / kOutPerTask) / (int)gridDim.x : 0;  // waves of COMPLETE groups only
  const int tail_begin = full_waves * (int)gridDim.x * kOutPerTask;           // first output of the tail
  int n_lo = 0, n_hi = 0, ngroups;
  if (deal == 0) {
    const int groups_total = (N + kOutPerTask - 1) / kOutPerTask;
    ngroups = (groups_total - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;
    if (ngroups < 0) ngroups = 0;
  } else {
    // this CTA's contiguous share [n_lo, n_hi) of the outputs from tail_begin on (deal 1: of all outputs)
    n_lo = tail_begin + (int)((long long)(N - tail_begin) * blockIdx.x / gridDim.x);
    n_hi = tail_begin + (int)((long long)(N - tail_begin) * (blockIdx.x + 1) / gridDim.x);
    ngroups = full_waves + (n_hi - n_lo + kOutPerTask - 1) / kOutPerTask;
  }
  // group g of this CTA -> first output n0 and its number of outputs cnt (1..kOutPerTask)
  auto group = [&](int g, int& n0, int& cnt) {
    if (deal == 0 || g < full_waves) {
      n0 = (g * (int)gridDim.x + (int)blockIdx.x) * kOutPerTask;
      cnt = min(kOutPerTask, N - n0);
    } else {
      const int gs = ngroups - full_waves, gi = g - full_waves;  // sizes differ by at most one: 10 -> 4,3,3
      n0 = n_lo + (int)((long long)(n_hi - n_lo) * gi / gs);
      cnt = n_lo + (int)((long long)(n_hi - n_lo) * (gi + 1) / gs) - n0;
    }
  };

  // rows of group [n0, n0 + cnt): row r is live when its output index r % kOutPerTask < cnt
  auto row_ptrs = [&](int n0, int cnt, const uint4* (&wrow)[kGemvRows]) {
#pragma unroll
    for (int r = 0; r < kGemvRows; ++r) {
      size_t row;
      if (kGated) {
        const int n = n0 + min(r % kOutPerTask, cnt - 1);
        row = gated_row(n, r / kOutPerTask);
      } else {
        row = (size_t)(n0 + min(r, cnt - 1));
      }
      wrow[r] = reinterpret_cast<const uint4*>(w + row * (size_t)K);
    }
  };

  const uint4* wrow[kGemvRows];
  if (pf_l2 && ngroups > 0 && tid < kGemvRows) {
    // Before waiting for the previous kernel: pull the head of this CTA's first weight rows into L2
    // with one bulk prefetch per row (no registers, shared memory or L1 footprint), so HBM keeps
    // streaming through the kernel boundary.  Worth ~1.5 us per layer on the 4-GEMV chain.
    int n00, cnt0;
    group(0, n00, cnt0);
    row_ptrs(n00, cnt0, wrow);
    const uint32_t bytes = (uint32_t)min((size_t)K * 2, (size_t)pf_l2) & ~15u;
    const uint4* mine = wrow[0];
#pragma unroll
    for (int r = 1; r < kGemvRows; ++r)
      if (tid == r) mine = wrow[r];  // (no dynamic indexing: keeps wrow in registers)
    if (tid % kOutPerTask < cnt0)
      asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(mine), "r"(bytes) : "memory");
  }
  pdl_wait();
  pdl_trigger();  // the grid is resident at once (grid <= SMs x occupancy): successors may queue behind us

  const uint4* xv = reinterpret_cast<const uint4*>(x);
  if (kNorm) {
    // RMSNorm prologue (TF:models/llama/modeling_llama.py:62-67): fp32 mean of squares over
    // the row, hn = bf16(x * rstd), xn = bf16(ln_w * hn); result parked in shared memory.
    float ss[M];
#pragma unroll
    for (int m = 0; m < M; ++m) ss[m] = 0.f;
    for (int i = tid; i < kvec; i += kGemvThreads) {
#pragma unroll
      for (int m = 0; m < M; ++m) {
        float f[8];
        unpack8(ld_16B(xv + (size_t)m * kvec + i), f);
#pragma unroll
        for (int e = 0; e < 8; ++e) ss[m] += f[e] * f[e];
      }
    }
#pragma unroll
    for (int m = 0; m < M; ++m) {
      ss[m] = warp_sum(ss[m]);
      if (lane == 0) red[warp][m] = ss[m];
    }
    __syncthreads();
    float rstd[M];
#pragma unroll
    for (int m = 0; m < M; ++m) {
      float t = 0.f;
#pragma unroll
      for (int ww = 0; ww < kGemvWarps; ++ww) t += red[ww][m];
      rstd[m] = rsqrtf(t / (float)K + eps);
    }
    uint4* xs = reinterpret_cast<uint4*>(xnorm_raw);
    const uint4* lw = reinterpret_cast<const uint4*>(ln_w);
    for (int i = tid; i < kvec; i += kGemvThreads) {
      float g[8];
      unpack8(ld_16B(lw + i), g);
#pragma unroll
      for (int m = 0; m < M; ++m) {
        float f[8];
        unpack8(ld_16B(xv + (size_t)m * kvec + i), f);
#pragma unroll
        for (int e = 0; e < 8; ++e) f[e] = g[e] * round_bf16(f[e] * rstd[m]);
        xs[(size_t)m * kvec + i] = pack8(f);
      }
    }
    __syncthreads();
    xv = xs;  // generic pointer into shared memory
  }

  // fused greedy choice (lm_head): this CTA's best (bf16-rounded logit, output index) per x row,
  // kept by lane 0 of the warp that finishes each output and merged through `red` at the end
  float best_v[M];
  int best_i[M];
#pragma unroll
  for (int m = 0; m < M; ++m) { best_v[m] = -INFINITY; best_i[m] = 0x7fffffff; }

  int parity = 0;
  for (int g = 0; g < ngroups; ++g, parity ^= 1) {
    int n0, cnt;
    group(g, n0, cnt);
    row_ptrs(n0, cnt, wrow);
    bool live[kGemvRows];
#pragma unroll
    for (int r = 0; r < kGemvRows; ++r) live[r] = (r % kOutPerTask) < cnt;
    float acc[kGemvRows][M];
#pragma unroll
    for (int r = 0; r < kGemvRows; ++r)
#pragma unroll
      for (int m = 0; m < M; ++m) acc[r][m] = 0.f;

    for (int j = 0; j < iters; j += 2) {
      const int i0 = tid + j * kGemvThreads, i1 = i0 + kGemvThreads;
      const bool p0 = i0 < kvec, p1 = i1 < kvec;
      uint4 wv0[kGemvRows], wv1[kGemvRows];
#pragma unroll
      for (int r = 0; r < kGemvRows; ++r) {
        wv0[r] = (p0 && live[r]) ? ld_stream_16B(wrow[r] + i0) : zero;
        wv1[r] = (p1 && live[r]) ? ld_stream_16B(wrow[r] + i1) : zero;
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
      if (lane == 0 && c < cnt) {
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
        const bf16 yb = __float2bfloat16_rn(v);
        y[(size_t)m * N + n] = yb;
        if (cand) {  // outputs of a warp arrive in increasing n: strict > keeps the lowest index on ties
          const float yv = __bfloat162float(yb);
#pragma unroll
          for (int mm = 0; mm < M; ++mm)
            if (mm == m && (yv > best_v[mm] || best_i[mm] == 0x7fffffff)) { best_v[mm] = yv; best_i[mm] = n; }
        }
      }
    }
    // the other parity buffer is only rewritten after the next __syncthreads()
  }
  if (cand) {
    // per-CTA candidate -> cand[(cta * M + m) * 2 + {0,1}] = (logit, index bits); a CTA with no
    // outputs writes (-inf, INT_MAX).  dli_argmax_merge picks the global best, lowest index on ties.
    __shared__ float cv[kGemvWarps][M];
    __shared__ int ci[kGemvWarps][M];
    if (lane == 0) {
#pragma unroll
      for (int m = 0; m < M; ++m) { cv[warp][m] = best_v[m]; ci[warp][m] = best_i[m]; }
    }
    __syncthreads();
    if (tid < M) {
      float bv = -INFINITY;
      int bi = 0x7fffffff;
      for (int ww = 0; ww < kGemvWarps; ++ww) {
        const float v = cv[ww][tid];
        const int i = ci[ww][tid];
        if (i != 0x7fffffff && (bi == 0x7fffffff || v > bv || (v == bv && i < bi))) { bv = v; bi = i; }
      }
      cand[((size_t)blockIdx.x * M + tid) * 2 + 0] = bv;
      cand[((size_t)blockIdx.x * M + tid) * 2 + 1] = __int_as_float(bi);
    }
  }
}

template <int M, bool kGated, bool kNorm>
static int launch_gemv(const bf16* x, const bf16* w, const bf16* bias, const bf16* residual, bf16* y, int N, int K,
                       int epilogue, const bf16* ln_w, float eps, cudaStream_t s, float* cand = nullptr,
                       int* grid_out = nullptr) {
  constexpr int kOutPerTask = kGated ? kGemvRows / 2 : kGemvRows;
  const int ngroups_total = (N + kOutPerTask - 1) / kOutPerTask;
  const int sms = sm_count_cached();
  if (sms <= 0) return fail(DLI_ERR_CUDA, "gemv: no CUDA device");
  const size_t dyn = kNorm ? (size_t)M * K * 2 : 0;
  // resident CTAs per SM for this instantiation on this device, redone when a call needs more shared memory
  static PerDeviceOnce once;
  {
    const int d = current_device_slot();
    if (d < 0) return DLI_ERR_CUDA;
    std::lock_guard<std::mutex> lock(once.mu);
    if (dyn > once.bytes[d]) once.done[d] = false;
  }
  DLI_ONCE_PER_DEVICE(once, dev, {
    DLI_CUDA_CALL(cudaFuncSetAttribute(gemv_kernel<M, kGated, kNorm>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)(dyn > 131072 ? dyn : 131072)));
    int o = 0;
    DLI_CUDA_CALL(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&o, gemv_kernel<M, kGated, kNorm>, kGemvThreads, dyn));
    once.value[dev] = o > 0 ? o : 1;
    once.bytes[dev] = dyn;
  });
  const int occ = once.value[dev];
  const int grid = ngroups_total < sms * occ ? ngroups_total : sms * occ;  // one resident wave
  if (grid_out) *grid_out = grid;
  // bytes per row prefetched into L2 across the kernel boundary (the whole row, or 8 KB)
  static const int pf_l2 = getenv("DLI_GEMV_PF_L2") ? atoi(getenv("DLI_GEMV_PF_L2")) : 8192;
  static const int deal = getenv("DLI_GEMV_DEAL") ? atoi(getenv("DLI_GEMV_DEAL")) : 0;  // measurement knob, see kernel
  DLI_LAUNCH("gemv_kernel", gemv_kernel<M, kGated, kNorm>, dim3(grid), dim3(kGemvThreads), dyn, s, x, w, bias, residual,
             y, N, K, epilogue, ln_w, eps, pf_l2, cand, deal);
  return DLI_OK;
}

template <bool kGated, bool kNorm>
static int gemv_dispatch(const void* x, const void* w, const void* bias, const void* residual, void* y, int M, int N,
                         int K, int epilogue, const void* ln_w, float eps, dli_stream_t stream,
                         float* cand = nullptr, int* grid_out = nullptr) {
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  // M > 4 is served in passes of <= 4 rows (weights re-streamed once per pass).
  for (int m0 = 0; m0 < M; m0 += 4) {
    const int mm = (M - m0) < 4 ? (M - m0) : 4;
    const bf16* xp = (const bf16*)x + (size_t)m0 * K;
    const bf16* rp = residual ? (const bf16*)residual + (size_t)m0 * N : nullptr;
    bf16* yp = (bf16*)y + (size_t)m0 * N;
    const bf16 *wp = (const bf16*)w, *bp = (const bf16*)bias, *lp = (const bf16*)ln_w;
    int rc;
    switch (mm) {
      case 1: rc = launch_gemv<1, kGated, kNorm>(xp, wp, bp, rp, yp, N, K, epilogue, lp, eps, s, cand, grid_out); break;
      case 2: rc = launch_gemv<2, kGated, kNorm>(xp, wp, bp, rp, yp, N, K, epilogue, lp, eps, s, cand, grid_out); break;
      case 3: rc = launch_gemv<3, kGated, kNorm>(xp, wp, bp, rp, yp, N, K, epilogue, lp, eps, s, cand, grid_out); break;
      default: rc = launch_gemv<4, kGated, kNorm>(xp, wp, bp, rp, yp, N, K, epilogue, lp, eps, s, cand, grid_out); break;
    }
    if (rc != DLI_OK) return rc;
  }
  return DLI_OK;
}

// elementwise.cu: one warp per row reduces [n][B] (logit, index) candidates
int launch_argmax_merge(const float* cands, int n, int B, int idx_offset, int32_t* tokens, float* cand_out,
                        cudaStream_t s, int n_stride = 0);

static int check_gemv(const char* name, const void* x, const void* w, const void* y, int M, int N, int K) {
  DLI_CHECK_ARG(x && w && y, "%s: null pointer", name);
  DLI_CHECK_ARG(M >= 0 && N > 0 && K > 0 && K % 8 == 0, "%s: need K %% 8 == 0 (got M=%d N=%d K=%d)", name, M, N, K);
  DLI_CHECK_ARG(aligned16(x) && aligned16(w), "%s: x and w must be 16-byte aligned", name);
  return DLI_OK;
}

}  // namespace dli

extern "C" {

int dli_gemv_bf16(const void* x, const void* w, const void* bias, const void* residual, void* y,
                  int M, int N, int K, int epilogue, dli_stream_t stream) {
  using namespace dli;
  int rc = check_gemv("dli_gemv_bf16", x, w, y, M, N, K);
  if (rc != DLI_OK) return rc;
  DLI_CHECK_ARG(!(epilogue & DLI_EPI_BIAS) || bias, "dli_gemv_bf16: DLI_EPI_BIAS without bias");
  DLI_CHECK_ARG(!(epilogue & DLI_EPI_RESIDUAL) || residual, "dli_gemv_bf16: DLI_EPI_RESIDUAL without residual");
  if (M == 0) return DLI_OK;
  return gemv_dispatch<false, false>(x, w, bias, residual, y, M, N, K, epilogue, nullptr, 0.f, stream);
}

int dli_gemv_gated_silu_bf16(const void* x, const void* w_gu, void* y, int M, int F, int K,
                             dli_stream_t stream) {
  using namespace dli;
  int rc = check_gemv("dli_gemv_gated_silu_bf16", x, w_gu, y, M, F, K);
  if (rc != DLI_OK) return rc;
  DLI_CHECK_ARG(F % 128 == 0, "dli_gemv_gated_silu_bf16: need F %% 128 == 0 (got F=%d)", F);
  if (M == 0) return DLI_OK;
  return gemv_dispatch<true, false>(x, w_gu, nullptr, nullptr, y, M, F, K, 0, nullptr, 0.f, stream);
}

int dli_rmsnorm_gemv_bf16(const void* h, const void* ln_w, float eps, const void* w, const void* bias,
                          const void* residual, void* y, int M, int N, int K, int epilogue, dli_stream_t stream) {
  using namespace dli;
  int rc = check_gemv("dli_rmsnorm_gemv_bf16", h, w, y, M, N, K);
  if (rc != DLI_OK) return rc;
  DLI_CHECK_ARG(ln_w && aligned16(ln_w), "dli_rmsnorm_gemv_bf16: ln_w must be a 16-byte aligned pointer");
  DLI_CHECK_ARG(K <= 16384, "dli_rmsnorm_gemv_bf16: K <= 16384 (the normalised rows live in shared memory)");
  DLI_CHECK_ARG(!(epilogue & DLI_EPI_BIAS) || bias, "dli_rmsnorm_gemv_bf16: DLI_EPI_BIAS without bias");
  DLI_CHECK_ARG(!(epilogue & DLI_EPI_RESIDUAL) || residual, "dli_rmsnorm_gemv_bf16: DLI_EPI_RESIDUAL without residual");
  if (M == 0) return DLI_OK;
  return gemv_dispatch<false, true>(h, w, bias, residual, y, M, N, K, epilogue, ln_w, eps, stream);
}

int dli_rmsnorm_gemv_gated_silu_bf16(const void* h, const void* ln_w, float eps, const void* w_gu, void* y, int M,
                                     int F, int K, dli_stream_t stream) {
  using namespace dli;
  int rc = check_gemv("dli_rmsnorm_gemv_gated_silu_bf16", h, w_gu, y, M, F, K);
  if (rc != DLI_OK) return rc;
  DLI_CHECK_ARG(ln_w && aligned16(ln_w), "dli_rmsnorm_gemv_gated_silu_bf16: ln_w must be a 16-byte aligned pointer");
  DLI_CHECK_ARG(F % 128 == 0 && K <= 16384, "dli_rmsnorm_gemv_gated_silu_bf16: need F %% 128 == 0, K <= 16384");
  if (M == 0) return DLI_OK;
  return gemv_dispatch<true, true>(h, w_gu, nullptr, nullptr, y, M, F, K, 0, ln_w, eps, stream);
}

size_t dli_gemv_argmax_scratch_bytes(int M) {
  const int sms = dli::sm_count_cached();
  if (M <= 0 || M > 4 || sms <= 0) return 0;
  return (size_t)sms * 8 /* resident CTAs per SM, upper bound */ * M * 2 * sizeof(float);
}

int dli_gemv_argmax_bf16(const void* h, const void* ln_w, float eps, const void* w, void* logits, int M, int N, int K,
                         int idx_offset, int32_t* tokens, float* cand_out, void* scratch, size_t scratch_bytes,
                         dli_stream_t stream) {
  using namespace dli;
  int rc = check_gemv("dli_gemv_argmax_bf16", h, w, logits, M, N, K);
  if (rc != DLI_OK) return rc;
  DLI_CHECK_ARG(M <= 4, "dli_gemv_argmax_bf16: at most 4 rows per call (got %d); larger batches use "
                "dli_linear_bf16 + dli_argmax_bf16", M);
  DLI_CHECK_ARG((tokens != nullptr) != (cand_out != nullptr),
                "dli_gemv_argmax_bf16: give exactly one of tokens / cand_out");
  DLI_CHECK_ARG(!ln_w || (aligned16(ln_w) && K <= 16384), "dli_gemv_argmax_bf16: ln_w must be 16-byte aligned, K <= 16384");
  DLI_CHECK_ARG(idx_offset >= 0 && scratch && scratch_bytes >= dli_gemv_argmax_scratch_bytes(M) && aligned16(scratch),
                "dli_gemv_argmax_bf16: scratch must hold dli_gemv_argmax_scratch_bytes(M) = %zu bytes (got %zu)",
                dli_gemv_argmax_scratch_bytes(M), scratch_bytes);
  if (M == 0) return DLI_OK;
  int grid = 0;
  float* cand = static_cast<float*>(scratch);
  rc = ln_w ? gemv_dispatch<false, true>(h, w, nullptr, nullptr, logits, M, N, K, 0, ln_w, eps, stream, cand, &grid)
            : gemv_dispatch<false, false>(h, w, nullptr, nullptr, logits, M, N, K, 0, nullptr, 0.f, stream, cand, &grid);
  if (rc != DLI_OK) return rc;
  return launch_argmax_merge(cand, grid, M, idx_offset, tokens, cand_out, static_cast<cudaStream_t>(stream));
}

}  // extern "C"
