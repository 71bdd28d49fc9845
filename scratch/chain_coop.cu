// Experiment: the 4 weight-streaming GEMVs of a decoder layer (batch 1) as ONE cooperative kernel
// with grid barriers between the phases, next phase's first weights requested BEFORE the barrier,
// against the library's PDL chain of 4 kernels per layer.  Llama-3-8B shapes, 8 weight sets cycled.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I../distributed-llm-inferencing_b200/csrc \
//        -I../include chain_coop.cu -L../distributed-llm-inferencing_b200 -ldli_b200 -o chain_coop
#include <cooperative_groups.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "dli.h"
namespace cg = cooperative_groups;
typedef __nv_bfloat16 bf16;

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s:%d %s\n", __FILE__, __LINE__, cudaGetErrorString(e_)); exit(1);} } while (0)

__device__ __forceinline__ uint4 ld_stream(const void* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
  return r;
}
__device__ __forceinline__ float lo(uint32_t u) { return __uint_as_float(u << 16); }
__device__ __forceinline__ float hi(uint32_t u) { return __uint_as_float(u & 0xffff0000u); }
__device__ __forceinline__ void unpack8(const uint4& u, float* f) {
  f[0] = lo(u.x); f[1] = hi(u.x); f[2] = lo(u.y); f[3] = hi(u.y); f[4] = lo(u.z); f[5] = hi(u.z); f[6] = lo(u.w); f[7] = hi(u.w);
}
__device__ __forceinline__ float rbf(float x) { return __bfloat162float(__float2bfloat16_rn(x)); }
__device__ __forceinline__ uint32_t pk(float a, float b) { __nv_bfloat162 t = __floats2bfloat162_rn(a, b); return *reinterpret_cast<uint32_t*>(&t); }
__device__ __forceinline__ float wsum(float v) { for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o); return v; }

constexpr int T = 256, W = 8, R = 4;
struct Smem {
  float part[2][R][T];
  float red[W];
};

struct Pre { uint4 a[R], b[R]; };

template <bool kGated>
__device__ __forceinline__ void rows_of(int task, const bf16* w, int N, int K, const uint4* (&wrow)[R]) {
  constexpr int OPT = kGated ? R / 2 : R;
  const int n0 = task * OPT;
#pragma unroll
  for (int r = 0; r < R; ++r) {
    size_t row;
    if (kGated) { const int n = min(n0 + (r % OPT), N - 1); row = (size_t)(n >> 7) * 256 + (n & 127) + ((r / OPT) ? 128 : 0); }
    else row = (size_t)min(n0 + r, N - 1);
    wrow[r] = reinterpret_cast<const uint4*>(w + row * (size_t)K);
  }
}

// request the first iteration of this CTA's first task (no dependency on activations)
template <bool kGated>
__device__ __forceinline__ void prefetch(Pre& p, const bf16* w, int N, int K, int ntasks) {
  const uint4 zero = make_uint4(0, 0, 0, 0);
  const int kvec = K >> 3, tid = threadIdx.x;
  const uint4* wrow[R];
  if ((int)blockIdx.x < ntasks) {
    rows_of<kGated>(blockIdx.x, w, N, K, wrow);
#pragma unroll
    for (int r = 0; r < R; ++r) {
      p.a[r] = tid < kvec ? ld_stream(wrow[r] + tid) : zero;
      p.b[r] = tid + T < kvec ? ld_stream(wrow[r] + tid + T) : zero;
    }
  }
}

template <bool kGated, bool kNorm>
__device__ __forceinline__ void phase(Smem& sm, uint4* xs, const bf16* x, const bf16* w, const bf16* residual, bf16* y,
                                      int N, int K, const bf16* ln_w, float eps, int ntasks, Pre& pre, bool use_pre) {
  constexpr int OPT = kGated ? R / 2 : R;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int kvec = K >> 3, iters = (kvec + T - 1) / T;
  const uint4 zero = make_uint4(0, 0, 0, 0);
  const uint4* xv = reinterpret_cast<const uint4*>(x);
  if (kNorm) {
    float ss = 0.f;
    for (int i = tid; i < kvec; i += T) { float f[8]; unpack8(__ldcg(xv + i), f); for (int e = 0; e < 8; ++e) ss += f[e] * f[e]; }
    ss = wsum(ss);
    if (lane == 0) sm.red[warp] = ss;
    __syncthreads();
    float t = 0.f;
    for (int ww = 0; ww < W; ++ww) t += sm.red[ww];
    const float rstd = rsqrtf(t / (float)K + eps);
    const uint4* lw = reinterpret_cast<const uint4*>(ln_w);
    for (int i = tid; i < kvec; i += T) {
      float g[8], f[8];
      unpack8(lw[i], g); unpack8(__ldcg(xv + i), f);
      for (int e = 0; e < 8; ++e) f[e] = g[e] * rbf(f[e] * rstd);
      uint4 u; u.x = pk(f[0], f[1]); u.y = pk(f[2], f[3]); u.z = pk(f[4], f[5]); u.w = pk(f[6], f[7]);
      xs[i] = u;
    }
    __syncthreads();
    xv = xs;
  }
  int parity = 0;
  bool first = use_pre;
  for (int task = blockIdx.x; task < ntasks; task += gridDim.x, parity ^= 1) {
    const int n0 = task * OPT;
    const uint4* wrow[R];
    rows_of<kGated>(task, w, N, K, wrow);
    float acc[R];
#pragma unroll
    for (int r = 0; r < R; ++r) acc[r] = 0.f;
    for (int j = 0; j < iters; j += 2) {
      const int i0 = tid + j * T, i1 = i0 + T;
      const bool p0 = i0 < kvec, p1 = i1 < kvec;
      uint4 wv0[R], wv1[R];
      if (first && j == 0) {
#pragma unroll
        for (int r = 0; r < R; ++r) { wv0[r] = pre.a[r]; wv1[r] = pre.b[r]; }
      } else {
#pragma unroll
        for (int r = 0; r < R; ++r) { wv0[r] = p0 ? ld_stream(wrow[r] + i0) : zero; wv1[r] = p1 ? ld_stream(wrow[r] + i1) : zero; }
      }
      // activations were written by other CTAs in the previous phase: L2 loads (kNorm: shared memory)
      const uint4 xa = p0 ? (kNorm ? xv[i0] : __ldcg(xv + i0)) : zero;
      const uint4 xb = p1 ? (kNorm ? xv[i1] : __ldcg(xv + i1)) : zero;
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        float xf[8];
        unpack8(h ? xb : xa, xf);
#pragma unroll
        for (int r = 0; r < R; ++r) {
          float wf[8];
          unpack8(h ? wv1[r] : wv0[r], wf);
#pragma unroll
          for (int e = 0; e < 8; ++e) acc[r] = fmaf(wf[e], xf[e], acc[r]);
        }
      }
    }
    first = false;
#pragma unroll
    for (int r = 0; r < R; ++r) sm.part[parity][r][tid] = acc[r];
    __syncthreads();
    for (int o = warp; o < OPT; o += W) {
      float s = 0.f, s2 = 0.f;
#pragma unroll
      for (int ww = 0; ww < W; ++ww) {
        s += sm.part[parity][o][ww * 32 + lane];
        if (kGated) s2 += sm.part[parity][o + OPT][ww * 32 + lane];
      }
      s = wsum(s);
      if (kGated) s2 = wsum(s2);
      const int n = n0 + o;
      if (lane == 0 && n < N) {
        float v;
        if (kGated) { const float g = rbf(s), u = rbf(s2); v = rbf(g / (1.0f + expf(-g))) * u; }
        else { v = rbf(s); if (residual) v += __bfloat162float(__ldcg(residual + n)); }
        y[n] = __float2bfloat16_rn(v);
      }
    }
  }
}

struct Layer { const bf16 *ln1, *wqkv, *wo, *ln2, *wgu, *wdown; };

template <int MINB>
__global__ void __launch_bounds__(T, MINB)
chain_kernel(const Layer* layers, int nsets, int L, bf16* h, bf16* qkv, const bf16* attn, bf16* act, int d, int nqkv, int ff,
             float eps, int use_pre) {
  cg::grid_group grid = cg::this_grid();
  __shared__ Smem sm;
  extern __shared__ __align__(16) uint8_t dyn[];
  uint4* xs = reinterpret_cast<uint4*>(dyn);
  Pre pre;
  const int tq = (nqkv + R - 1) / R, to = (d + R - 1) / R, tg = (ff + 1) / 2, td = (d + R - 1) / R;
  if (use_pre) prefetch<false>(pre, layers[0].wqkv, nqkv, d, tq);
  for (int l = 0; l < L; ++l) {
    const Layer& Lw = layers[l % nsets];
    const Layer& Ln = layers[(l + 1) % nsets];
    phase<false, true>(sm, xs, h, Lw.wqkv, nullptr, qkv, nqkv, d, Lw.ln1, eps, tq, pre, use_pre);
    if (use_pre) prefetch<false>(pre, Lw.wo, d, d, to);
    grid.sync();
    phase<false, false>(sm, xs, attn, Lw.wo, h, h, d, d, nullptr, eps, to, pre, use_pre);
    if (use_pre) prefetch<true>(pre, Lw.wgu, ff, d, tg);
    grid.sync();
    phase<true, true>(sm, xs, h, Lw.wgu, nullptr, act, ff, d, Lw.ln2, eps, tg, pre, use_pre);
    if (use_pre) prefetch<false>(pre, Lw.wdown, d, ff, td);
    grid.sync();
    phase<false, false>(sm, xs, act, Lw.wdown, h, h, d, ff, nullptr, eps, td, pre, use_pre);
    if (use_pre) prefetch<false>(pre, Ln.wqkv, nqkv, d, tq);
    grid.sync();
  }
}

#include <map>
static bf16* rnd(size_t n, float scale, float shift) {
  static std::map<size_t, std::vector<bf16>> cache;  // same values for every buffer of a size: only addresses matter
  std::vector<bf16>& hbuf = cache[n + (size_t)(shift * 7)];
  if (hbuf.empty()) {
    hbuf.resize(n);
    uint32_t s = 12345u + (uint32_t)n;
    for (size_t i = 0; i < n; ++i) { s = s * 1664525u + 1013904223u; hbuf[i] = __float2bfloat16(shift + scale * ((float)(s >> 8) / 8388608.f - 1.f)); }
  }
  bf16* p; CK(cudaMalloc(&p, n * 2)); CK(cudaMemcpy(p, hbuf.data(), n * 2, cudaMemcpyHostToDevice));
  return p;
}

int main() {
  const int d = 4096, ff = 14336, nqkv = 6144, nsets = 8, L = 32;
  std::vector<Layer> hl(nsets);
  for (int i = 0; i < nsets; ++i) {
    hl[i].ln1 = rnd(d, 0.1f, 1.f); hl[i].ln2 = rnd(d, 0.1f, 1.f);
    hl[i].wqkv = rnd((size_t)nqkv * d, 0.02f, 0.f); hl[i].wo = rnd((size_t)d * d, 0.02f, 0.f);
    hl[i].wgu = rnd((size_t)2 * ff * d, 0.02f, 0.f); hl[i].wdown = rnd((size_t)d * ff, 0.02f, 0.f);
  }
  Layer* dl; CK(cudaMalloc(&dl, sizeof(Layer) * nsets)); CK(cudaMemcpy(dl, hl.data(), sizeof(Layer) * nsets, cudaMemcpyHostToDevice));
  bf16* h0 = rnd(d, 1.f, 0.f); bf16* attn = rnd(d, 1.f, 0.f);
  bf16 *h, *qkv, *act, *h2;
  CK(cudaMalloc(&h, d * 2)); CK(cudaMalloc(&h2, d * 2)); CK(cudaMalloc(&qkv, nqkv * 2)); CK(cudaMalloc(&act, ff * 2));
  cudaStream_t s; CK(cudaStreamCreate(&s));
  cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  const double mb = ((double)nqkv * d + (double)d * d + 2.0 * ff * d + (double)d * ff) * 2 / 1e6;

  // ---- cooperative chain --------------------------------------------------------------
  const size_t dynb = d * 2;
  int occ = 0; CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, chain_kernel<3>, T, dynb));
  cudaDeviceProp prop; CK(cudaGetDeviceProperties(&prop, 0));
  printf("coop occupancy %d CTAs/SM, %d SMs\n", occ, prop.multiProcessorCount);
  for (int use_pre = 0; use_pre < 2; ++use_pre) {
    for (int ctas_per_sm = 2; ctas_per_sm <= occ && ctas_per_sm <= 3; ++ctas_per_sm) {
      const int grid = prop.multiProcessorCount * ctas_per_sm;
      float best = 1e9f;
      for (int rep = 0; rep < 6; ++rep) {
        CK(cudaMemcpyAsync(h, h0, d * 2, cudaMemcpyDeviceToDevice, s));
        int nsets_ = nsets, L_ = L, d_ = d, nq_ = nqkv, ff_ = ff, up = use_pre; float eps = 1e-5f;
        const Layer* dlp = dl; const bf16* attnp = attn;
        void* args[] = {&dlp, &nsets_, &L_, &h, &qkv, &attnp, &act, &d_, &nq_, &ff_, &eps, &up};
        CK(cudaEventRecord(e0, s));
        CK(cudaLaunchCooperativeKernel(ctas_per_sm == 3 ? (void*)chain_kernel<3> : (void*)chain_kernel<2>, dim3(grid), dim3(T), args, dynb, s));
        CK(cudaEventRecord(e1, s)); CK(cudaEventSynchronize(e1));
        float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
        if (rep >= 2 && ms < best) best = ms;
      }
      printf("cooperative chain  pre=%d ctas/SM=%d: %7.2f us/layer  %.2f TB/s\n", use_pre, ctas_per_sm, best * 1e3 / L, mb / (best * 1e3 / L));
    }
  }
  std::vector<bf16> hc(d); CK(cudaMemcpy(hc.data(), h, d * 2, cudaMemcpyDeviceToHost));

  // ---- library PDL chain in a CUDA graph ------------------------------------------------
  for (int pdl = 0; pdl < 2; ++pdl) {
    dli_set_pdl(pdl);
    auto chain = [&]() {
      for (int l = 0; l < L; ++l) {
        const Layer& Lw = hl[l % nsets];
        int rc = dli_rmsnorm_gemv_bf16(h2, Lw.ln1, 1e-5f, Lw.wqkv, nullptr, nullptr, qkv, 1, nqkv, d, 0, s);
        rc |= dli_gemv_bf16(attn, Lw.wo, nullptr, h2, h2, 1, d, d, DLI_EPI_RESIDUAL, s);
        rc |= dli_rmsnorm_gemv_gated_silu_bf16(h2, Lw.ln2, 1e-5f, Lw.wgu, act, 1, ff, d, s);
        rc |= dli_gemv_bf16(act, Lw.wdown, nullptr, h2, h2, 1, d, ff, DLI_EPI_RESIDUAL, s);
        if (rc) { printf("dli error: %s\n", dli_last_error()); exit(1); }
      }
    };
    CK(cudaMemcpyAsync(h2, h0, d * 2, cudaMemcpyDeviceToDevice, s));
    chain(); CK(cudaStreamSynchronize(s));
    cudaGraph_t g; cudaGraphExec_t ge;
    CK(cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal));
    chain();
    CK(cudaStreamEndCapture(s, &g)); CK(cudaGraphInstantiate(&ge, g, 0));
    float best = 1e9f;
    for (int rep = 0; rep < 6; ++rep) {
      CK(cudaMemcpyAsync(h2, h0, d * 2, cudaMemcpyDeviceToDevice, s));
      CK(cudaEventRecord(e0, s)); CK(cudaGraphLaunch(ge, s)); CK(cudaEventRecord(e1, s)); CK(cudaEventSynchronize(e1));
      float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
      if (rep >= 2 && ms < best) best = ms;
    }
    printf("library chain (graph) pdl=%d:          %7.2f us/layer  %.2f TB/s\n", pdl, best * 1e3 / L, mb / (best * 1e3 / L));
  }
  std::vector<bf16> hg(d); CK(cudaMemcpy(hg.data(), h2, d * 2, cudaMemcpyDeviceToHost));
  int diff = 0; for (int i = 0; i < d; ++i) diff += (*(uint16_t*)&hc[i] != *(uint16_t*)&hg[i]);
  printf("residual stream after %d layers: %d of %d elements differ between the two chains\n", L, diff, d);
  return 0;
}
