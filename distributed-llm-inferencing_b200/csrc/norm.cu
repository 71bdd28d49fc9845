// Residual-add + RMSNorm / LayerNorm, one CTA per row, row held in registers.
//
// HBM-bound: algorithmic bytes per row = d*2 (x) [+ d*2 residual] read,
// d*2 (y) [+ d*2 resid_out] written, + d*2 weight (L2-resident after row 0).
// Rounding points follow HF exactly:
//   h  = bf16(residual + x)                       modeling_llama.py:325,:331
//   hn = bf16(float(h) * rsqrt(mean(h^2) + eps))  modeling_llama.py:64-66
//   y  = bf16(weight * hn)                        modeling_llama.py:67
#include "common.cuh"

namespace dli {

constexpr int kNormThreads = 256;
constexpr int kNormMaxVec = 8;  // 256 threads * 8 vec * 8 elems = d <= 16384

__device__ __forceinline__ float block_sum(float v, float* red) {
  v = warp_sum(v);
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  if (lane == 0) red[wid] = v;
  __syncthreads();
  float t = (lane < (kNormThreads / 32)) ? red[lane] : 0.f;
  t = warp_sum(t);
  __syncthreads();  // red[] may be reused by the caller
  return t;
}

template <bool kLayerNorm>
__global__ void __launch_bounds__(kNormThreads)
add_norm_kernel(const bf16* __restrict__ x, const bf16* __restrict__ residual,
                const bf16* __restrict__ weight, const bf16* __restrict__ bias,
                bf16* resid_out, bf16* __restrict__ y, int d, float eps) {
  __shared__ float red[kNormThreads / 32];
  pdl_wait();
  pdl_trigger();
  const size_t row = blockIdx.x;
  const int nvec = d >> 3;
  const uint4* xr = reinterpret_cast<const uint4*>(x + row * d);
  const uint4* rr = residual ? reinterpret_cast<const uint4*>(residual + row * d) : nullptr;
  uint4* ro = resid_out ? reinterpret_cast<uint4*>(resid_out + row * d) : nullptr;

  float h[kNormMaxVec][8];
  float sum = 0.f, sumsq = 0.f;
#pragma unroll
  for (int i = 0; i < kNormMaxVec; ++i) {
    const int v = threadIdx.x + i * kNormThreads;
    if (v < nvec) {
      uint4 ux = ld_16B(xr + v);
      unpack8(ux, h[i]);
      if (rr) {
        float r[8];
        uint4 ur = ld_16B(rr + v);
        unpack8(ur, r);
#pragma unroll
        for (int j = 0; j < 8; ++j) h[i][j] = round_bf16(r[j] + h[i][j]);
        ux = pack8(h[i]);
      }
      if (ro) st_16B(ro + v, ux);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        sum += h[i][j];
        sumsq += h[i][j] * h[i][j];
      }
    }
  }
  float mean = 0.f, rstd;
  if (kLayerNorm) {
    // two-pass variance on the register copy, as at::native_layer_norm does
    mean = block_sum(sum, red) / (float)d;
    float sq = 0.f;
#pragma unroll
    for (int i = 0; i < kNormMaxVec; ++i) {
      const int v = threadIdx.x + i * kNormThreads;
      if (v < nvec) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float c = h[i][j] - mean;
          sq += c * c;
        }
      }
    }
    rstd = rsqrtf(block_sum(sq, red) / (float)d + eps);
  } else {
    rstd = rsqrtf(block_sum(sumsq, red) / (float)d + eps);
  }

  const uint4* wv = reinterpret_cast<const uint4*>(weight);
  const uint4* bv = reinterpret_cast<const uint4*>(bias);
  uint4* yr = reinterpret_cast<uint4*>(y + row * d);
#pragma unroll
  for (int i = 0; i < kNormMaxVec; ++i) {
    const int v = threadIdx.x + i * kNormThreads;
    if (v < nvec) {
      float w[8], o[8];
      unpack8(ld_16B(wv + v), w);
      if (kLayerNorm) {
        float b[8];
        unpack8(ld_16B(bv + v), b);
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = (h[i][j] - mean) * rstd * w[j] + b[j];
      } else {
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = w[j] * round_bf16(h[i][j] * rstd);
      }
      st_16B(yr + v, pack8(o));
    }
  }
}

// RMSNorm for MANY rows (prefill): one WARP per row, eight rows per CTA, no block barrier — every load of a row is
// requested before any is used and the reduction is shuffles only (the CTA-per-row kernel above runs 2 048 tiny
// CTAs with two block barriers each: 12.6 us for 2 048 x 4 096 where the bytes take 5).  The sum of squares is
// formed in EXACTLY the order of the CTA-per-row kernel, so both give the same bits and a token normalised in a
// prompt pass equals the same token normalised in a decode step: lane l plays the threads l, l + 32, ...,
// l + 224 of that kernel one after the other (thread t holds vectors t, t + 256, ...), each of the eight partial
// sums goes through the same shuffle tree, and the eight per-warp sums are added as that kernel's second tree adds
// them: ((r0 + r4) + (r2 + r6)) + ((r1 + r5) + (r3 + r7)).
template <int I>  // vectors per thread of the CTA-per-row kernel: d = 2048 * I
__global__ void __launch_bounds__(kNormThreads)
rmsnorm_rows_kernel(const bf16* __restrict__ x, const bf16* __restrict__ residual, const bf16* __restrict__ weight,
                    bf16* resid_out, bf16* __restrict__ y, int rows, float eps) {
  constexpr int d = 2048 * I;
  const int lane = threadIdx.x & 31;
  const size_t row = (size_t)blockIdx.x * (kNormThreads / 32) + (threadIdx.x >> 5);
  pdl_wait();
  pdl_trigger();
  if (row >= (size_t)rows) return;
  const uint4* xr = reinterpret_cast<const uint4*>(x + row * d);
  const uint4* rr = residual ? reinterpret_cast<const uint4*>(residual + row * d) : nullptr;
  uint4* ro = resid_out ? reinterpret_cast<uint4*>(resid_out + row * d) : nullptr;
  uint4 hp[8][I];  // h = bf16(residual + x), packed
#pragma unroll
  for (int w = 0; w < 8; ++w)
#pragma unroll
    for (int i = 0; i < I; ++i) hp[w][i] = ld_16B(xr + lane + 32 * w + 256 * i);
  if (rr) {
#pragma unroll
    for (int w = 0; w < 8; ++w)
#pragma unroll
      for (int i = 0; i < I; ++i) {
        float a[8], r[8];
        unpack8(hp[w][i], a);
        unpack8(ld_16B(rr + lane + 32 * w + 256 * i), r);
#pragma unroll
        for (int j = 0; j < 8; ++j) a[j] = round_bf16(r[j] + a[j]);
        hp[w][i] = pack8(a);
      }
  }
  float part[8];
#pragma unroll
  for (int w = 0; w < 8; ++w) {
    float sumsq = 0.f;
#pragma unroll
    for (int i = 0; i < I; ++i) {
      float h[8];
      unpack8(hp[w][i], h);
      if (ro) st_16B(ro + lane + 32 * w + 256 * i, hp[w][i]);
#pragma unroll
      for (int j = 0; j < 8; ++j) sumsq += h[j] * h[j];
    }
    part[w] = warp_sum(sumsq);
  }
  const float total = ((part[0] + part[4]) + (part[2] + part[6])) + ((part[1] + part[5]) + (part[3] + part[7]));
  const float rstd = rsqrtf(total / (float)d + eps);
  const uint4* wv = reinterpret_cast<const uint4*>(weight);
  uint4* yr = reinterpret_cast<uint4*>(y + row * d);
#pragma unroll
  for (int w = 0; w < 8; ++w)
#pragma unroll
    for (int i = 0; i < I; ++i) {
      float h[8], wt[8], o[8];
      unpack8(hp[w][i], h);
      unpack8(ld_16B(wv + lane + 32 * w + 256 * i), wt);
#pragma unroll
      for (int j = 0; j < 8; ++j) o[j] = wt[j] * round_bf16(h[j] * rstd);
      st_16B(yr + lane + 32 * w + 256 * i, pack8(o));
    }
}

static int launch_norm(bool ln, const void* x, const void* residual, const void* weight,
                       const void* bias, void* resid_out, void* y, int rows, int d,
                       float eps, dli_stream_t stream) {
  const char* name = ln ? "dli_add_layernorm_bf16" : "dli_add_rmsnorm_bf16";
  DLI_CHECK_ARG(x && weight && y, "%s: null pointer", name);
  DLI_CHECK_ARG(!ln || bias, "%s: bias is required", name);
  DLI_CHECK_ARG(rows >= 0 && d > 0 && d % 8 == 0 && d <= kNormThreads * kNormMaxVec * 8,
                "%s: need d %% 8 == 0 and d <= %d (got rows=%d d=%d)", name,
                kNormThreads * kNormMaxVec * 8, rows, d);
  DLI_CHECK_ARG(aligned16(x) && aligned16(weight) && aligned16(y) && aligned16(residual) &&
                    aligned16(resid_out) && aligned16(bias),
                "%s: pointers must be 16-byte aligned", name);
  if (rows == 0) return DLI_OK;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  constexpr int kRowsPerCta = kNormThreads / 32;
  if (!ln && rows >= 64 && (d == 2048 || d == 4096 || d == 8192)) {  // many rows: one warp per row (same bits)
    const dim3 grid((rows + kRowsPerCta - 1) / kRowsPerCta);
    if (d == 2048)
      DLI_LAUNCH(name, rmsnorm_rows_kernel<1>, grid, dim3(kNormThreads), (size_t)0, s, (const bf16*)x, (const bf16*)residual,
                 (const bf16*)weight, (bf16*)resid_out, (bf16*)y, rows, eps);
    else if (d == 4096)
      DLI_LAUNCH(name, rmsnorm_rows_kernel<2>, grid, dim3(kNormThreads), (size_t)0, s, (const bf16*)x, (const bf16*)residual,
                 (const bf16*)weight, (bf16*)resid_out, (bf16*)y, rows, eps);
    else
      DLI_LAUNCH(name, rmsnorm_rows_kernel<4>, grid, dim3(kNormThreads), (size_t)0, s, (const bf16*)x, (const bf16*)residual,
                 (const bf16*)weight, (bf16*)resid_out, (bf16*)y, rows, eps);
    return DLI_OK;
  }
  if (ln)
    DLI_LAUNCH(name, add_norm_kernel<true>, dim3(rows), dim3(kNormThreads), (size_t)0, s, (const bf16*)x,
               (const bf16*)residual, (const bf16*)weight, (const bf16*)bias, (bf16*)resid_out, (bf16*)y, d, eps);
  else
    DLI_LAUNCH(name, add_norm_kernel<false>, dim3(rows), dim3(kNormThreads), (size_t)0, s, (const bf16*)x,
               (const bf16*)residual, (const bf16*)weight, (const bf16*)nullptr, (bf16*)resid_out, (bf16*)y, d, eps);
  return DLI_OK;
}

}  // namespace dli

extern "C" {

int dli_add_rmsnorm_bf16(const void* x, const void* residual, const void* weight,
                         void* resid_out, void* y, int rows, int d, float eps,
                         dli_stream_t stream) {
  return dli::launch_norm(false, x, residual, weight, nullptr, resid_out, y, rows, d, eps, stream);
}

int dli_add_layernorm_bf16(const void* x, const void* residual, const void* weight,
                           const void* bias, void* resid_out, void* y, int rows, int d,
                           float eps, dli_stream_t stream) {
  return dli::launch_norm(true, x, residual, weight, bias, resid_out, y, rows, d, eps, stream);
}

}  // extern "C"
