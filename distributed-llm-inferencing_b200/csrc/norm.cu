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
