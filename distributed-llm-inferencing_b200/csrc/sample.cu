// Sampling epilogue: temperature -> top-k -> top-p -> multinomial, one CTA per row.
//
// This is what the reference worker asks of generate() by default
// (worker/app.py:297-305: do_sample=True, top_k=50, top_p=0.95, temperature=0.8), i.e.
// TemperatureLogitsWarper -> TopKLogitsWarper -> TopPLogitsWarper -> softmax ->
// torch.multinomial (TF:generation/logits_process.py, TF:generation/utils.py:2762-2793),
// on the bf16 logits cast to fp32:
//   top-k keeps every logit >= the k-th largest (ties at the threshold are all kept, as
//         `scores < topk(scores, k)[0][..., -1]` does);
//   top-p keeps, in descending order of probability, the tokens whose EXCLUSIVE prefix sum
//         of (top-k-renormalised) probabilities is < top_p — the same set as HF's ascending
//         `cumsum <= 1 - top_p` removal — and always at least one token;
//   the draw is inverse-CDF over the kept tokens with a Philox4x32-10 uniform keyed by
//         (seed, step, row).  torch.multinomial uses a different stream and algorithm, so
//         draws agree in distribution, not bit for bit; the kept set and its probabilities
//         are what the parity test compares.
//
// HBM-bound: the row (V * 2 bytes) is read three times (two radix-select passes over the
// 16-bit keys + the gather); 0.77 MB per row at V = 128 256, negligible next to lm_head.
#include "common.cuh"

namespace dli {

constexpr int kSampleThreads = 512;
constexpr int kMaxCand = 2048;  // candidates kept in shared memory (k plus ties)

// monotone key: larger bf16 value <-> larger 16-bit key
__device__ __forceinline__ uint32_t bf16_key(uint16_t b) { return (b & 0x8000u) ? (uint16_t)~b : (uint16_t)(b | 0x8000u); }
__device__ __forceinline__ float key_value(uint32_t key) {
  const uint16_t b = (key & 0x8000u) ? (uint16_t)(key & 0x7fffu) : (uint16_t)~key;
  return __uint_as_float((uint32_t)b << 16);
}

struct Philox {
  uint32_t c[4], k[2];
  __device__ __forceinline__ static void mulhilo(uint32_t a, uint32_t b, uint32_t& hi, uint32_t& lo) {
    const uint64_t p = (uint64_t)a * b;
    hi = (uint32_t)(p >> 32);
    lo = (uint32_t)p;
  }
  __device__ __forceinline__ void round() {
    uint32_t h0, l0, h1, l1;
    mulhilo(0xD2511F53u, c[0], h0, l0);
    mulhilo(0xCD9E8D57u, c[2], h1, l1);
    const uint32_t n0 = h1 ^ c[1] ^ k[0], n1 = l1, n2 = h0 ^ c[3] ^ k[1], n3 = l0;
    c[0] = n0; c[1] = n1; c[2] = n2; c[3] = n3;
    k[0] += 0x9E3779B9u;
    k[1] += 0xBB67AE85u;
  }
};
// uniform in [0, 1) from Philox4x32-10 keyed by seed, counter (step, row)
__device__ __forceinline__ float philox_uniform(uint64_t seed, uint32_t step, uint32_t row) {
  Philox p;
  p.c[0] = step; p.c[1] = row; p.c[2] = 0x6a09e667u; p.c[3] = 0xbb67ae85u;
  p.k[0] = (uint32_t)seed; p.k[1] = (uint32_t)(seed >> 32);
#pragma unroll
  for (int i = 0; i < 10; ++i) p.round();
  return (float)(p.c[0] >> 8) * (1.0f / 16777216.0f);
}

// Largest key t such that at least `k` elements have key >= t (two 8-bit radix passes).
__device__ uint32_t kth_largest_key(const uint16_t* __restrict__ row, int V, int k, uint32_t* hist /*256*/,
                                    uint32_t* sh /*2*/) {
  const int tid = threadIdx.x;
  uint32_t prefix = 0;
  int remaining = k;
  for (int pass = 0; pass < 2; ++pass) {
    for (int i = tid; i < 256; i += kSampleThreads) hist[i] = 0;
    __syncthreads();
    for (int i = tid; i < V; i += kSampleThreads) {
      const uint32_t key = bf16_key(row[i]);
      if (pass == 0) atomicAdd(&hist[key >> 8], 1u);
      else if ((key >> 8) == prefix) atomicAdd(&hist[key & 0xff], 1u);
    }
    __syncthreads();
    if (tid == 0) {
      int acc = 0, bin = 255;
      for (; bin > 0; --bin) {
        if (acc + (int)hist[bin] >= remaining) break;
        acc += hist[bin];
      }
      sh[0] = bin;
      sh[1] = remaining - acc;  // still needed inside this bin
    }
    __syncthreads();
    if (pass == 0) { prefix = sh[0]; remaining = sh[1]; }
    else prefix = (prefix << 8) | sh[0];
    __syncthreads();
  }
  return prefix;
}

__global__ void __launch_bounds__(kSampleThreads)
sample_kernel(const bf16* __restrict__ logits, int ld, int32_t* __restrict__ out, int V, float inv_temperature,
              int top_k, float top_p, const uint64_t* __restrict__ seed, const int32_t* __restrict__ step,
              float* __restrict__ dbg_probs, int32_t* __restrict__ dbg_ids, int dbg_cap, int per_row,
              const int32_t* __restrict__ row_mask) {
  __shared__ uint32_t hist[256];
  __shared__ uint32_t sh[2];
  __shared__ float cval[kMaxCand];
  __shared__ int cidx[kMaxCand];
  __shared__ int ncand;
  __shared__ float red[kSampleThreads / 32];
  pdl_wait();
  pdl_trigger();
  if (row_mask != nullptr && row_mask[blockIdx.x] == 0) return;  // this row keeps the token already in out[]
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const uint16_t* row = reinterpret_cast<const uint16_t*>(logits + (size_t)blockIdx.x * ld);
  const int k = top_k > 0 ? min(top_k, V) : V;

  // ---- top-k threshold and candidate gather ------------------------------------------------
  const uint32_t thr = kth_largest_key(row, V, min(k, kMaxCand), hist, sh);
  if (tid == 0) ncand = 0;
  __syncthreads();
  for (int i = tid; i < V; i += kSampleThreads) {
    const uint32_t key = bf16_key(row[i]);
    if (key >= thr) {
      const int slot = atomicAdd(&ncand, 1);
      if (slot < kMaxCand) { cval[slot] = key_value(key) * inv_temperature; cidx[slot] = i; }
    }
  }
  __syncthreads();
  const int n = min(ncand, kMaxCand);
  int npow = 1;
  while (npow < n) npow <<= 1;
  for (int i = n + tid; i < npow; i += kSampleThreads) { cval[i] = -INFINITY; cidx[i] = -1; }
  __syncthreads();
  // bitonic sort, descending by value, equal values by ascending token id.  Which of several
  // EQUAL-probability tokens survive at the nucleus boundary depends on torch.sort's tie order
  // in the reference (not specified); here the lower token ids stay.
  for (int size = 2; size <= npow; size <<= 1) {
    for (int stride = size >> 1; stride > 0; stride >>= 1) {
      for (int i = tid; i < npow; i += kSampleThreads) {
        const int j = i ^ stride;
        if (j > i) {
          const bool desc = (i & size) == 0;
          const float a = cval[i], b = cval[j];
          const int ia = cidx[i], ib = cidx[j];
          const bool a_first = (a > b) || (a == b && (unsigned)ia < (unsigned)ib);
          if (a_first != desc) { cval[i] = b; cval[j] = a; cidx[i] = ib; cidx[j] = ia; }
        }
      }
      __syncthreads();
    }
  }
  // ---- softmax over the top-k set, nucleus cut, renormalise, draw (n <= 2048: one warp scans) ----
  const float vmax = cval[0];
  float part = 0.f;
  for (int i = tid; i < n; i += kSampleThreads) part += __expf(cval[i] - vmax);
  part = warp_sum(part);
  if (lane == 0) red[warp] = part;
  __syncthreads();
  float denom = 0.f;
#pragma unroll
  for (int w = 0; w < kSampleThreads / 32; ++w) denom += red[w];
  for (int i = tid; i < n; i += kSampleThreads) cval[i] = __expf(cval[i] - vmax) / denom;  // probabilities, descending
  __syncthreads();
  if (warp == 0) {
    // exclusive prefix sums in descending order; keep while prefix < top_p (always the first)
    float carry = 0.f;
    int kept = 0;
    for (int base = 0; base < n; base += 32) {
      const int i = base + lane;
      const float p = i < n ? cval[i] : 0.f;
      float incl = p;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const float t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += t;
      }
      const float excl = carry + incl - p;
      const bool keep = i < n && (i == 0 || excl < top_p);
      const unsigned m = __ballot_sync(0xffffffffu, keep);
      kept += __popc(m);
      carry += __shfl_sync(0xffffffffu, incl, 31);
      if (m != 0xffffffffu) break;
    }
    kept = max(kept, 1);
    float mass = 0.f;
    for (int i = lane; i < kept; i += 32) mass += cval[i];
    mass = warp_sum(mass);
    if (dbg_probs) {
      for (int i = lane; i < dbg_cap; i += 32) {
        dbg_probs[(size_t)blockIdx.x * dbg_cap + i] = i < kept ? cval[i] / mass : 0.f;
        dbg_ids[(size_t)blockIdx.x * dbg_cap + i] = i < kept ? cidx[i] : -1;
      }
    }
    if (lane == 0) {
      // per_row: every row has its own (seed, step) and the row index stays out of the key, so a sequence's
      // draws do not depend on the batch row it happens to occupy
      const int r = per_row ? blockIdx.x : 0;
      const float u = philox_uniform(seed ? seed[r] : 0ull, step ? (uint32_t)step[r] : 0u,
                                     per_row ? 0u : blockIdx.x) * mass;
      float acc = 0.f;
      int pick = kept - 1;
      for (int i = 0; i < kept; ++i) {
        acc += cval[i];
        if (u < acc) { pick = i; break; }
      }
      out[blockIdx.x] = cidx[pick];
    }
  }
}

}  // namespace dli

extern "C" int dli_sample_bf16(const void* logits, int ld, int32_t* out, int B, int V, float temperature, int top_k,
                               float top_p, const uint64_t* seed, const int32_t* step, float* dbg_probs,
                               int32_t* dbg_ids, int dbg_cap, dli_stream_t stream) {
  using namespace dli;
  DLI_CHECK_ARG(logits && out, "dli_sample_bf16: null pointer");
  DLI_CHECK_ARG(B >= 0 && V > 0 && ld >= V, "dli_sample_bf16: bad sizes B=%d V=%d ld=%d", B, V, ld);
  DLI_CHECK_ARG(temperature > 0.f && top_p > 0.f && top_p <= 1.f && top_k >= 0,
                "dli_sample_bf16: need temperature > 0, 0 < top_p <= 1, top_k >= 0 (0 = off)");
  DLI_CHECK_ARG((top_k > 0 && top_k <= kMaxCand) || V <= kMaxCand,
                "dli_sample_bf16: top_k must be 1..%d when the vocabulary is larger than that", kMaxCand);
  DLI_CHECK_ARG((dbg_probs == nullptr) == (dbg_ids == nullptr) && dbg_cap >= 0, "dli_sample_bf16: bad debug buffers");
  if (B == 0) return DLI_OK;
  DLI_LAUNCH("dli_sample_bf16", sample_kernel, dim3(B), dim3(kSampleThreads), (size_t)0, static_cast<cudaStream_t>(stream),
             (const bf16*)logits, ld, out, V, 1.0f / temperature, top_k, top_p, seed, step, dbg_probs, dbg_ids, dbg_cap,
             0, (const int32_t*)nullptr);
  return DLI_OK;
}

extern "C" int dli_sample_rows_bf16(const void* logits, int ld, int32_t* out, int B, int V, float temperature,
                                    int top_k, float top_p, const uint64_t* seeds, const int32_t* steps,
                                    const int32_t* row_mask, dli_stream_t stream) {
  using namespace dli;
  DLI_CHECK_ARG(logits && out && seeds && steps, "dli_sample_rows_bf16: null pointer");
  DLI_CHECK_ARG(B >= 0 && V > 0 && ld >= V, "dli_sample_rows_bf16: bad sizes B=%d V=%d ld=%d", B, V, ld);
  DLI_CHECK_ARG(temperature > 0.f && top_p > 0.f && top_p <= 1.f && top_k >= 0,
                "dli_sample_rows_bf16: need temperature > 0, 0 < top_p <= 1, top_k >= 0 (0 = off)");
  DLI_CHECK_ARG((top_k > 0 && top_k <= kMaxCand) || V <= kMaxCand,
                "dli_sample_rows_bf16: top_k must be 1..%d when the vocabulary is larger than that", kMaxCand);
  if (B == 0) return DLI_OK;
  DLI_LAUNCH("dli_sample_rows_bf16", sample_kernel, dim3(B), dim3(kSampleThreads), (size_t)0,
             static_cast<cudaStream_t>(stream), (const bf16*)logits, ld, out, V, 1.0f / temperature, top_k, top_p, seeds,
             steps, (float*)nullptr, (int32_t*)nullptr, 0, 1, row_mask);
  return DLI_OK;
}
