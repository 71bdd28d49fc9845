// Embedding gather (+ learned positions) and the greedy argmax epilogue.
#include "common.cuh"

namespace dli {

// One CTA per token row; 16-byte gathers.  Bytes per row: d*2 read (+ d*2 for
// the position row), d*2 written.
__global__ void __launch_bounds__(128)
embedding_kernel(const int32_t* __restrict__ ids, const int32_t* __restrict__ positions,
                 const bf16* __restrict__ table, const bf16* __restrict__ pos_table,
                 bf16* __restrict__ out, int d, int vocab) {
  const size_t row = blockIdx.x;
  pdl_wait();
  pdl_trigger();
  int id = ids[row];
  id = id < 0 ? 0 : (id >= vocab ? vocab - 1 : id);
  const uint4* src = reinterpret_cast<const uint4*>(table + (size_t)id * d);
  const uint4* psrc = pos_table ? reinterpret_cast<const uint4*>(pos_table + (size_t)positions[row] * d) : nullptr;
  uint4* dst = reinterpret_cast<uint4*>(out + row * d);
  for (int v = threadIdx.x; v < (d >> 3); v += blockDim.x) {
    uint4 u = ld_16B(src + v);
    if (psrc) {
      float a[8], p[8];
      unpack8(u, a);
      unpack8(ld_16B(psrc + v), p);
#pragma unroll
      for (int e = 0; e < 8; ++e) a[e] += p[e];
      u = pack8(a);
    }
    st_16B(dst + v, u);
  }
}

// Argmax over a bf16 row, lowest index on ties (torch.argmax behaviour).
// One CTA per row; V*2 bytes read per row.
constexpr int kArgmaxThreads = 512;

__device__ __forceinline__ void argmax_combine(float& v, int& i, float ov, int oi) {
  if (ov > v || (ov == v && oi < i)) { v = ov; i = oi; }
}

__global__ void __launch_bounds__(kArgmaxThreads)
argmax_kernel(const bf16* __restrict__ logits, int ld, int32_t* __restrict__ out, int V,
              float* __restrict__ cand, int idx_offset) {
  __shared__ float sv[kArgmaxThreads / 32];
  __shared__ int si[kArgmaxThreads / 32];
  pdl_wait();
  pdl_trigger();
  const bf16* row = logits + (size_t)blockIdx.x * ld;
  float best = -INFINITY;
  int besti = 0x7fffffff;
  // rows are only guaranteed 2-byte aligned relative to each other when ld % 8 != 0
  const bool vec_ok = ((reinterpret_cast<uintptr_t>(row) & 15u) == 0);
  const int nvec = vec_ok ? (V >> 3) : 0;
  for (int v = threadIdx.x; v < nvec; v += kArgmaxThreads) {
    float f[8];
    unpack8(ld_16B(reinterpret_cast<const uint4*>(row) + v), f);
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      // NaN never wins here; torch would return a NaN's index, logits are finite in practice
      if (f[e] > best) { best = f[e]; besti = v * 8 + e; }
    }
  }
  for (int i = nvec * 8 + threadIdx.x; i < V; i += kArgmaxThreads) {
    const float f = __bfloat162float(row[i]);
    if (f > best || (f == best && i < besti)) { best = f; besti = i; }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float ov = __shfl_xor_sync(0xffffffffu, best, o);
    const int oi = __shfl_xor_sync(0xffffffffu, besti, o);
    argmax_combine(best, besti, ov, oi);
  }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (lane == 0) { sv[warp] = best; si[warp] = besti; }
  __syncthreads();
  if (warp == 0) {
    best = lane < kArgmaxThreads / 32 ? sv[lane] : -INFINITY;
    besti = lane < kArgmaxThreads / 32 ? si[lane] : 0x7fffffff;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float ov = __shfl_xor_sync(0xffffffffu, best, o);
      const int oi = __shfl_xor_sync(0xffffffffu, besti, o);
      argmax_combine(best, besti, ov, oi);
    }
    if (lane == 0) {
      const int idx = besti == 0x7fffffff ? 0 : besti;
      if (cand) {  // vocabulary-parallel head: (best logit, global token id) of this slice
        cand[blockIdx.x * 2] = best;
        reinterpret_cast<int32_t*>(cand)[blockIdx.x * 2 + 1] = idx + idx_offset;
      } else {
        out[blockIdx.x] = idx;
      }
    }
  }
}

// Bookkeeping between two decode steps, on device so a captured CUDA graph can be
// replayed without host involvement: log the tokens just sampled, then advance
// every sequence's position / cached length and the step counter.
__global__ void decode_advance_kernel(const int32_t* __restrict__ tokens, int32_t* __restrict__ token_log,
                                      int32_t* __restrict__ positions, int32_t* __restrict__ seq_lens,
                                      int32_t* __restrict__ step, int B, int log_capacity) {
  pdl_wait();
  pdl_trigger();
  const int s = *step;
  for (int b = threadIdx.x; b < B; b += blockDim.x) {
    if (token_log && s < log_capacity) token_log[(size_t)s * B + b] = tokens[b];
    // a row with seq_lens == 0 is an idle slot of a continuously batched step: it stays idle
    const bool active = !seq_lens || seq_lens[b] > 0;
    if (positions && active) positions[b] += 1;
    if (seq_lens && active) seq_lens[b] += 1;
  }
  __syncthreads();
  if (threadIdx.x == 0) *step = s + 1;
}

// Global argmax from candidates [n][B][2] = (best logit, index bits) — the per-slice candidates of a
// vocabulary-parallel head, or the per-CTA candidates of the lm_head GEMV's fused epilogue: highest logit,
// lowest index on ties — the same winner a single argmax over the whole row picks.  One warp per row;
// lanes stride over the n candidates, then a shuffle tree.  Writes tokens[b] = idx_offset + index, or
// (cand_out) the row's (logit, idx_offset + index) for a further merge across stages.
__global__ void __launch_bounds__(256)
argmax_merge_kernel(const float* __restrict__ cands, int n, int B, int n_stride, int idx_offset,
                    int32_t* __restrict__ out, float* __restrict__ cand_out) {
  pdl_wait();
  pdl_trigger();
  const int lane = threadIdx.x & 31;
  for (int b = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; b < B; b += (gridDim.x * blockDim.x) >> 5) {
    float best = -INFINITY;
    int besti = 0x7fffffff;
    for (int s = lane; s < n; s += 32) {
      const float v = cands[(size_t)s * n_stride + b * 2];
      const int i = reinterpret_cast<const int32_t*>(cands)[(size_t)s * n_stride + b * 2 + 1];
      if (i != 0x7fffffff && (besti == 0x7fffffff || v > best || (v == best && i < besti))) { best = v; besti = i; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float ov = __shfl_xor_sync(0xffffffffu, best, o);
      const int oi = __shfl_xor_sync(0xffffffffu, besti, o);
      if (oi != 0x7fffffff && (besti == 0x7fffffff || ov > best || (ov == best && oi < besti))) { best = ov; besti = oi; }
    }
    if (lane == 0) {
      const int idx = (besti == 0x7fffffff ? 0 : besti) + idx_offset;
      if (cand_out) {
        cand_out[b * 2] = best;
        reinterpret_cast<int32_t*>(cand_out)[b * 2 + 1] = idx;
      } else {
        out[b] = idx;
      }
    }
  }
}

int launch_argmax_merge(const float* cands, int n, int B, int idx_offset, int32_t* tokens, float* cand_out,
                        cudaStream_t s, int n_stride) {
  const int warps = 8;
  DLI_LAUNCH("argmax_merge_kernel", argmax_merge_kernel, dim3((B + warps - 1) / warps), dim3(warps * 32), (size_t)0, s,
             cands, n, B, n_stride > 0 ? n_stride : 2 * B, idx_offset, tokens, cand_out);
  return DLI_OK;
}

}  // namespace dli

extern "C" {

int dli_argmax_candidate_bf16(const void* logits, int ld, int idx_offset, float* cand, int B, int V,
                              dli_stream_t stream) {
  using namespace dli;
  DLI_CHECK_ARG(logits && cand, "dli_argmax_candidate_bf16: null pointer");
  DLI_CHECK_ARG(B >= 0 && V > 0 && ld >= V && idx_offset >= 0, "dli_argmax_candidate_bf16: bad sizes B=%d V=%d ld=%d", B, V, ld);
  if (B == 0) return DLI_OK;
  DLI_LAUNCH("dli_argmax_candidate_bf16", argmax_kernel, dim3(B), dim3(kArgmaxThreads), (size_t)0,
             static_cast<cudaStream_t>(stream), (const bf16*)logits, ld, (int32_t*)nullptr, V, cand, idx_offset);
  return DLI_OK;
}

int dli_argmax_merge(const float* cands, int n, int B, int32_t* out, dli_stream_t stream) {
  using namespace dli;
  DLI_CHECK_ARG(cands && out && n > 0 && B >= 0, "dli_argmax_merge: bad arguments");
  if (B == 0) return DLI_OK;
  return launch_argmax_merge(cands, n, B, 0, out, nullptr, static_cast<cudaStream_t>(stream), 0);
}

int dli_argmax_merge_strided(const float* cands, int n, int B, int n_stride, int32_t* out, dli_stream_t stream) {
  using namespace dli;
  DLI_CHECK_ARG(cands && out && n > 0 && B >= 0 && n_stride >= 2 * B, "dli_argmax_merge_strided: bad arguments");
  if (B == 0) return DLI_OK;
  return launch_argmax_merge(cands, n, B, 0, out, nullptr, static_cast<cudaStream_t>(stream), n_stride);
}

int dli_decode_advance(const int32_t* tokens, int32_t* token_log, int32_t* positions, int32_t* seq_lens,
                       int32_t* step, int B, int log_capacity, dli_stream_t stream) {
  using namespace dli;
  DLI_CHECK_ARG(step && (tokens || !token_log), "dli_decode_advance: null pointer");
  DLI_CHECK_ARG(B >= 0 && log_capacity >= 0, "dli_decode_advance: bad sizes B=%d cap=%d", B, log_capacity);
  DLI_LAUNCH("dli_decode_advance", decode_advance_kernel, dim3(1), dim3(256), (size_t)0, static_cast<cudaStream_t>(stream),
             tokens, token_log, positions, seq_lens, step, B, log_capacity);
  return DLI_OK;
}

int dli_embedding_bf16(const int32_t* ids, const int32_t* positions, const void* table,
                       const void* pos_table, void* out, int rows, int d, int vocab,
                       dli_stream_t stream) {
  using namespace dli;
  if (rows == 0) return DLI_OK;
  DLI_CHECK_ARG(ids && table && out, "dli_embedding_bf16: null pointer");
  DLI_CHECK_ARG(!pos_table || positions, "dli_embedding_bf16: pos_table needs positions");
  DLI_CHECK_ARG(rows >= 0 && d > 0 && d % 8 == 0 && vocab > 0,
                "dli_embedding_bf16: need d %% 8 == 0 (got rows=%d d=%d vocab=%d)", rows, d, vocab);
  DLI_CHECK_ARG(aligned16(table) && aligned16(pos_table) && aligned16(out),
                "dli_embedding_bf16: pointers must be 16-byte aligned");
  if (rows == 0) return DLI_OK;
  DLI_LAUNCH("dli_embedding_bf16", embedding_kernel, dim3(rows), dim3(128), (size_t)0, static_cast<cudaStream_t>(stream),
             ids, positions, (const bf16*)table, (const bf16*)pos_table, (bf16*)out, d, vocab);
  return DLI_OK;
}

int dli_argmax_bf16(const void* logits, int ld, int32_t* out, int B, int V, dli_stream_t stream) {
  using namespace dli;
  DLI_CHECK_ARG(logits && out, "dli_argmax_bf16: null pointer");
  DLI_CHECK_ARG(B >= 0 && V > 0 && ld >= V, "dli_argmax_bf16: bad sizes B=%d V=%d ld=%d", B, V, ld);
  if (B == 0) return DLI_OK;
  DLI_LAUNCH("dli_argmax_bf16", argmax_kernel, dim3(B), dim3(kArgmaxThreads), (size_t)0, static_cast<cudaStream_t>(stream),
             (const bf16*)logits, ld, out, V, (float*)nullptr, 0);
  return DLI_OK;
}

}  // extern "C"
