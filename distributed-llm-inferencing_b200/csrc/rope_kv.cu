// RoPE (rotate-half) applied in place to q and k of a packed qkv row, fused
// with the paged KV-cache append.  One CTA per token row.
//
// HBM-bound: per row reads (H+2KV)*hd*2 bytes, writes (H+KV)*hd*2 in place and
// 2*KV*hd*2 into the cache pages; cos/sin rows are hd bytes each.
// Rounding follows apply_rotary_pos_emb on bf16 tensors
// (TF:models/llama/modeling_llama.py:138-168): each product and the sum are
// separately rounded to bf16:  q' = bf16(bf16(q*cos) + bf16(rot(q)*sin)).
#include "common.cuh"

namespace dli {

constexpr int kRopeThreads = 128;

__global__ void __launch_bounds__(kRopeThreads)
rope_kv_append_kernel(bf16* __restrict__ qkv, const bf16* __restrict__ cos_t,
                      const bf16* __restrict__ sin_t, const int32_t* __restrict__ positions,
                      const int32_t* __restrict__ seq_ids,
                      const int32_t* __restrict__ page_table, int max_pages,
                      bf16* __restrict__ k_cache, bf16* __restrict__ v_cache, int H, int KV,
                      int hd, int page_shift) {
  const size_t row = blockIdx.x;
  pdl_wait();
  pdl_trigger();
  const int pos = positions[row];
  const int seq = seq_ids[row];
  const int page_size = 1 << page_shift;
  const int page = page_table[(size_t)seq * max_pages + (pos >> page_shift)];
  const int slot = pos & (page_size - 1);
  const int half = hd >> 1;
  const int vph = half >> 3;  // vector pairs per head
  bf16* r = qkv + row * (size_t)(H + 2 * KV) * hd;

  const int n_rot = (H + KV) * vph;
  const int n_v = KV * (hd >> 3);
  for (int it = threadIdx.x; it < n_rot + n_v; it += kRopeThreads) {
    if (it < n_rot) {
      const int head = it / vph, j = it - head * vph;  // head in [0, H+KV): q heads then k heads
      bf16* p = r + (size_t)head * hd + j * 8;
      float x1[8], x2[8], o1[8], o2[8];
      unpack8(ld_16B(p), x1);
      unpack8(ld_16B(p + half), x2);
      if (cos_t) {
        float c[8], s[8];
        unpack8(ld_16B(cos_t + (size_t)pos * half + j * 8), c);
        unpack8(ld_16B(sin_t + (size_t)pos * half + j * 8), s);
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          // rotate_half(x) = cat(-x2, x1); cos/sin repeat over the two halves
          o1[e] = round_bf16(x1[e] * c[e]) + round_bf16(-x2[e] * s[e]);
          o2[e] = round_bf16(x2[e] * c[e]) + round_bf16(x1[e] * s[e]);
        }
      } else {
#pragma unroll
        for (int e = 0; e < 8; ++e) { o1[e] = x1[e]; o2[e] = x2[e]; }
      }
      const uint4 u1 = pack8(o1), u2 = pack8(o2);
      if (cos_t) { st_16B(p, u1); st_16B(p + half, u2); }
      if (head >= H) {
        const int kvh = head - H;
        bf16* dst = k_cache + (((size_t)page * KV + kvh) * page_size + slot) * hd + j * 8;
        st_16B(dst, u1);
        st_16B(dst + half, u2);
      }
    } else {
      const int iv = it - n_rot;
      const int kvh = iv / (hd >> 3), j = iv - kvh * (hd >> 3);
      const uint4 u = ld_16B(r + (size_t)(H + KV + kvh) * hd + j * 8);
      st_16B(v_cache + (((size_t)page * KV + kvh) * page_size + slot) * hd + j * 8, u);
    }
  }
}

}  // namespace dli

extern "C" int dli_rope_kv_append_bf16(void* qkv, const void* cos, const void* sin,
                                       const int32_t* positions, const int32_t* seq_ids,
                                       const int32_t* page_table, int max_pages,
                                       void* k_cache, void* v_cache, int rows, int H, int KV,
                                       int hd, int page_size, dli_stream_t stream) {
  using namespace dli;
  DLI_CHECK_ARG(qkv && positions && seq_ids && page_table && k_cache && v_cache,
                "dli_rope_kv_append_bf16: null pointer");
  DLI_CHECK_ARG((cos == nullptr) == (sin == nullptr),
                "dli_rope_kv_append_bf16: cos and sin must both be given or both NULL");
  DLI_CHECK_ARG(hd == 64 || hd == 128, "dli_rope_kv_append_bf16: hd must be 64 or 128 (got %d)", hd);
  DLI_CHECK_ARG(page_size >= 8 && (page_size & (page_size - 1)) == 0,
                "dli_rope_kv_append_bf16: page_size must be a power of two >= 8 (got %d)", page_size);
  DLI_CHECK_ARG(rows >= 0 && H > 0 && KV > 0 && max_pages > 0,
                "dli_rope_kv_append_bf16: bad sizes rows=%d H=%d KV=%d max_pages=%d", rows, H, KV, max_pages);
  DLI_CHECK_ARG(aligned16(qkv) && aligned16(cos) && aligned16(sin) && aligned16(k_cache) && aligned16(v_cache),
                "dli_rope_kv_append_bf16: pointers must be 16-byte aligned");
  if (rows == 0) return DLI_OK;
  int shift = 0;
  while ((1 << shift) < page_size) ++shift;
  DLI_LAUNCH("dli_rope_kv_append_bf16", rope_kv_append_kernel, dim3(rows), dim3(kRopeThreads), (size_t)0,
             static_cast<cudaStream_t>(stream), (bf16*)qkv, (const bf16*)cos, (const bf16*)sin, positions, seq_ids,
             page_table, max_pages, (bf16*)k_cache, (bf16*)v_cache, H, KV, hd, shift);
  return DLI_OK;
}
