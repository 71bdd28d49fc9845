// Paged-KV decode attention: one query token per sequence, split-KV, GQA-aware.
//
// Grid = (num_splits, KV heads, sequences).  A CTA owns one KV head of one
// sequence and a contiguous range of 64-token KV tiles; the G = H/KV query
// heads that share this KV head form the (zero-padded) 16-row M tile, so every
// K/V byte is read from HBM exactly once for all of them.  K/V tiles are staged
// page by page into shared memory with 16-byte cp.async (each page row segment
// is a contiguous hd*2-byte run, fully coalesced), double-buffered; scores and
// P.V use warp-level m16n8k16 MMAs, the softmax is an online (running max/sum)
// softmax in fp32 with quad shuffles.  The 4 warps split each tile's tokens and
// are merged through shared memory at the end; with num_splits > 1 the CTA
// writes an un-normalised partial (O, max, sum) to scratch and a second tiny
// kernel merges the splits.
//
// Fused variant (cos/sin/new-token pointers given): q is read from the packed qkv row
// of the new token and rotated in shared memory; the CTA that owns the sequence's
// last KV tile also rotates the new token's k, writes k and v into their cache page
// and patches them into its staged tile, so the separate RoPE/append kernel
// disappears from the decode step (bit-identical rounding to rope_kv.cu).
//
// Algorithmic bytes per launch: sum_b min(S_b, window) * KV * hd * 2 (K and V)
// * 2 bytes; q/out/scratch are negligible.
#include <stdlib.h>

#include "mma_utils.cuh"

namespace dli {

constexpr int kAttnThreads = 128;  // split-merge kernel; the decode kernel runs NW warps (template)

// cp.async.bulk.prefetch.L2: one instruction moves a whole piece from HBM into L2.
__device__ __forceinline__ void l2_prefetch_slice(const uint8_t* ptr, size_t bytes, size_t part, size_t nparts) {
  constexpr size_t kPiece = 4096;
  if (ptr == nullptr || bytes == 0) return;
  const size_t per = ((bytes + nparts - 1) / nparts + kPiece - 1) / kPiece * kPiece;
  const size_t begin = part * per;
  const size_t end = begin + per < bytes ? begin + per : bytes;
  for (size_t off = begin + (size_t)threadIdx.x * kPiece; off < end; off += (size_t)blockDim.x * kPiece) {
    const uint32_t n = (uint32_t)((end - off < kPiece ? end - off : kPiece) & ~(size_t)15);
    if (n) asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(ptr + off), "r"(n) : "memory");
  }
}

// NW warps per CTA, each owning 16 tokens of a (NW*16)-token KV tile; NS cp.async stages.
template <int HD, int NS, int NW>
struct DecodeSmem {
  static constexpr int kThreads = NW * 32;
  static constexpr int kTile = NW * 16;
  static constexpr int kRowBytes = HD * 2;
  static constexpr int kQBytes = 16 * kRowBytes;
  static constexpr int kTileBytes = kTile * kRowBytes;
  static constexpr int kStageTotal = kQBytes + 2 * NS * kTileBytes;  // Q + NS stages x (K, V)
  // NS == 4, NW == 4 (the cluster-merge configuration) also holds this CTA's merged partial
  // (16 x HD fp32 + max/sum) where the other CTAs of its cluster can read it (DSMEM)
  static constexpr int kClusterBytes = (NS == 4 && NW == 4) ? 16 * HD * 4 + 128 : 0;
  // cross-warp merge area (re-uses the K/V staging memory): per warp 16 x HD partial O + 16 x (max, sum)
  static constexpr int kRedBytes = NW * 16 * HD * 4 + NW * 16 * 2 * 4;
  static constexpr int kTotal =
      (kStageTotal > kQBytes + kRedBytes ? kStageTotal : kQBytes + kRedBytes) + kClusterBytes;
  static constexpr int kClusterOff = kTotal - kClusterBytes;
};

// ---- thread-block cluster helpers (distributed shared memory) ------------------------------------
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ float ld_dsmem_f32(uint32_t local_smem_addr, uint32_t cta_rank) {
  uint32_t remote;
  float v;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(local_smem_addr), "r"(cta_rank));
  asm volatile("ld.shared::cluster.f32 %0, [%1];" : "=f"(v) : "r"(remote) : "memory");
  return v;
}

// NS = cp.async stages, NW = warps.  Three configurations are used:
//  * NW = 4, NS = 3 (two CTAs per SM, two 64-token tiles each in flight): the large-batch,
//    bandwidth-bound case — 5.5-5.6 TB/s at 32-64 sequences x 2k tokens with ONE split.
//  * NW = 16, NS = 1 / NW = 8, NS = 2 (one CTA per SM at most, a whole 256-token split in flight):
//    the few-CTA case, i.e. a batch-1 step.  That case is not bandwidth- but ISSUE-bound: ncu on
//    the former 4-warp form showed one warp per scheduler, stalled on fixed-latency dependencies
//    between its ~3 000 instructions (address arithmetic of the cp.async staging, ldmatrix, MMA),
//    14 k cycles per CTA for 117 KB of K/V.  Sixteen warps split the same tile four ways finer,
//    so four warps per scheduler cover each other's latencies.
//  * NW = 4, NS = 4: only for the thread-block-cluster merge (no scratch given).
template <int HD, int NS, int NW>
__global__ void __launch_bounds__(NW * 32)
paged_attn_decode_kernel(const bf16* __restrict__ q, int q_stride,
                         const bf16* __restrict__ k_cache, const bf16* __restrict__ v_cache,
                         const int32_t* __restrict__ page_table, int max_pages,
                         const int32_t* __restrict__ seq_lens, bf16* __restrict__ out,
                         float* __restrict__ part_o, float* __restrict__ part_ml, int H, int KV,
                         int page_shift, int num_splits, float scale_log2, int window,
                         int fused, const bf16* __restrict__ cos_t, const bf16* __restrict__ sin_t,
                         bf16* k_cache_w, bf16* v_cache_w, int cluster_merge, const uint8_t* __restrict__ pf_ptr,
                         size_t pf_bytes) {
  using L = DecodeSmem<HD, NS, NW>;
  constexpr int RB = L::kRowBytes;
  constexpr int CH = HD / 8;  // 16-byte chunks per row
  constexpr int kAttnThreads = L::kThreads;  // shadows the merge kernel's constant on purpose
  constexpr int kAttnTile = L::kTile;
  constexpr bool kStreamQ = NW > 8;  // 512 threads leave 128 registers: Q fragments are re-read per use
  extern __shared__ __align__(128) uint8_t smem[];
  const uint32_t sQ = smem_u32(smem);
  const uint32_t sK = sQ + L::kQBytes;
  const uint32_t sV = sK + NS * L::kTileBytes;

  const int split = blockIdx.x, kvh = blockIdx.y, b = blockIdx.z;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int G = H / KV;
  if (split >= num_splits) {
    // Prefetch role (extra CTAs on SMs the attention grid leaves idle): pull the weights of
    // the NEXT projection into L2 while this latency-bound kernel runs, so HBM keeps
    // streaming.  Weights are constants: no dependency wait needed.
    l2_prefetch_slice(pf_ptr, pf_bytes, ((size_t)b * KV + kvh) * (gridDim.x - num_splits) + (split - num_splits),
                      (size_t)(gridDim.x - num_splits) * KV * gridDim.z);
    return;
  }
  // Fused form: only q and the new token's k, v come from the kernel just before this one (the qkv projection);
  // the sequence length, the page table and the cache pages of the OLD tokens were written at least one kernel
  // earlier, and that kernel is complete once this one runs at all (every kernel of the library waits for its
  // predecessor before it lets its successor start).  So the dependency wait moves below the K/V requests: the
  // sequence-length -> page-table -> K/V chain (2.7 us of the 6.5 us this kernel took after the wait at batch 1,
  // scratch/trace_decode_run.py) runs while the projection drains.  The unfused form reads K/V that the rope
  // kernel just before it appended: it waits first.
  if (!fused) {
    pdl_wait();
    pdl_trigger();
  }
  const int seq_len = seq_lens[b];
  const int first_tok = (window > 0 && seq_len > window) ? seq_len - window : 0;
  const int tile_lo = first_tok / kAttnTile;
  const int tile_hi = (seq_len + kAttnTile - 1) / kAttnTile;
  const int tps = (tile_hi - tile_lo + num_splits - 1) / num_splits;
  const int t_begin = tile_lo + split * tps;
  const int t_end = min(t_begin + tps, tile_hi);
  const int page_mask = (1 << page_shift) - 1;
  const int32_t* pt = page_table + (size_t)b * max_pages;

  auto load_tile = [&](int tile, int stage) {
    const uint32_t dK = sK + stage * L::kTileBytes, dV = sV + stage * L::kTileBytes;
#pragma unroll
    for (int i = 0; i < kAttnTile * CH / kAttnThreads; ++i) {
      const int id = tid + i * kAttnThreads;
      const int row = id / CH, c = id - row * CH;
      const int tok = tile * kAttnTile + row;
      const bool valid = tok < seq_len;
      size_t off = 0;
      if (valid) {
        const int page = pt[tok >> page_shift];
        off = ((((size_t)page * KV + kvh) << page_shift) + (tok & page_mask)) * HD + c * 8;
      }
      cp_async_16B(dK + swz<RB>(row, c), k_cache + off, valid ? 16 : 0);
      cp_async_16B(dV + swz<RB>(row, c), v_cache + off, valid ? 16 : 0);
    }
  };

  float m_run[2] = {-INFINITY, -INFINITY};  // rows g and g+8 (raw score units)
  float l_run[2] = {0.f, 0.f};
  float o[HD / 8][4];
#pragma unroll
  for (int i = 0; i < HD / 8; ++i) o[i][0] = o[i][1] = o[i][2] = o[i][3] = 0.f;
  uint32_t qf[kStreamQ ? 1 : HD / 16][4];

  // one commit group per stage (empty groups keep the count uniform)
#pragma unroll
  for (int st = 0; st < NS; ++st) {
    if (t_begin + st < t_end) load_tile(t_begin + st, st);
    cp_async_commit();
  }
  if (fused) {
    pdl_wait();  // q and the new token's k, v
    pdl_trigger();
  }
  // ---- stage Q (rows >= G are zero): its own commit group, the newest; the first tile waits for all groups ----
  for (int id = tid; id < 16 * CH; id += kAttnThreads) {
    const int row = id / CH, c = id - row * CH;
    const bool valid = row < G;
    const bf16* src = q + (size_t)b * q_stride + (size_t)(kvh * G + (valid ? row : 0)) * HD + c * 8;
    cp_async_16B(sQ + swz<RB>(row, c), src, valid ? 16 : 0);
  }
  cp_async_commit();

  // Fused form: cos/sin of the new position, and on the CTA that owns the last tile the new token's k, v and cache
  // page, are needed only once the first K/V tile has landed.  Requested here, their round trips hide under that
  // wait instead of following it (the kernel is bound by its chain of dependent loads, not by bandwidth).
  // Every thread rotates elements i = tid % (HD/2) only (the block size is a multiple of HD/2), so one pair does.
  float pre_c = 1.f, pre_s = 0.f, pre_k1 = 0.f, pre_k2 = 0.f;
  uint4 pre_v = make_uint4(0u, 0u, 0u, 0u);
  int pre_page = 0;
  if (fused && t_begin < t_end) {
    const int pos = seq_len - 1, half = HD / 2;
    if (cos_t) {
      pre_c = __bfloat162float(cos_t[(size_t)pos * half + (tid % half)]);
      pre_s = __bfloat162float(sin_t[(size_t)pos * half + (tid % half)]);
    }
    if (t_end == tile_hi && tid < half + CH) {
      pre_page = pt[pos >> page_shift];
      if (tid < half) {
        const bf16* knew = q + (size_t)b * q_stride + (size_t)(H + kvh) * HD;
        pre_k1 = __bfloat162float(knew[tid]);
        pre_k2 = __bfloat162float(knew[tid + half]);
      } else {
        pre_v = ld_16B(q + (size_t)b * q_stride + (size_t)(H + KV + kvh) * HD + (tid - half) * 8);
      }
    }
  }

  for (int tile = t_begin; tile < t_end; ++tile) {
    const int stage = (tile - t_begin) % NS;
    if (tile == t_begin) cp_async_wait<0>();  // Q (the newest group) and every stage requested so far
    else cp_async_wait<NS - 1>();             // all but the newest NS-1 groups have landed: this tile is in
    __syncthreads();
    const uint32_t cK = sK + stage * L::kTileBytes, cV = sV + stage * L::kTileBytes;
    if (fused && tile == tile_hi - 1) {
      // this CTA owns the tile of the new token (position seq_len-1): rotate its k, store
      // k and v to the cache page and patch them into the staged tile
      const int pos = seq_len - 1, row = pos - tile * kAttnTile, half = HD / 2;
      const size_t coff = ((((size_t)pre_page * KV + kvh) << page_shift) + (pos & page_mask)) * HD;
      if (tid < half) {
        const int i = tid;
        float x1 = pre_k1, x2 = pre_k2;
        if (cos_t) {
          const float c = pre_c, sn = pre_s;
          const float o1 = round_bf16(x1 * c) + round_bf16(-x2 * sn);
          const float o2 = round_bf16(x2 * c) + round_bf16(x1 * sn);
          x1 = o1; x2 = o2;
        }
        const bf16 b1 = __float2bfloat16_rn(x1), b2 = __float2bfloat16_rn(x2);
        k_cache_w[coff + i] = b1;
        k_cache_w[coff + i + half] = b2;
        *reinterpret_cast<bf16*>(smem + (cK - sQ) + swz<RB>(row, i >> 3) + (i & 7) * 2) = b1;
        *reinterpret_cast<bf16*>(smem + (cK - sQ) + swz<RB>(row, (i + half) >> 3) + (i & 7) * 2) = b2;
      } else if (tid < half + CH) {
        const int c = tid - half;
        const uint4 u = pre_v;
        st_16B(v_cache_w + coff + c * 8, u);
        *reinterpret_cast<uint4*>(smem + (cV - sQ) + swz<RB>(row, c)) = u;
      }
      __syncthreads();
    }
    if (tile == t_begin) {
      if (fused && cos_t) {
        // rotate the G query rows in shared memory (rotate-half, bf16 op by op)
        const int half = HD / 2;
        for (int idx = tid; idx < G * half; idx += kAttnThreads) {
          const int g = idx / half, i = idx - g * half;
          bf16* p1 = reinterpret_cast<bf16*>(smem + swz<RB>(g, i >> 3) + (i & 7) * 2);
          bf16* p2 = reinterpret_cast<bf16*>(smem + swz<RB>(g, (i + half) >> 3) + (i & 7) * 2);
          const float x1 = __bfloat162float(*p1), x2 = __bfloat162float(*p2);
          const float c = pre_c, sn = pre_s;  // i == tid % half on every iteration
          *p1 = __float2bfloat16_rn(round_bf16(x1 * c) + round_bf16(-x2 * sn));
          *p2 = __float2bfloat16_rn(round_bf16(x2 * c) + round_bf16(x1 * sn));
        }
        __syncthreads();
      }
      if (!kStreamQ) {
#pragma unroll
        for (int kk = 0; kk < HD / 16; ++kk) {
          const int mi = lane >> 3;
          ldmatrix_x4(qf[kStreamQ ? 0 : kk], sQ + swz<RB>((mi & 1) * 8 + (lane & 7), kk * 2 + (mi >> 1)));
        }
      }
    }

    // ---- S = Q K^T for this warp's 16 tokens ---------------------------------------
    float s[2][4];
#pragma unroll
    for (int j = 0; j < 2; ++j) s[j][0] = s[j][1] = s[j][2] = s[j][3] = 0.f;
#pragma unroll
    for (int kk = 0; kk < HD / 16; ++kk) {
      uint32_t kf[4];
      const int mi = lane >> 3;
      ldmatrix_x4(kf, cK + swz<RB>(warp * 16 + (mi >> 1) * 8 + (lane & 7), kk * 2 + (mi & 1)));
      if (kStreamQ) ldmatrix_x4(qf[0], sQ + swz<RB>((mi & 1) * 8 + (lane & 7), kk * 2 + (mi >> 1)));
      mma_16816(s[0], qf[kStreamQ ? 0 : kk], kf[0], kf[1]);
      mma_16816(s[1], qf[kStreamQ ? 0 : kk], kf[2], kf[3]);
    }
    // ---- mask + online softmax --------------------------------------------------------
    const int tok0 = tile * kAttnTile + warp * 16 + 2 * (lane & 3);
#pragma unroll
    for (int j = 0; j < 2; ++j)
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const int tok = tok0 + j * 8 + (e & 1);
        if (tok >= seq_len || tok < first_tok) s[j][e] = -INFINITY;
      }
    uint32_t pa[4];
#pragma unroll
    for (int r = 0; r < 2; ++r) {
      float mx = fmaxf(fmaxf(s[0][2 * r], s[0][2 * r + 1]), fmaxf(s[1][2 * r], s[1][2 * r + 1]));
      mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 1));
      mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 2));
      const float m_new = fmaxf(m_run[r], mx);
      const float m_use = (m_new == -INFINITY) ? 0.f : m_new;
      const float corr = ex2_ftz((m_run[r] - m_use) * scale_log2);
      m_run[r] = m_new;
      float p[4];
      p[0] = ex2_ftz((s[0][2 * r] - m_use) * scale_log2);
      p[1] = ex2_ftz((s[0][2 * r + 1] - m_use) * scale_log2);
      p[2] = ex2_ftz((s[1][2 * r] - m_use) * scale_log2);
      p[3] = ex2_ftz((s[1][2 * r + 1] - m_use) * scale_log2);
      l_run[r] = l_run[r] * corr + (p[0] + p[1] + p[2] + p[3]);
      pa[r] = pack_bf16(p[0], p[1]);      // a0 (row g) / a1 (row g+8): tokens 2t,2t+1
      pa[2 + r] = pack_bf16(p[2], p[3]);  // a2 / a3: tokens 8+2t, 8+2t+1
#pragma unroll
      for (int i = 0; i < HD / 8; ++i) {
        o[i][2 * r] *= corr;
        o[i][2 * r + 1] *= corr;
      }
    }
    // ---- O += P V ------------------------------------------------------------------------
#pragma unroll
    for (int jp = 0; jp < HD / 16; ++jp) {
      uint32_t vf[4];
      const int mi = lane >> 3;
      ldmatrix_x4_trans(vf, cV + swz<RB>(warp * 16 + (mi & 1) * 8 + (lane & 7), jp * 2 + (mi >> 1)));
      mma_16816(o[2 * jp], pa, vf[0], vf[1]);
      mma_16816(o[2 * jp + 1], pa, vf[2], vf[3]);
    }
    __syncthreads();  // every warp is done with this stage: refill it
    if (tile + NS < t_end) load_tile(tile + NS, stage);
    cp_async_commit();
  }
  cp_async_wait<0>();
  __syncthreads();

  // ---- merge the 4 warps (reuse the K/V staging memory) ------------------------------
#pragma unroll
  for (int r = 0; r < 2; ++r) {
    l_run[r] += __shfl_xor_sync(0xffffffffu, l_run[r], 1);
    l_run[r] += __shfl_xor_sync(0xffffffffu, l_run[r], 2);
  }
  float* red_o = reinterpret_cast<float*>(smem + L::kQBytes);  // [NW][16][HD]
  float* red_ml = red_o + NW * 16 * HD;                        // [NW][16][2]
  {
    const int g = lane >> 2, t = lane & 3;
#pragma unroll
    for (int i = 0; i < HD / 8; ++i) {  // only the G query rows are ever read back
      float* p0 = red_o + ((size_t)(warp * 16 + g) * HD + i * 8 + 2 * t);
      float* p1 = red_o + ((size_t)(warp * 16 + g + 8) * HD + i * 8 + 2 * t);
      if (g < G) { p0[0] = o[i][0]; p0[1] = o[i][1]; }
      if (g + 8 < G) { p1[0] = o[i][2]; p1[1] = o[i][3]; }
    }
    if (t == 0) {
      red_ml[(warp * 16 + g) * 2 + 0] = m_run[0];
      red_ml[(warp * 16 + g) * 2 + 1] = l_run[0];
      red_ml[(warp * 16 + g + 8) * 2 + 0] = m_run[1];
      red_ml[(warp * 16 + g + 8) * 2 + 1] = l_run[1];
    }
  }
  __syncthreads();
  for (int idx = tid; idx < G * HD; idx += kAttnThreads) {
    const int row = idx / HD, col = idx - row * HD;
    float mmax = -INFINITY;
#pragma unroll
    for (int w = 0; w < NW; ++w) mmax = fmaxf(mmax, red_ml[(w * 16 + row) * 2]);
    float acc = 0.f, lsum = 0.f;
    if (mmax != -INFINITY) {
#pragma unroll
      for (int w = 0; w < NW; ++w) {
        const float f = ex2_ftz((red_ml[(w * 16 + row) * 2] - mmax) * scale_log2);
        acc += f * red_o[(size_t)(w * 16 + row) * HD + col];
        lsum += f * red_ml[(w * 16 + row) * 2 + 1];
      }
    }
    const int h = kvh * G + row;
    if (num_splits == 1) {
      out[((size_t)b * H + h) * HD + col] = __float2bfloat16_rn(lsum > 0.f ? acc / lsum : 0.f);
    } else if (cluster_merge) {
      float* cpo = reinterpret_cast<float*>(smem + L::kClusterOff);
      cpo[idx] = acc;
      if (col == 0) {
        cpo[16 * HD + row * 2 + 0] = mmax;
        cpo[16 * HD + row * 2 + 1] = lsum;
      }
    } else {
      const size_t slot = ((size_t)b * H + h) * num_splits + split;
      part_o[slot * HD + col] = acc;
      if (col == 0) {
        part_ml[slot * 2 + 0] = mmax;
        part_ml[slot * 2 + 1] = lsum;
      }
    }
  }
  if (num_splits > 1 && cluster_merge) {
    // The num_splits CTAs of this (sequence, kv head) are one thread-block cluster: after a
    // cluster barrier each CTA merges its share of the G*HD outputs by reading every CTA's
    // partial through distributed shared memory, in split order (deterministic) — no scratch
    // buffer in HBM, no second kernel.
    cluster_sync_all();
    const uint32_t cpo_addr = smem_u32(smem + L::kClusterOff);
    const int total = G * HD;
    const int per = (total + num_splits - 1) / num_splits;
    const int lo = split * per, hi = min(lo + per, total);
    for (int idx = lo + tid; idx < hi; idx += kAttnThreads) {
      const int row = idx / HD, col = idx - row * HD;
      float mmax = -INFINITY;
      for (int sp = 0; sp < num_splits; ++sp)
        mmax = fmaxf(mmax, ld_dsmem_f32(cpo_addr + (16 * HD + row * 2) * 4, sp));
      float acc = 0.f, lsum = 0.f;
      if (mmax != -INFINITY) {
        for (int sp = 0; sp < num_splits; ++sp) {
          const float f = ex2_ftz((ld_dsmem_f32(cpo_addr + (16 * HD + row * 2) * 4, sp) - mmax) * scale_log2);
          acc += f * ld_dsmem_f32(cpo_addr + idx * 4, sp);
          lsum += f * ld_dsmem_f32(cpo_addr + (16 * HD + row * 2 + 1) * 4, sp);
        }
      }
      out[((size_t)b * H + kvh * G + row) * HD + col] = __float2bfloat16_rn(lsum > 0.f ? acc / lsum : 0.f);
    }
    cluster_sync_all();  // nobody leaves while a peer may still read its shared memory
  }
}

// Merge the per-split partials: one warp per (sequence, head).  Lanes first take one
// split each (max / weights / denominator by shuffles), then every lane accumulates its
// HD/32 output columns over the splits with independent, unrolled loads.
template <int HD>
__global__ void __launch_bounds__(kAttnThreads)
attn_split_merge_kernel(const float* __restrict__ part_o, const float* __restrict__ part_ml,
                        bf16* __restrict__ out, int n_bh, int num_splits, float scale_log2, int merge_ctas,
                        const uint8_t* __restrict__ pf_ptr, size_t pf_bytes) {
  constexpr int VEC = HD / 32;
  if ((int)blockIdx.x >= merge_ctas) {  // prefetch role, see paged_attn_decode_kernel
    l2_prefetch_slice(pf_ptr, pf_bytes, blockIdx.x - merge_ctas, gridDim.x - merge_ctas);
    return;
  }
  const int bh = blockIdx.x * (kAttnThreads / 32) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  pdl_wait();
  pdl_trigger();
  if (bh >= n_bh) return;
  const float* ml = part_ml + (size_t)bh * num_splits * 2;
  const float* po = part_o + (size_t)bh * num_splits * HD + lane * VEC;
  // the first 8 splits' partial outputs are requested together with the (max, sum) pairs: one L2 round trip for
  // a 2k-context merge (9 splits) instead of two dependent ones
  constexpr int kPre = 8;
  float pv[kPre][VEC];
#pragma unroll
  for (int s = 0; s < kPre; ++s) {
    if (s < num_splits) {
      if (VEC == 4) {
        const float4 t = *reinterpret_cast<const float4*>(po + (size_t)s * HD);
        pv[s][0] = t.x; pv[s][1] = t.y; pv[s][2] = t.z; pv[s][VEC - 1] = t.w;
      } else {
        const float2 t = *reinterpret_cast<const float2*>(po + (size_t)s * HD);
        pv[s][0] = t.x; pv[s][VEC - 1] = t.y;
      }
    } else {
#pragma unroll
      for (int i = 0; i < VEC; ++i) pv[s][i] = 0.f;
    }
  }
  // up to 64 splits: lane owns splits lane and lane + 32
  float m0 = -INFINITY, l0 = 0.f, m1 = -INFINITY, l1 = 0.f;
  if (lane < num_splits) { m0 = ml[lane * 2]; l0 = ml[lane * 2 + 1]; }
  if (lane + 32 < num_splits) { m1 = ml[(lane + 32) * 2]; l1 = ml[(lane + 32) * 2 + 1]; }
  const float mmax = warp_max(fmaxf(m0, m1));
  float f0 = 0.f, f1 = 0.f;
  if (mmax != -INFINITY) {
    f0 = ex2_ftz((m0 - mmax) * scale_log2);
    f1 = ex2_ftz((m1 - mmax) * scale_log2);
  }
  const float lsum = warp_sum(f0 * l0 + f1 * l1);
  float acc[VEC];
#pragma unroll
  for (int i = 0; i < VEC; ++i) acc[i] = 0.f;
#pragma unroll
  for (int s = 0; s < kPre; ++s) {  // (splits beyond num_splits hold zeros and a zero weight)
    const float f = __shfl_sync(0xffffffffu, f0, s);
#pragma unroll
    for (int i = 0; i < VEC; ++i) acc[i] += (s < num_splits ? f : 0.f) * pv[s][i];
  }
#pragma unroll 8
  for (int s = kPre; s < num_splits; ++s) {
    const float f = __shfl_sync(0xffffffffu, s < 32 ? f0 : f1, s & 31);
    float v[VEC];
    if (VEC == 4) {
      const float4 t = *reinterpret_cast<const float4*>(po + (size_t)s * HD);
      v[0] = t.x; v[1] = t.y; v[2] = t.z; v[VEC - 1] = t.w;
    } else {
      const float2 t = *reinterpret_cast<const float2*>(po + (size_t)s * HD);
      v[0] = t.x; v[VEC - 1] = t.y;
    }
#pragma unroll
    for (int i = 0; i < VEC; ++i) acc[i] += f * v[i];
  }
  const float inv = lsum > 0.f ? 1.f / lsum : 0.f;
  bf16* o = out + (size_t)bh * HD + lane * VEC;
#pragma unroll
  for (int i = 0; i < VEC; i += 2)
    *reinterpret_cast<uint32_t*>(o + i) = pack_bf16(acc[i] * inv, acc[i + 1] * inv);
}

template <int HD, int NS, int NW>
static int launch_decode_ns(const void* q, int q_stride, const void* k_cache, const void* v_cache,
                            const int32_t* page_table, int max_pages, const int32_t* seq_lens, void* out,
                            void* scratch, int B, int H, int KV, int page_shift, int num_splits, float scale,
                            int window, cudaStream_t s, int fused, const void* cos_t, const void* sin_t,
                            const void* pf_a, size_t pf_a_bytes, const void* pf_b, size_t pf_b_bytes) {
  using L = DecodeSmem<HD, NS, NW>;
  static PerDeviceOnce once;
  DLI_ONCE_PER_DEVICE(once, dev, DLI_CUDA_CALL(cudaFuncSetAttribute(paged_attn_decode_kernel<HD, NS, NW>,
                                                                    cudaFuncAttributeMaxDynamicSharedMemorySize,
                                                                    L::kTotal)));
  const float scale_log2 = scale * 1.4426950408889634f;
  float* part_o = reinterpret_cast<float*>(scratch);
  float* part_ml = part_o ? part_o + (size_t)B * H * num_splits * HD : nullptr;
  const int sms = sm_count_cached();
  const int attn_ctas = num_splits * KV * B;
  if constexpr (NS == 4 && NW == 4) {
    if (num_splits > 1 && num_splits <= 8 && scratch == nullptr) {
      // no scratch given: the splits of a (sequence, kv head) run as one thread-block cluster and
      // merge through distributed shared memory inside the kernel
      cudaError_t e = launch_kernel_cluster(
          paged_attn_decode_kernel<HD, NS, NW>, dim3(num_splits, KV, B), dim3(L::kThreads), (size_t)L::kTotal, s,
          (unsigned)num_splits, (const bf16*)q, q_stride, (const bf16*)k_cache, (const bf16*)v_cache, page_table,
          max_pages, seq_lens, (bf16*)out, (float*)nullptr, (float*)nullptr, H, KV, page_shift, num_splits, scale_log2,
          window, fused, (const bf16*)cos_t, (const bf16*)sin_t, (bf16*)k_cache, (bf16*)v_cache, 1,
          (const uint8_t*)nullptr, (size_t)0);
      if (e != cudaSuccess) return fail(DLI_ERR_CUDA, "paged_attn_decode_kernel (cluster): %s", cudaGetErrorString(e));
      return DLI_OK;
    }
  }
  // prefetch CTAs only when the attention grid leaves SMs idle
  int extra = 0;
  if (pf_a && pf_a_bytes && attn_ctas < sms) extra = (sms - attn_ctas) / (KV * B);
  dim3 grid(num_splits + extra, KV, B);
  DLI_LAUNCH("paged_attn_decode_kernel", paged_attn_decode_kernel<HD, NS, NW>, grid, dim3(L::kThreads),
             (size_t)L::kTotal, s, (const bf16*)q, q_stride, (const bf16*)k_cache, (const bf16*)v_cache, page_table,
             max_pages, seq_lens, (bf16*)out, part_o, part_ml, H, KV, page_shift, num_splits, scale_log2, window,
             fused, (const bf16*)cos_t, (const bf16*)sin_t, (bf16*)k_cache, (bf16*)v_cache, 0,
             (const uint8_t*)(extra ? pf_a : nullptr), extra ? pf_a_bytes : (size_t)0);
  if (num_splits > 1) {
    const int n_bh = B * H;
    const int per_cta = kAttnThreads / 32;
    const int merge_ctas = (n_bh + per_cta - 1) / per_cta;
    const int pf_ctas = (pf_b && pf_b_bytes && merge_ctas < sms) ? sms - merge_ctas : 0;
    DLI_LAUNCH("attn_split_merge_kernel", attn_split_merge_kernel<HD>, dim3(merge_ctas + pf_ctas), dim3(kAttnThreads),
               (size_t)0, s, (const float*)part_o, (const float*)part_ml, (bf16*)out, n_bh, num_splits, scale_log2,
               merge_ctas, (const uint8_t*)(pf_ctas ? pf_b : nullptr), pf_ctas ? pf_b_bytes : (size_t)0);
  }
  return DLI_OK;
}

// DLI_ATTN_WIDE: warps per CTA of the few-CTA configuration (0 / 4 = the former 4-warp, 4-stage kernel;
// 8; 16; default: chosen from the split length).  Measurement knob.
static int attn_wide_env() {
  static const int v = getenv("DLI_ATTN_WIDE") ? atoi(getenv("DLI_ATTN_WIDE")) : -1;
  return v;
}

template <int HD>
static int launch_decode(const void* q, int q_stride, const void* k_cache, const void* v_cache,
                         const int32_t* page_table, int max_pages, const int32_t* seq_lens,
                         void* out, void* scratch, int B, int H, int KV, int page_shift,
                         int num_splits, float scale, int window, cudaStream_t s, int fused = 0,
                         const void* cos_t = nullptr, const void* sin_t = nullptr, const void* pf_a = nullptr,
                         size_t pf_a_bytes = 0, const void* pf_b = nullptr, size_t pf_b_bytes = 0) {
#define DLI_DECODE_ARGS                                                                                            \
  q, q_stride, k_cache, v_cache, page_table, max_pages, seq_lens, out, scratch, B, H, KV, page_shift, num_splits, \
      scale, window, s, fused, cos_t, sin_t, pf_a, pf_a_bytes, pf_b, pf_b_bytes
  // many CTAs: occupancy (two 4-warp CTAs per SM, 3 stages)
  if ((size_t)num_splits * KV * B > (size_t)sm_count_cached()) return launch_decode_ns<HD, 3, 4>(DLI_DECODE_ARGS);
  // few CTAs (batch-1 style): at most one CTA per SM, the whole split in flight, many warps
  const bool cluster = num_splits > 1 && num_splits <= 8 && scratch == nullptr;
  int wide = attn_wide_env();
  if (wide < 0) {
    // longest split the page table allows, in tokens: one 256-token tile -> 16 warps on a single stage;
    // longer splits loop over 128-token tiles with two stages
    const long long max_tokens = ((long long)max_pages << page_shift);
    const long long bound = window > 0 && window + 256 < max_tokens ? window + 256 : max_tokens;
    wide = (bound + num_splits - 1) / num_splits <= 256 ? 16 : 8;
  }
  if (cluster || wide <= 4) return launch_decode_ns<HD, 4, 4>(DLI_DECODE_ARGS);
  if (wide >= 16) return launch_decode_ns<HD, 1, 16>(DLI_DECODE_ARGS);
  return launch_decode_ns<HD, 2, 8>(DLI_DECODE_ARGS);
#undef DLI_DECODE_ARGS
}

}  // namespace dli

extern "C" {

size_t dli_paged_attn_decode_scratch_bytes(int B, int H, int hd, int num_splits) {
  if (B <= 0 || H <= 0 || hd <= 0 || num_splits <= 1) return 0;
  return sizeof(float) * (size_t)B * H * num_splits * (hd + 2);
}

int dli_paged_attn_decode_bf16(const void* q, int q_stride, const void* k_cache,
                               const void* v_cache, const int32_t* page_table, int max_pages,
                               const int32_t* seq_lens, void* out, void* scratch,
                               size_t scratch_bytes, int B, int H, int KV, int hd, int page_size,
                               int num_splits, float scale, int window, dli_stream_t stream) {
  using namespace dli;
  DLI_CHECK_ARG(q && k_cache && v_cache && page_table && seq_lens && out,
                "dli_paged_attn_decode_bf16: null pointer");
  DLI_CHECK_ARG(hd == 64 || hd == 128, "dli_paged_attn_decode_bf16: hd must be 64 or 128 (got %d)", hd);
  DLI_CHECK_ARG(B >= 0 && H > 0 && KV > 0 && H % KV == 0 && H / KV <= 16,
                "dli_paged_attn_decode_bf16: need H %% KV == 0 and H/KV <= 16 (got H=%d KV=%d)", H, KV);
  DLI_CHECK_ARG(page_size >= 8 && (page_size & (page_size - 1)) == 0,
                "dli_paged_attn_decode_bf16: page_size must be a power of two >= 8 (got %d)", page_size);
  DLI_CHECK_ARG(num_splits >= 1 && num_splits <= 64 && KV <= 65535 && B <= 65535,
                "dli_paged_attn_decode_bf16: num_splits must be 1..64 (got %d)", num_splits);
  DLI_CHECK_ARG(q_stride >= H * hd && q_stride % 8 == 0 && max_pages > 0 && window >= 0,
                "dli_paged_attn_decode_bf16: bad q_stride/max_pages/window");
  DLI_CHECK_ARG(aligned16(q) && aligned16(k_cache) && aligned16(v_cache) && aligned16(scratch),
                "dli_paged_attn_decode_bf16: pointers must be 16-byte aligned");

  if (num_splits > 1) {
    const size_t need = dli_paged_attn_decode_scratch_bytes(B, H, hd, num_splits);
    const bool cluster_ok = scratch == nullptr && num_splits <= 8 && num_splits * KV * B <= sm_count_cached();
    DLI_CHECK_ARG(cluster_ok || (scratch && scratch_bytes >= need),
                  "dli_paged_attn_decode_bf16: scratch too small (%zu < %zu bytes); scratch may only be NULL for <= 8 splits and "
                  "num_splits*KV*B <= number of SMs (cluster merge)", scratch_bytes, need);
  }
  if (B == 0) return DLI_OK;
  int shift = 0;
  while ((1 << shift) < page_size) ++shift;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (hd == 128)
    return launch_decode<128>(q, q_stride, k_cache, v_cache, page_table, max_pages, seq_lens, out,
                              scratch, B, H, KV, shift, num_splits, scale, window, s);
  return launch_decode<64>(q, q_stride, k_cache, v_cache, page_table, max_pages, seq_lens, out,
                           scratch, B, H, KV, shift, num_splits, scale, window, s);
}

}  // extern "C"

extern "C" int dli_paged_attn_decode_fused_bf16(void* qkv, int row_stride, const void* cos, const void* sin,
                                                void* k_cache, void* v_cache, const int32_t* page_table,
                                                int max_pages, const int32_t* seq_lens, void* out, void* scratch,
                                                size_t scratch_bytes, int B, int H, int KV, int hd, int page_size,
                                                int num_splits, float scale, int window, const void* l2_prefetch_a,
                                                size_t l2_prefetch_a_bytes, const void* l2_prefetch_b,
                                                size_t l2_prefetch_b_bytes, dli_stream_t stream) {
  using namespace dli;
  DLI_CHECK_ARG(qkv && k_cache && v_cache && page_table && seq_lens && out,
                "dli_paged_attn_decode_fused_bf16: null pointer");
  DLI_CHECK_ARG(aligned16(l2_prefetch_a) && aligned16(l2_prefetch_b),
                "dli_paged_attn_decode_fused_bf16: prefetch ranges must be 16-byte aligned");
  DLI_CHECK_ARG((cos == nullptr) == (sin == nullptr), "dli_paged_attn_decode_fused_bf16: cos and sin go together");
  DLI_CHECK_ARG(hd == 64 || hd == 128, "dli_paged_attn_decode_fused_bf16: hd must be 64 or 128 (got %d)", hd);
  DLI_CHECK_ARG(B >= 0 && H > 0 && KV > 0 && H % KV == 0 && H / KV <= 16,
                "dli_paged_attn_decode_fused_bf16: need H %% KV == 0 and H/KV <= 16 (got H=%d KV=%d)", H, KV);
  DLI_CHECK_ARG(page_size >= 8 && (page_size & (page_size - 1)) == 0,
                "dli_paged_attn_decode_fused_bf16: page_size must be a power of two >= 8 (got %d)", page_size);
  DLI_CHECK_ARG(num_splits >= 1 && num_splits <= 64 && KV <= 65535 && B <= 65535,
                "dli_paged_attn_decode_fused_bf16: num_splits must be 1..64");
  DLI_CHECK_ARG(row_stride >= (H + 2 * KV) * hd && row_stride % 8 == 0 && max_pages > 0 && window >= 0,
                "dli_paged_attn_decode_fused_bf16: bad row_stride/max_pages/window");
  DLI_CHECK_ARG(aligned16(qkv) && aligned16(k_cache) && aligned16(v_cache) && aligned16(scratch) &&
                    aligned16(cos) && aligned16(sin),
                "dli_paged_attn_decode_fused_bf16: pointers must be 16-byte aligned");

  if (num_splits > 1) {
    const size_t need = dli_paged_attn_decode_scratch_bytes(B, H, hd, num_splits);
    const bool cluster_ok = scratch == nullptr && num_splits <= 8 && num_splits * KV * B <= sm_count_cached();
    DLI_CHECK_ARG(cluster_ok || (scratch && scratch_bytes >= need),
                  "dli_paged_attn_decode_fused_bf16: scratch too small (%zu < %zu bytes); scratch may only be NULL for <= 8 splits and "
                  "num_splits*KV*B <= number of SMs (cluster merge)", scratch_bytes, need);
  }
  if (B == 0) return DLI_OK;
  int shift = 0;
  while ((1 << shift) < page_size) ++shift;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (hd == 128)
    return launch_decode<128>(qkv, row_stride, k_cache, v_cache, page_table, max_pages, seq_lens, out, scratch, B,
                              H, KV, shift, num_splits, scale, window, s, 1, cos, sin, l2_prefetch_a,
                              l2_prefetch_a_bytes, l2_prefetch_b, l2_prefetch_b_bytes);
  return launch_decode<64>(qkv, row_stride, k_cache, v_cache, page_table, max_pages, seq_lens, out, scratch, B, H,
                           KV, shift, num_splits, scale, window, s, 1, cos, sin, l2_prefetch_a, l2_prefetch_a_bytes,
                           l2_prefetch_b, l2_prefetch_b_bytes);
}

// ---------------------------------------------------------------------------------------
// Split-KV decode attention for callers that cannot allocate device memory.
//
// Some users of this library hold only the buffers they were handed: a serving process
// that gave all of HBM to its weight and KV pools at start-up (a later cudaMalloc fails or
// forces an eviction), a foreign-language binding of this C-ABI with no device allocator of
// its own, or a per-token loop that cannot afford the device-wide synchronisation
// cudaMalloc/cudaFree imply.  For them the split partials live in the fixed array below
// instead of a caller-provided scratch buffer.  The array is part of the library image, so
// every GPU that loads the library holds these bytes until the process exits.
constexpr size_t kStaticScratchBytes = 8u << 20;  // 8 MiB: e.g. B=32, H=32, hd=128, 15 splits
static __device__ __align__(256) unsigned char g_static_attn_scratch[kStaticScratchBytes];

namespace dli {

// In host code `g_static_attn_scratch` is only a host-side stand-in for the symbol, not a
// device address, so the address is looked up once per device (with that device current)
// and cached.  Returns nullptr after recording the error.
static void* static_scratch_for_current_device() {
  static void* cached[kMaxDevices] = {nullptr};
  static std::mutex mu;
  const int dev = current_device_slot();  // records its own message (CUDA error, or ordinal out of range)
  if (dev < 0) return nullptr;
  std::lock_guard<std::mutex> lock(mu);
  if (cached[dev] == nullptr) {
    void* ptr = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(&ptr, g_static_attn_scratch);
    if (e != cudaSuccess || ptr == nullptr) {
      fail(DLI_ERR_CUDA, "dli_paged_attn_decode_static_bf16: cudaGetSymbolAddress: %s", cudaGetErrorString(e));
      return nullptr;
    }
    cached[dev] = ptr;
  }
  return cached[dev];
}

}  // namespace dli

extern "C" {

size_t dli_paged_attn_decode_static_max_bytes(void) { return kStaticScratchBytes; }

int dli_paged_attn_decode_static_bf16(const void* q, int q_stride, const void* k_cache,
                                      const void* v_cache, const int32_t* page_table, int max_pages,
                                      const int32_t* seq_lens, void* out, int B, int H, int KV, int hd,
                                      int page_size, int num_splits, float scale, int window,
                                      dli_stream_t stream) {
  using namespace dli;
  const size_t need = dli_paged_attn_decode_scratch_bytes(B, H, hd, num_splits);
  DLI_CHECK_ARG(need <= kStaticScratchBytes,
                "dli_paged_attn_decode_static_bf16: B=%d H=%d hd=%d num_splits=%d needs %zu bytes of split "
                "scratch, the built-in array holds %zu; use fewer splits or dli_paged_attn_decode_bf16 with "
                "your own scratch",
                B, H, hd, num_splits, need, (size_t)kStaticScratchBytes);
  void* scratch = nullptr;
  if (num_splits > 1) {
    scratch = static_scratch_for_current_device();
    if (scratch == nullptr) return DLI_ERR_CUDA;
  }
  // Nothing in the array has to start at zero: every (sequence, head, split) slot this call
  // later reads is first written by this call, empty splits included.
  return dli_paged_attn_decode_bf16(q, q_stride, k_cache, v_cache, page_table, max_pages, seq_lens, out,
                                    scratch, need, B, H, KV, hd, page_size, num_splits, scale, window, stream);
}

}  // extern "C"
