// Causal prefill attention on the 5th-generation tensor cores (head_dim 128).
//
// Work item = 128 query rows of one head; KV tiles of 128 tokens; one persistent CTA per SM walks a
// length-sorted list of items (see TcItem below).  Both GEMMs of flash attention are tcgen05.mma
// with fp32 accumulators in TMEM:
//   S = Q K^T   : A = Q tile, B = K tile, both K-major (hd contiguous), SWIZZLE_128B, via TMA
//   O += P V    : A = P (bf16, two per 32-bit column, written to TENSOR memory by the softmax warps with
//                 tcgen05.st; the TS form of the MMA reads it there), B = V tile as an MN-major operand
//                 (hd contiguous, tokens along K), exactly as TMA lands the [tokens, hd] rows
// TMEM: S0 | S1 (two score buffers, 128 columns each) | O (128 columns) | P0 | P1 (64 columns each).
// Roles (576 threads): warp 0 = TMA producer, warp 1 = MMA issuer + TMEM owner, warps 2-17 =
// softmax + output pass: four threads per query row (TMEM lane), 32 columns each, fp32 online softmax.
// The MMA warp issues S(j+1) before it waits for the softmax of tile j, so the tensor pipe
// works on the next scores while the SFUs exponentiate the current ones; P/V of tile j are
// multiplied as soon as P(j) is in tensor memory and (rarely, see the lazy maximum) O has been rescaled.
// The T x T score matrix never leaves the SM.  FLOPs per head: 4 * T^2/2 * hd (causal).
#include "tcgen05_utils.cuh"

namespace dli {

constexpr int kRowThreads = 4;                 // softmax threads per query row (each owns 128 / kRowThreads columns)
constexpr int kSmWarps = 4 * kRowThreads;      // softmax warps: kRowThreads per TMEM lane quadrant
constexpr int kSmCols = 128 / kRowThreads;     // score / output columns per softmax thread
constexpr int kTcThreads = 64 + 32 * kSmWarps;  // TMA warp + MMA warp + softmax warps
static_assert(kRowThreads == 4, "the P store writes 16 packed columns per thread; the TMEM map assumes 4 x 32 columns");
constexpr int kTcTile = 128;      // query rows per item and tokens per KV tile
constexpr int kTcHd = 128;
constexpr int kSpanBytes = 128 * 64 * 2;  // one [128 rows x 64 elements] swizzled sub-tile: 16 KB
constexpr int kTileBytes2 = 2 * kSpanBytes;  // a [128 x 128] bf16 operand tile: 32 KB

// smem: Q | K0 K1 | V0 V1 | barriers
// + alignment slack, barriers, row max exchange [2][kRowThreads][128], row sum exchange [kRowThreads][128]
constexpr int kTcSmem = 5 * kTileBytes2 + 1024 + 256 + 3 * kRowThreads * 128 * 4;

__device__ __forceinline__ void tmem_st_32x32(uint32_t taddr, const uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
      "{%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,"
      "%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31,%32};" ::"r"(taddr),
      "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]),
      "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]), "r"(r[16]), "r"(r[17]), "r"(r[18]),
      "r"(r[19]), "r"(r[20]), "r"(r[21]), "r"(r[22]), "r"(r[23]), "r"(r[24]), "r"(r[25]), "r"(r[26]), "r"(r[27]),
      "r"(r[28]), "r"(r[29]), "r"(r[30]), "r"(r[31])
      : "memory");
}
__device__ __forceinline__ void tmem_st_32x16(uint32_t taddr, const uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], "
      "{%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};" ::"r"(taddr),
      "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]),
      "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15])
      : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

// MN-major SWIZZLE_128B operand (V as B: N = hd contiguous, K = tokens): 64-element MN groups are
// LBO apart (the two hd halves = two TMA boxes, 16 KB), 8-row K groups are SBO = 1024 B apart.
__device__ __forceinline__ uint64_t make_smem_desc_mn(uint32_t saddr) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr & 0x3ffffu) >> 4);
  d |= (uint64_t)(kSpanBytes >> 4) << 16;  // leading (MN) byte offset
  d |= (uint64_t)64 << 32;                 // stride (K) byte offset = 1024 B
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}
// D=f32, A=bf16 K-major, B=bf16 (K-major or MN-major), M=128, N=128
__host__ __device__ constexpr uint32_t make_idesc_attn(bool b_mn_major) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((b_mn_major ? 1u : 0u) << 16) | ((uint32_t)(128 >> 3) << 17) |
         ((uint32_t)(128 >> 4) << 24);
}

// One work item = one 128-row query tile of one head of one sequence.  The kernel is PERSISTENT: a CTA per SM walks a
// static list of items, so the TMEM allocation, the barrier set-up and the launch of a CTA are paid once per SM and
// the output pass of an item (O from TMEM, normalise, store) runs while the producer loads the next item's Q, K, V and
// the MMA warp computes its first two score tiles.  The list is sorted by length (the last query tile of every head
// first: it has the most key tiles) and dealt to the CTAs in serpentine order — round r gives CTA c item
// r*G + c, the next round item r*G + (G-1-c) — which evens out the sum of lengths per CTA without any counter in
// global memory.  Measured on B200, T = 2048 / 4096, 32 heads: 65.0 / 190.1 us with one CTA per item.
struct TcItem {
  int h, b, kvh, q0, qpos0, t_begin, t_end, row_base;
};

// kPaged (chunked prefill): the T query rows are a chunk at positions prefix .. prefix+T-1; K and V tiles come
// from the paged cache through two more tensor maps (the cache seen as [pages*KV*page_size, hd] rows, box = one
// page x 64 columns): the 32 lanes of the producer warp each issue one page-sized box of a 128-token tile.
// A box whose first token lies beyond the context is requested out of bounds, i.e. zero-filled.
template <bool kPaged>
__global__ void __launch_bounds__(kTcThreads, 1)
attn_prefill_tc_kernel(const __grid_constant__ CUtensorMap tm, const __grid_constant__ CUtensorMap tmK,
                       const __grid_constant__ CUtensorMap tmV, bf16* __restrict__ out, int T, int H, int KV,
                       float scale_log2, int window, const int32_t* __restrict__ page_table, int max_pages,
                       int page_shift, int prefix, int cache_rows, int B, int n_qtiles) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  const uint32_t sQ = smem_addr(smem);
  const uint32_t sK = sQ + kTileBytes2;          // 2 stages
  const uint32_t sV = sK + 2 * kTileBytes2;      // 2 stages
  uint8_t* bar_mem = smem + 5 * kTileBytes2;
  const uint32_t bars = smem_addr(bar_mem);
  // barrier slots (8 B each)
  const uint32_t q_full = bars, k_full = bars + 8, k_empty = bars + 24, v_full = bars + 40, v_empty = bars + 56,
                 s_full = bars + 72, s_empty = bars + 88, p_full = bars + 104, pv_done = bars + 112 /* 2 slots */,
                 q_empty = bars + 128;
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(bar_mem + 136);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n_hb = H * B;
  const int n_items = n_hb * n_qtiles;
  const int G = gridDim.x, cta = blockIdx.x;

  // item r of this CTA (serpentine deal of the length-sorted list); false when the list is exhausted
  auto get_item = [&](int r, TcItem& it) -> bool {
    const int idx = r * G + ((r & 1) ? G - 1 - cta : cta);
    if (idx >= n_items) return false;
    const int level = idx / n_hb, hb = idx - level * n_hb;
    const int qtile = n_qtiles - 1 - level;  // longest first
    it.b = hb / H;
    it.h = hb - it.b * H;
    it.kvh = it.h / (H / KV);
    it.q0 = qtile * kTcTile;
    it.qpos0 = (kPaged ? prefix : 0) + it.q0;  // absolute position of the tile's first query row
    it.t_begin = 0;
    if (window > 0 && it.qpos0 - window + 1 > 0) it.t_begin = (it.qpos0 - window + 1) / kTcTile;
    // last key tile: the one holding the last query row's own position (the diagonal tile when prefix == 0)
    it.t_end = kPaged ? min((prefix + T + kTcTile - 1) / kTcTile, (it.qpos0 + kTcTile - 1) / kTcTile + 1) : qtile + 1;
    it.row_base = it.b * T;  // row of token 0 of this sequence in the packed qkv matrix
    return true;
  };

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tm);
    mbar_init(q_full, 1);
    mbar_init(q_empty, 1);
    for (int i = 0; i < 2; ++i) {
      mbar_init(k_full + 8 * i, 1);
      mbar_init(k_empty + 8 * i, 1);
      mbar_init(v_full + 8 * i, 1);
      mbar_init(v_empty + 8 * i, 1);
      mbar_init(s_full + 8 * i, 1);
      mbar_init(s_empty + 8 * i, kSmWarps);  // one arrive per softmax warp
    }
    mbar_init(p_full, kSmWarps);
    mbar_init(pv_done, 1);
    mbar_init(pv_done + 8, 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(smem_addr(tmem_ptr_smem), 512);
  pdl_wait();
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;
  // S0 | S1 | O (128 columns each) | P0 | P1 (64 columns each: 128 bf16 probabilities per row, two per column)
  const uint32_t tS = tmem_base, tO = tmem_base + 256, tP = tmem_base + 384;

  // Running counters, the same in every role by construction: n = items this CTA has started, g = key tiles it has
  // started over all items.  The K/V/S buffers alternate with g and the barrier parities follow g (Q: n).
  // O needs no second buffer: the first P.V of an item waits for P, which the softmax warps write only after they
  // have read the previous item's O in its output pass.
  if (warp == 0 && kPaged) {
    // ---------------- TMA producer, paged K/V: one page-sized box per lane ----------------
    if (lane == 0) {
      tma_prefetch_desc(&tmK);
      tma_prefetch_desc(&tmV);
    }
    const int page_size = 1 << page_shift;
    const int nb = kTcTile >> page_shift;  // boxes per 64-column half of a tile (<= 16)
    const int half = lane / nb, bi = lane - half * nb;
    const int total = prefix + T;
    int g = 0;
    TcItem w;
    for (int n = 0; get_item(n, w); ++n) {
      if (lane == 0) {
        mbar_wait(q_empty, (n & 1) ^ 1);  // the last S = Q K^T of the previous item has read Q
        mbar_expect_tx(q_full, kTileBytes2);
        tma_load_2d(sQ, &tm, q_full, w.h * kTcHd, w.row_base + w.q0);
        tma_load_2d(sQ + kSpanBytes, &tm, q_full, w.h * kTcHd + 64, w.row_base + w.q0);
      }
      for (int j = w.t_begin; j < w.t_end; ++j, ++g) {
        const int st = g & 1;
        const uint32_t ph = (g >> 1) & 1;
        const int tok0 = j * kTcTile + bi * page_size;
        int row = cache_rows;  // out of bounds: the box is zero-filled
        if (lane < 2 * nb && tok0 < total)
          row = (page_table[(size_t)w.b * max_pages + (tok0 >> page_shift)] * KV + w.kvh) << page_shift;
        const uint32_t dst = (uint32_t)(st * kTileBytes2 + half * kSpanBytes + bi * page_size * 128);
        if (lane == 0) {
          mbar_wait(k_empty + 8 * st, ph ^ 1);
          mbar_expect_tx(k_full + 8 * st, kTileBytes2);
        }
        __syncwarp();
        if (lane < 2 * nb) tma_load_2d(sK + dst, &tmK, k_full + 8 * st, half * 64, row);
        if (lane == 0) {
          mbar_wait(v_empty + 8 * st, ph ^ 1);
          mbar_expect_tx(v_full + 8 * st, kTileBytes2);
        }
        __syncwarp();
        if (lane < 2 * nb) tma_load_2d(sV + dst, &tmV, v_full + 8 * st, half * 64, row);
      }
    }
  } else if (warp == 0) {
    if (lane == 0) {
      // ---------------- TMA producer ----------------
      int g = 0;
      TcItem w;
      for (int n = 0; get_item(n, w); ++n) {
        mbar_wait(q_empty, (n & 1) ^ 1);
        mbar_expect_tx(q_full, kTileBytes2);
        tma_load_2d(sQ, &tm, q_full, w.h * kTcHd, w.row_base + w.q0);
        tma_load_2d(sQ + kSpanBytes, &tm, q_full, w.h * kTcHd + 64, w.row_base + w.q0);
        for (int j = w.t_begin; j < w.t_end; ++j, ++g) {
          const int st = g & 1;
          const uint32_t ph = (g >> 1) & 1;
          const int krow = w.row_base + j * kTcTile;
          mbar_wait(k_empty + 8 * st, ph ^ 1);
          mbar_expect_tx(k_full + 8 * st, kTileBytes2);
          tma_load_2d(sK + st * kTileBytes2, &tm, k_full + 8 * st, (H + w.kvh) * kTcHd, krow);
          tma_load_2d(sK + st * kTileBytes2 + kSpanBytes, &tm, k_full + 8 * st, (H + w.kvh) * kTcHd + 64, krow);
          mbar_wait(v_empty + 8 * st, ph ^ 1);
          mbar_expect_tx(v_full + 8 * st, kTileBytes2);
          tma_load_2d(sV + st * kTileBytes2, &tm, v_full + 8 * st, (H + KV + w.kvh) * kTcHd, krow);
          tma_load_2d(sV + st * kTileBytes2 + kSpanBytes, &tm, v_full + 8 * st, (H + KV + w.kvh) * kTcHd + 64, krow);
        }
      }
    }
  } else if (warp == 1) {
    {
      // ---------------- MMA issuer: the whole warp runs the loop (uniform values stay in uniform registers), one
      // elected lane issues each tcgen05 instruction ----------------
      constexpr uint32_t idesc_s = make_idesc_attn(false), idesc_o = make_idesc_attn(true);
      // operand descriptors: the 14-bit start-address field (bytes >> 4) advances with the k-step, so a k-step is
      // an add on the low word.  Q/K: 16 elements = 32 B inside a 64-element span, spans kSpanBytes apart;
      // V (MN-major): 16 tokens = two 8-row groups = 2048 B
      const uint64_t qdesc = make_smem_desc(sQ);
      const uint64_t kdesc0 = make_smem_desc(sK), vdesc0 = make_smem_desc_mn(sV);
      auto issue_s = [&](int g, bool last_of_item) {  // S(g) = Q K^T into score buffer g & 1
        const int st = g & 1;
        mbar_wait(k_full + 8 * st, (g >> 1) & 1);
        mbar_wait(s_empty + 8 * st, ((g >> 1) & 1) ^ 1);  // the softmax of tile g-2 has read this buffer
        tcgen05_fence_after();
        const uint64_t kdesc = kdesc0 + (uint64_t)(st * (kTileBytes2 >> 4));
        if (elect_one()) {
#pragma unroll
          for (int kk = 0; kk < kTcHd / kUmmaK; ++kk) {
            const uint32_t off = ((kk >> 2) * kSpanBytes + (kk & 3) * 32) >> 4;
            umma_bf16(tS + st * 128, qdesc + off, kdesc + off, idesc_s, kk > 0 ? 1u : 0u);
          }
          umma_commit(s_full + 8 * st);
          umma_commit(k_empty + 8 * st);
          if (last_of_item) umma_commit(q_empty);  // Q may be replaced by the next item's
        }
        __syncwarp();
      };
      int g = 0;
      TcItem w;
      for (int n = 0; get_item(n, w); ++n) {
        const int nt = w.t_end - w.t_begin;
        mbar_wait(q_full, n & 1);
        issue_s(g, nt == 1);
        for (int it = 0; it < nt; ++it, ++g) {
          // S of the next tile goes out before this tile's P is waited for: the tensor pipe works on the next
          // scores while the softmax warps exponentiate the current ones
          if (it + 1 < nt) issue_s(g + 1, it + 2 == nt);
          const int st = g & 1;
          mbar_wait(p_full, g & 1);                 // P(g) in tensor memory, O rescaled (or, it == 0, read out)
          mbar_wait(v_full + 8 * st, (g >> 1) & 1);
          tcgen05_fence_after();
          const uint64_t vdesc = vdesc0 + (uint64_t)(st * (kTileBytes2 >> 4));
          if (elect_one()) {
#pragma unroll
            for (int kk = 0; kk < kTcTile / kUmmaK; ++kk) {
              // A = P from tensor memory: k-step = 16 tokens = 8 columns; B = V, 16 tokens further per k-step
              umma_bf16_ts(tO, tP + st * 64 + kk * 8, vdesc + (uint64_t)(kk * (2048 >> 4)), idesc_o,
                           (it > 0 || kk > 0) ? 1u : 0u);
            }
            umma_commit(pv_done + 8 * st);
            umma_commit(v_empty + 8 * st);
          }
          __syncwarp();
        }
      }
    }
  } else {
    // ---------------- softmax + output pass: kSmWarps warps; a query row is shared by kRowThreads threads ------------
    // The kRowThreads warps w, w+4, ... own the same TMEM lane quadrant (w % 4); `part` selects the thread's kSmCols
    // score/output columns.  Only the row maximum has to be exchanged per tile (shared memory + one named barrier
    // per quadrant); each thread keeps the running sum of its own columns until the output pass.  With four warps
    // per scheduler instead of two the TMEM loads, exchanges and fences of one warp hide behind the arithmetic of
    // the others (the loop was latency-bound at two).
    const int sw = warp - 2;
    const int quad = warp & 3, part = sw >> 2;
    const int row = quad * 32 + lane;           // TMEM lane
    const uint32_t lane_off = (uint32_t)(quad * 32) << 16;
    const uint32_t col0 = part * kSmCols;
    float* xch = reinterpret_cast<float*>(bar_mem + 256);                        // row maxima: [2 parities][parts][128 rows]
    float* xch_l = xch + 2 * kRowThreads * 128;                                  // row sums of the output pass: [parts][128 rows]
    int g = 0;
    TcItem w;
    for (int n = 0; get_item(n, w); ++n) {
      const int qrel = w.q0 + row;                // row of the chunk / prompt
      const int qrow = w.qpos0 + row;             // absolute position in the sequence
      float m_run = -INFINITY, l_run = 0.f;
      for (int j = w.t_begin; j < w.t_end; ++j, ++g) {
        const bool first = j == w.t_begin;
        const int st = g & 1;
        mbar_wait(s_full + 8 * st, (g >> 1) & 1);
        tcgen05_fence_after();
        float s[kSmCols];
        {
          uint32_t r[kSmCols / 32][32];  // all pieces requested before the one wait: one TMEM round trip per tile
#pragma unroll
          for (int c = 0; c < kSmCols / 32; ++c) tmem_ld_32x32(tS + st * 128 + lane_off + col0 + c * 32, r[c]);
          tmem_ld_wait();
#pragma unroll
          for (int c = 0; c < kSmCols / 32; ++c)
#pragma unroll
            for (int i = 0; i < 32; ++i) s[c * 32 + i] = __uint_as_float(r[c][i]);
        }
        tcgen05_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(s_empty + 8 * st);  // score buffer may be overwritten by S(g+2)
        const bool need_mask = (j * kTcTile + kTcTile - 1 > w.qpos0) ||
                               (window > 0 && j * kTcTile < w.qpos0 + kTcTile - window + 1);
        if (need_mask) {
#pragma unroll
          for (int i = 0; i < kSmCols; ++i) {
            const int tok = j * kTcTile + (int)col0 + i;
            if (tok > qrow || (window > 0 && tok <= qrow - window)) s[i] = -INFINITY;
          }
        }
        float mx4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};  // four chains: one long chain is latency-bound
#pragma unroll
        for (int i = 0; i < kSmCols; i += 4) {
          mx4[0] = fmaxf(mx4[0], s[i]);
          mx4[1] = fmaxf(mx4[1], s[i + 1]);
          mx4[2] = fmaxf(mx4[2], s[i + 2]);
          mx4[3] = fmaxf(mx4[3], s[i + 3]);
        }
        float mx = fmaxf(fmaxf(mx4[0], mx4[1]), fmaxf(mx4[2], mx4[3]));
        float* xrow = xch + (g & 1) * (kRowThreads * 128) + row;
        xrow[part * 128] = mx;
        // only the warps that share a TMEM lane quadrant meet (named barrier 1 + quad): the four quadrants do not
        // wait for each other
        asm volatile("bar.sync %0, %1;" ::"r"(1 + quad), "n"(32 * kRowThreads) : "memory");
#pragma unroll
        for (int q = 0; q < kRowThreads; ++q) mx = fmaxf(mx, xrow[q * 128]);
        // Lazy reference maximum: the exponentials only need a COMMON reference per row, not the true running
        // maximum, so m_run moves only when the tile's maximum exceeds it by more than 2^kLazyLog2 (p stays
        // <= 256, exact in bf16's range; sums are fp32).  Then corr == 1 on almost every tile and the
        // TMEM round trip that rescales O below is skipped.  (-inf - -inf = NaN compares false: a fully
        // masked tile leaves m_run alone.)
        constexpr float kLazyLog2 = 8.f;
        const float m_new = ((mx - m_run) * scale_log2 > kLazyLog2) ? mx : m_run;
        const float m_use = (m_new == -INFINITY) ? 0.f : m_new;
        const float corr = ex2_ftz((m_run - m_use) * scale_log2);
        m_run = m_new;
        // p = 2^(s*scale - m*scale): one FFMA feeds each exp2; four partial sums
        const float ms = m_use * scale_log2;
        float rs4[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
        for (int i = 0; i < kSmCols; i += 4) {
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            s[i + u] = ex2_ftz(fmaf(s[i + u], scale_log2, -ms));
            rs4[u] += s[i + u];
          }
        }
        const float rs = (rs4[0] + rs4[1]) + (rs4[2] + rs4[3]);
        l_run = l_run * corr + rs;
        // P (bf16, two tokens per 32-bit column) -> tensor memory, buffer g & 1: the A operand of this tile's P.V.
        // The buffer's previous reader, P.V(g-2), has finished: S(g) was issued after it and is complete.
        {
          static_assert(kSmCols == 32, "P store below writes 16 packed columns per thread");
          uint32_t pk[16];
#pragma unroll
          for (int c = 0; c < 16; ++c) pk[c] = pack_bf16(s[2 * c], s[2 * c + 1]);
          tmem_st_32x16(tP + st * 64 + lane_off + (col0 >> 1), pk);
        }
        // (skipped when no row of this warp moved its maximum: corr == 1 exactly, nothing to do)
        if (!first && !__all_sync(0xffffffffu, corr == 1.f)) {
          // O *= corr on this thread's output columns; TMEM -> registers -> TMEM.  The previous P.V of this item
          // must have finished first.  P.V(t) completes barrier pv_done[t & 1]; the softmax warps do not wait for
          // every P.V any more, so the parity test must be safe without that: S(g) complete implies P.V(g-3)
          // complete (issued earlier, same pipe), and P.V(g+1) cannot start before this warp delivers P(g+1) —
          // barrier (g-1) & 1 is in the phase of P.V(g-1) or just past it, never further.
          mbar_wait(pv_done + 8 * ((g - 1) & 1), ((g - 1) >> 1) & 1);
          tcgen05_fence_after();
#pragma unroll
          for (int c = 0; c < kSmCols / 32; ++c) {
            uint32_t r[32];
            tmem_ld_32x32(tO + lane_off + col0 + c * 32, r);
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 32; ++i) r[i] = __float_as_uint(__uint_as_float(r[i]) * corr);
            tmem_st_32x32(tO + lane_off + col0 + c * 32, r);
          }
        }
        tmem_st_wait();            // P (and O) are in tensor memory before the MMA warp is told
        tcgen05_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(p_full);
      }
      // ---- output pass: O / l -> bf16 (the threads of a row add their partial sums, all in the same order) ----
      xch_l[part * 128 + row] = l_run;
      asm volatile("bar.sync %0, %1;" ::"r"(1 + quad), "n"(32 * kRowThreads) : "memory");
      l_run = xch_l[row];
#pragma unroll
      for (int q = 1; q < kRowThreads; ++q) l_run += xch_l[q * 128 + row];
      mbar_wait(pv_done + 8 * ((g - 1) & 1), ((g - 1) >> 1) & 1);  // the item's last P.V (same argument as above)
      tcgen05_fence_after();
      const float inv = l_run > 0.f ? 1.f / l_run : 0.f;
      bf16* dst = out + ((size_t)(w.row_base + qrel)) * ((size_t)H * kTcHd) + (size_t)w.h * kTcHd + col0;
      {
        uint32_t r[kSmCols / 32][32];
#pragma unroll
        for (int c = 0; c < kSmCols / 32; ++c) tmem_ld_32x32(tO + lane_off + col0 + c * 32, r[c]);
        tmem_ld_wait();
        if (qrel < T) {
#pragma unroll
          for (int c = 0; c < kSmCols / 32; ++c)
#pragma unroll
            for (int gq = 0; gq < 4; ++gq) {
              float v[8];
#pragma unroll
              for (int i = 0; i < 8; ++i) v[i] = __uint_as_float(r[c][gq * 8 + i]) * inv;
              st_16B(dst + c * 32 + gq * 8, pack8(v));
            }
        }
      }
      tcgen05_fence_before();  // O has been read: the next item's first P.V (after its P) may overwrite it
    }
  }
  __syncthreads();
  if (warp == 1) {
    tcgen05_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

}  // namespace dli

extern "C" int dli_attn_prefill_tc_bf16(const void* qkv, void* out, int B, int T, int H, int KV, int hd, float scale,
                                        int window, dli_stream_t stream) {
  using namespace dli;
  DLI_CHECK_ARG(qkv && out, "dli_attn_prefill_tc_bf16: null pointer");
  DLI_CHECK_ARG(hd == 128, "dli_attn_prefill_tc_bf16: head_dim must be 128 (got %d); use dli_attn_prefill_bf16", hd);
  DLI_CHECK_ARG(B >= 0 && T >= 0 && H > 0 && KV > 0 && H % KV == 0 && window >= 0,
                "dli_attn_prefill_tc_bf16: bad sizes B=%d T=%d H=%d KV=%d", B, T, H, KV);
  DLI_CHECK_ARG((long long)((T + 127) / 128) * H * B <= 0x7fffffffLL, "dli_attn_prefill_tc_bf16: too many query tiles");
  DLI_CHECK_ARG(aligned16(qkv) && aligned16(out), "dli_attn_prefill_tc_bf16: pointers must be 16-byte aligned");
  if (B == 0 || T == 0) return DLI_OK;
  static PerDeviceOnce once;
  DLI_ONCE_PER_DEVICE(once, dev, DLI_CUDA_CALL(cudaFuncSetAttribute(
                                     attn_prefill_tc_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kTcSmem)));
  CUtensorMap tm;
  const int cols = (H + 2 * KV) * hd;
  // the packed qkv matrix [B*T, cols]; box = 128 rows x 64 columns (one swizzle span)
  int rc = make_tmap(&tm, qkv, B * T, cols, 128);
  if (rc != DLI_OK) return rc;
  const int n_qtiles = (T + kTcTile - 1) / kTcTile;
  const long long n_items = (long long)n_qtiles * H * B;
  dim3 grid((unsigned)(n_items < sm_count_cached() ? n_items : sm_count_cached()));  // persistent: one CTA per SM
  DLI_LAUNCH("attn_prefill_tc_kernel", attn_prefill_tc_kernel<false>, grid, dim3(kTcThreads), (size_t)kTcSmem,
             static_cast<cudaStream_t>(stream), tm, tm, tm, (bf16*)out, T, H, KV, scale * 1.4426950408889634f, window,
             (const int32_t*)nullptr, 0, 0, 0, 0, B, n_qtiles);
  return DLI_OK;
}

namespace dli {

// tcgen05 form of dli_attn_prefill_paged_bf16 (head_dim 128, 8 <= page_size <= 128); called from attn_prefill.cu.
int launch_attn_prefill_tc_paged(const void* qkv, void* out, const void* k_cache, const void* v_cache,
                                 const int32_t* page_table, int max_pages, int num_pages, int B, int T, int prefix,
                                 int H, int KV, int page_size, float scale, int window, cudaStream_t s) {
  static PerDeviceOnce once;
  DLI_ONCE_PER_DEVICE(once, dev, DLI_CUDA_CALL(cudaFuncSetAttribute(
                                     attn_prefill_tc_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kTcSmem)));
  CUtensorMap tm, tmK, tmV;
  const int cols = (H + 2 * KV) * kTcHd;
  int rc = make_tmap(&tm, qkv, B * T, cols, 128);
  if (rc != DLI_OK) return rc;
  // the cache pools as [num_pages * KV * page_size, 128] rows; box = one page x 64 columns
  const long long rows = (long long)num_pages * KV * page_size;
  if (rows <= 0 || rows > 0x7fffffffLL) return fail(DLI_ERR_INVALID, "dli_attn_prefill_paged_bf16: cache too large for one tensor map");
  rc = make_tmap(&tmK, k_cache, (int)rows, kTcHd, page_size);
  if (rc != DLI_OK) return rc;
  rc = make_tmap(&tmV, v_cache, (int)rows, kTcHd, page_size);
  if (rc != DLI_OK) return rc;
  int shift = 0;
  while ((1 << shift) < page_size) ++shift;
  const int n_qtiles = (T + kTcTile - 1) / kTcTile;
  const long long n_items = (long long)n_qtiles * H * B;
  dim3 grid((unsigned)(n_items < sm_count_cached() ? n_items : sm_count_cached()));
  DLI_LAUNCH("attn_prefill_tc_kernel", attn_prefill_tc_kernel<true>, grid, dim3(kTcThreads), (size_t)kTcSmem, s, tm,
             tmK, tmV, (bf16*)out, T, H, KV, scale * 1.4426950408889634f, window, page_table, max_pages, shift, prefix,
             (int)rows, B, n_qtiles);
  return DLI_OK;
}

}  // namespace dli
