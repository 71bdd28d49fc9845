// bf16 GEMM on the 5th-generation tensor cores: tcgen05.mma with the fp32
// accumulator in TMEM, operands staged in shared memory by TMA, hand-written
// in inline PTX (no CUTLASS/CuTe).
//
//   y[M,N] = epi( x[M,K] . w[N,K]^T )          both operands K-major (row-major,
//                                              K contiguous) = HF Linear layout
//
// One 128-row UMMA tile per CTA, BLOCK_K = 64 (one 128-byte swizzle span of
// bf16), a kStages-deep TMA->MMA ring of mbarriers, and three warp roles:
//   warp 0  : TMA producer (one elected lane issues cp.async.bulk.tensor)
//   warp 1  : TMEM allocator + MMA issuer (one elected lane issues tcgen05.mma,
//             tcgen05.commit releases smem slots / publishes the accumulator)
//   warps 2-5: epilogue — tcgen05.ld 32 lanes x 32 columns per warp, fused
//             bias / gelu / residual / silu*up, bf16 pack, 16-byte stores.
//
// Two operand assignments (template kSwap):
//   kSwap = false (prefill, M large):  A = x tile [128 m, 64 k], B = w tile
//       [BN n, 64 k]; TMEM lane = output row m, column = n.
//   kSwap = true  (batched decode, M <= 128): A = w tile [128 n, 64 k], B = x
//       tile [BN >= M rows m, 64 k] (TMA zero-fills rows >= M); TMEM lane = n,
//       column = m, so no MMA work is wasted on padding and the kernel is a pure
//       weight stream; the epilogue writes y[m, n] with the warp's 32 lanes on
//       32 consecutive n.
// Gated MLP (template kGated): w is the [gate128 | up128]-interleaved matrix;
//   non-swap uses one BN=256 tile (TMEM columns 0-127 gate, 128-255 up), swap
//   issues two M=128 MMAs per k-step into two accumulators.  The epilogue
//   computes bf16(silu(bf16(g))) * bf16(u) — the HF rounding points.
//
// Tensor-bound for prefill: 2*M*N*K FLOPs per launch; HBM-bound weight stream
// for decode: N*K*2 bytes per launch.
#include <stdlib.h>

#include "tcgen05_utils.cuh"

namespace dli {

constexpr int kGemmThreads = 192;

template <int BN, bool kSwap, bool kGated>
struct GemmCfg {
  // A tile rows: 128, or 256 for the swap-gated case (gate128 | up128 in one box)
  static constexpr int kARows = (kSwap && kGated) ? 256 : 128;
  static constexpr int kABytes = kARows * kBlockK * 2;
  static constexpr int kBBytes = BN * kBlockK * 2;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kStagesRaw = (220 * 1024) / kStageBytes;
  static constexpr int kStages = kStagesRaw > 8 ? 8 : kStagesRaw;
  // accumulator columns: BN, or 2*BN for the two swap-gated accumulators
  static constexpr int kAccCols = (kSwap && kGated) ? 2 * BN : BN;
  static constexpr int kTmemCols = kAccCols <= 32 ? 32 : kAccCols <= 64 ? 64 : kAccCols <= 128 ? 128 : kAccCols <= 256 ? 256 : 512;
  static constexpr int kSmemBytes = kStages * kStageBytes + 1024 /*align slack*/ + 256 /*barriers*/;
  static_assert(kStageBytes % 1024 == 0, "stages must keep 1024-byte alignment");
  static_assert(BN % 16 == 0 && BN >= 16 && BN <= 256, "UMMA N");
  static_assert(!kGated || kSwap || BN == 256, "non-swap gated tile is gate128|up128");
};

template <int BN, bool kSwap, bool kGated>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_tcgen05_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                    const bf16* __restrict__ bias, const bf16* residual, bf16* y, int M, int N, int K,
                    int epilogue, float* splitk_partial, int* splitk_counters) {
  using Cfg = GemmCfg<BN, kSwap, kGated>;
  extern __shared__ uint8_t smem_raw[];
  // SWIZZLE_128B operands need 1024-byte alignment
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* bar_mem = smem + Cfg::kStages * Cfg::kStageBytes;
  const uint32_t full_bar = smem_addr(bar_mem);               // kStages x 8 B
  const uint32_t empty_bar = full_bar + 8 * Cfg::kStages;     // kStages x 8 B
  const uint32_t tmem_full_bar = empty_bar + 8 * Cfg::kStages;
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(bar_mem + 8 * (2 * Cfg::kStages + 1));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // split-K (gridDim.z > 1, swap mode): this CTA reduces k-blocks [kb0, kb0 + num_kb)
  const int total_kb = K / kBlockK;
  const int kb_per = (total_kb + gridDim.z - 1) / gridDim.z;
  const int kb0 = blockIdx.z * kb_per;
  const int num_kb = min(kb_per, total_kb - kb0);
  // blockIdx.x walks the activation-row tiles fastest so CTAs sharing a weight
  // tile run together and the weight tile is fetched from HBM once.
  const int a_row0 = (kSwap ? blockIdx.y : blockIdx.x) * Cfg::kARows;  // rows of the A-operand matrix
  const int b_row0 = (kSwap ? blockIdx.x : blockIdx.y) * BN;           // rows of the B-operand matrix

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    for (int s = 0; s < Cfg::kStages; ++s) {
      mbar_init(full_bar + 8 * s, 1);
      mbar_init(empty_bar + 8 * s, 1);
    }
    mbar_init(tmem_full_bar, 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(smem_addr(tmem_ptr_smem), Cfg::kTmemCols);
  // barrier init / TMEM allocation overlap the previous kernel's tail; the warps that touch
  // data written by the previous kernel call pdl_wait() themselves (producer: after the
  // first ring of WEIGHT loads is in flight; epilogue: before reading residual / scratch)
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;

  if (warp == 0) {
    if (lane == 0) {
      // the operand that holds WEIGHTS (A in swap mode, B otherwise) does not depend on the
      // previous kernel: its first ring of loads goes out before the dependency wait
      const int pre = min((int)Cfg::kStages, num_kb);
      for (int kb = 0; kb < pre; ++kb) {
        const uint32_t sA = smem_addr(smem + kb * Cfg::kStageBytes);
        mbar_expect_tx(full_bar + 8 * kb, Cfg::kStageBytes);
        if (kSwap) tma_load_2d(sA, &tmA, full_bar + 8 * kb, (kb0 + kb) * kBlockK, a_row0);
        else tma_load_2d(sA + Cfg::kABytes, &tmB, full_bar + 8 * kb, (kb0 + kb) * kBlockK, b_row0);
      }
      pdl_wait();
      for (int kb = 0; kb < num_kb; ++kb) {
        const int s = kb % Cfg::kStages;
        const uint32_t ph = (kb / Cfg::kStages) & 1;
        const uint32_t sA = smem_addr(smem + s * Cfg::kStageBytes);
        const uint32_t sB = sA + Cfg::kABytes;
        if (kb < pre) {  // weights in flight, expect_tx armed: only the activation tile is missing
          if (kSwap) tma_load_2d(sB, &tmB, full_bar + 8 * s, (kb0 + kb) * kBlockK, b_row0);
          else tma_load_2d(sA, &tmA, full_bar + 8 * s, (kb0 + kb) * kBlockK, a_row0);
          continue;
        }
        mbar_wait(empty_bar + 8 * s, ph ^ 1);
        mbar_expect_tx(full_bar + 8 * s, Cfg::kStageBytes);
        tma_load_2d(sA, &tmA, full_bar + 8 * s, (kb0 + kb) * kBlockK, a_row0);
        tma_load_2d(sB, &tmB, full_bar + 8 * s, (kb0 + kb) * kBlockK, b_row0);
      }
    }
  } else if (warp == 1) {
    // MMA issuer: the whole warp walks the loop with warp-uniform values (descriptors stay in uniform registers),
    // one elected lane issues
    constexpr uint32_t idesc = make_idesc(BN);
    const uint64_t adesc0 = make_smem_desc(smem_addr(smem));
    const uint64_t bdesc0 = make_smem_desc(smem_addr(smem) + Cfg::kABytes);
    for (int kb = 0; kb < num_kb; ++kb) {
      const int s = kb % Cfg::kStages;
      const uint32_t ph = (kb / Cfg::kStages) & 1;
      mbar_wait(full_bar + 8 * s, ph);
      tcgen05_fence_after();
      const uint64_t adesc = adesc0 + (uint64_t)(s * (Cfg::kStageBytes >> 4));  // start-address field: bytes >> 4
      const uint64_t bdesc = bdesc0 + (uint64_t)(s * (Cfg::kStageBytes >> 4));
      if (elect_one()) {
#pragma unroll
        for (int k = 0; k < kBlockK / kUmmaK; ++k) {
          // advancing 16 elements (32 bytes) along K inside the swizzle span: +2 in the >>4 address field
          const uint32_t acc = (kb > 0 || k > 0) ? 1u : 0u;
          umma_bf16(tmem_base, adesc + 2 * k, bdesc + 2 * k, idesc, acc);
          if (kSwap && kGated) {
            // second accumulator: the up rows (A rows 128..255 start 128*128 B further)
            umma_bf16(tmem_base + BN, adesc + (128 * 128 >> 4) + 2 * k, bdesc + 2 * k, idesc, acc);
          }
        }
        umma_commit(empty_bar + 8 * s);  // frees the smem slot when these MMAs retire
        if (kb == num_kb - 1) umma_commit(tmem_full_bar);  // accumulator complete
      }
      __syncwarp();
    }
    if (num_kb <= 0 && elect_one()) umma_commit(tmem_full_bar);
  } else {
    // ---- epilogue warps 2..5: TMEM lane quadrant = warp % 4 -----------------------------
    const int quad = warp & 3;
    pdl_wait();
    mbar_wait(tmem_full_bar, 0);
    tcgen05_fence_after();
    const uint32_t lane_base = tmem_base + ((uint32_t)(quad * 32) << 16);
    if (!kSwap) {
      const int m = blockIdx.x * kBlockM + quad * 32 + lane;  // this thread's output row
      constexpr int kOutCols = kGated ? 128 : BN;             // output columns of this tile
      const int n_tile0 = blockIdx.y * kOutCols;
#pragma unroll 1
      for (int c0 = 0; c0 < kOutCols; c0 += 16) {
        uint32_t r[16], r2[16];
        tmem_ld_32x16(lane_base + c0, r);
        if (kGated) tmem_ld_32x16(lane_base + 128 + c0, r2);
        tmem_ld_wait();
        const int n0 = n_tile0 + c0;
        if (m < M && n0 < N) {
          float v[16];
#pragma unroll
          for (int j = 0; j < 16; ++j) v[j] = __uint_as_float(r[j]);
          if (kGated) {
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              const float gte = round_bf16(v[j]), up = round_bf16(__uint_as_float(r2[j]));
              v[j] = round_bf16(silu_f(gte)) * up;
            }
          } else {
            if (epilogue & DLI_EPI_BIAS) {
#pragma unroll
              for (int j = 0; j < 16; ++j) v[j] += (n0 + j < N) ? __bfloat162float(bias[n0 + j]) : 0.f;
            }
#pragma unroll
            for (int j = 0; j < 16; ++j) v[j] = round_bf16(v[j]);
            if (epilogue & DLI_EPI_GELU_NEW) {
#pragma unroll
              for (int j = 0; j < 16; ++j) v[j] = round_bf16(gelu_new_f(v[j]));
            }
          }
          const int n_out = kGated ? N / 2 : N;  // gated: N counts gate+up rows of w
          bf16* dst = y + (size_t)m * n_out + n0;
#pragma unroll
          for (int h = 0; h < 2; ++h) {
            if (n0 + h * 8 < n_out) {  // N % 8 == 0 is checked on the host
              if (!kGated && (epilogue & DLI_EPI_RESIDUAL)) {
                float rr[8];
                unpack8(ld_16B(residual + (size_t)m * n_out + n0 + h * 8), rr);
#pragma unroll
                for (int j = 0; j < 8; ++j) v[h * 8 + j] += rr[j];
              }
              st_16B(dst + h * 8, pack8(v + h * 8));
            }
          }
        }
      }
    } else {
      // swap: lane = weight row n, columns = activation rows m (only m < M are real)
      constexpr int kOutRows = 128;
      const int row = quad * 32 + lane;
      const int n = blockIdx.y * kOutRows + row;
      const int n_out = kGated ? N / 2 : N;
      const int m_tile0 = blockIdx.x * BN;
      const int splits = gridDim.z;
      const int tile = blockIdx.y * gridDim.x + blockIdx.x;
      bool finish = true;
      if (splits > 1) {
        // Deterministic split-K: every split parks its fp32 partial tile in the workspace
        // ([tile][split][column m][row n], lanes contiguous), the CTA that arrives last at the
        // tile's counter adds the partials in split order and runs the epilogue, then
        // re-zeroes the counter for the next launch.
        float* mine = splitk_partial + ((size_t)tile * splits + blockIdx.z) * Cfg::kAccCols * 128;
#pragma unroll 1
        for (int c0 = 0; c0 < BN; c0 += 16) {
          if (m_tile0 + c0 >= M) break;
          uint32_t r[16], r2[16];
          tmem_ld_32x16(lane_base + c0, r);
          if (kGated) tmem_ld_32x16(lane_base + BN + c0, r2);
          tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            if (m_tile0 + c0 + j < M) {
              mine[(size_t)(c0 + j) * 128 + row] = __uint_as_float(r[j]);
              if (kGated) mine[(size_t)(BN + c0 + j) * 128 + row] = __uint_as_float(r2[j]);
            }
          }
        }
        __threadfence();
        asm volatile("bar.sync 1, 128;" ::: "memory");  // the 4 epilogue warps
        volatile uint32_t* flag = tmem_ptr_smem + 2;
        if (warp == 2 && lane == 0) *flag = (atomicAdd(splitk_counters + tile, 1) == splits - 1) ? 1u : 0u;
        asm volatile("bar.sync 1, 128;" ::: "memory");
        finish = *flag != 0;
        if (finish) __threadfence();
      }
      if (finish && splits > 1) {
        // Reduction by the last-arriving CTA of the tile.  The partials are in global memory now, so
        // the work is re-dealt for memory parallelism: an item is 8 consecutive weight rows n of one
        // activation row m (two float4 per split, one 16-byte residual load, one 16-byte store), 16
        // consecutive threads cover the 128 rows of a y row, and a thread requests everything its
        // batch of items needs before it touches any of it (one L2 round trip per batch instead of
        // one per 16-column chunk).  Splits are still added in split order: deterministic.
        const float* all = splitk_partial + (size_t)tile * splits * Cfg::kAccCols * 128;
        constexpr int kItems = kGated ? 2 : 4;  // items per batch (register budget)
        const int et = (int)threadIdx.x - 64;   // 0..127 over the epilogue warps
        const int m_real = min(BN, M - m_tile0);
        const int items = m_real * 16;
        const int n_tile0 = blockIdx.y * kOutRows;
        for (int it0 = 0; it0 < items; it0 += 128 * kItems) {
          float acc1[kItems][8], acc2[kItems][8];
          uint4 resv[kItems];
#pragma unroll
          for (int q = 0; q < kItems; ++q) {
#pragma unroll
            for (int e = 0; e < 8; ++e) { acc1[q][e] = 0.f; acc2[q][e] = 0.f; }
            const int id = it0 + q * 128 + et;
            const int mc = id >> 4, ng = id & 15;
            const int nn = n_tile0 + ng * 8;
            const bool on = id < items && nn < n_out;
            resv[q] = make_uint4(0u, 0u, 0u, 0u);
            if (!kGated && on && (epilogue & DLI_EPI_RESIDUAL))
              resv[q] = ld_16B(residual + (size_t)(m_tile0 + mc) * n_out + nn);
          }
          for (int sp0 = 0; sp0 < splits; sp0 += 4) {
            float4 v[kItems][4][kGated ? 4 : 2];
#pragma unroll
            for (int q = 0; q < kItems; ++q) {
              const int id = it0 + q * 128 + et;
              const int mc = id >> 4, ng = id & 15;
              const bool on = id < items;
#pragma unroll
              for (int u = 0; u < 4; ++u) {
                const bool ld = on && sp0 + u < splits;
                const float* src = all + ((size_t)(ld ? sp0 + u : 0) * Cfg::kAccCols + (ld ? mc : 0)) * 128 + ng * 8;
                const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
                v[q][u][0] = ld ? __ldcg(reinterpret_cast<const float4*>(src)) : z;
                v[q][u][1] = ld ? __ldcg(reinterpret_cast<const float4*>(src) + 1) : z;
                if (kGated) {
                  v[q][u][2] = ld ? __ldcg(reinterpret_cast<const float4*>(src + (size_t)BN * 128)) : z;
                  v[q][u][3] = ld ? __ldcg(reinterpret_cast<const float4*>(src + (size_t)BN * 128) + 1) : z;
                }
              }
            }
#pragma unroll
            for (int q = 0; q < kItems; ++q)
#pragma unroll
              for (int u = 0; u < 4; ++u) {
                acc1[q][0] += v[q][u][0].x; acc1[q][1] += v[q][u][0].y; acc1[q][2] += v[q][u][0].z; acc1[q][3] += v[q][u][0].w;
                acc1[q][4] += v[q][u][1].x; acc1[q][5] += v[q][u][1].y; acc1[q][6] += v[q][u][1].z; acc1[q][7] += v[q][u][1].w;
                if (kGated) {
                  acc2[q][0] += v[q][u][2].x; acc2[q][1] += v[q][u][2].y; acc2[q][2] += v[q][u][2].z; acc2[q][3] += v[q][u][2].w;
                  acc2[q][4] += v[q][u][3].x; acc2[q][5] += v[q][u][3].y; acc2[q][6] += v[q][u][3].z; acc2[q][7] += v[q][u][3].w;
                }
              }
          }
#pragma unroll
          for (int q = 0; q < kItems; ++q) {
            const int id = it0 + q * 128 + et;
            const int mc = id >> 4, ng = id & 15;
            const int nn = n_tile0 + ng * 8;
            if (id < items && nn < n_out) {
              float o[8], rr[8];
              unpack8(resv[q], rr);
              float bvv[8];
#pragma unroll
              for (int e = 0; e < 8; ++e) bvv[e] = 0.f;
              if (!kGated && (epilogue & DLI_EPI_BIAS)) unpack8(ld_16B(bias + nn), bvv);
#pragma unroll
              for (int e = 0; e < 8; ++e) {
                if (kGated) {
                  const float gte = round_bf16(acc1[q][e]), up = round_bf16(acc2[q][e]);
                  o[e] = round_bf16(silu_f(gte)) * up;
                } else {
                  float vv = round_bf16(acc1[q][e] + bvv[e]);
                  if (epilogue & DLI_EPI_GELU_NEW) vv = round_bf16(gelu_new_f(vv));
                  o[e] = vv + rr[e];
                }
              }
              st_16B(y + (size_t)(m_tile0 + mc) * n_out + nn, pack8(o));
            }
          }
        }
      } else if (finish) {
#pragma unroll 1
        for (int c0 = 0; c0 < BN; c0 += 16) {
          if (m_tile0 + c0 >= M) break;  // warp-uniform
          float a1[16], a2[16];
          {
            uint32_t r[16], r2[16];
            tmem_ld_32x16(lane_base + c0, r);
            if (kGated) tmem_ld_32x16(lane_base + BN + c0, r2);
            tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < 16; ++j) { a1[j] = __uint_as_float(r[j]); a2[j] = kGated ? __uint_as_float(r2[j]) : 0.f; }
          }
          if (n < n_out) {
            const float bv = (!kGated && (epilogue & DLI_EPI_BIAS)) ? __bfloat162float(bias[n]) : 0.f;
            // y may alias residual: read all residual values of the chunk before the first store,
            // otherwise every load would have to wait for the previous store
            float rr[16];
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              const int m = m_tile0 + c0 + j;
              rr[j] = (!kGated && (epilogue & DLI_EPI_RESIDUAL) && m < M)
                          ? __bfloat162float(residual[(size_t)m * n_out + n]) : 0.f;
            }
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              const int m = m_tile0 + c0 + j;
              if (m < M) {
                float v;
                if (kGated) {
                  const float gte = round_bf16(a1[j]), up = round_bf16(a2[j]);
                  v = round_bf16(silu_f(gte)) * up;
                } else {
                  v = round_bf16(a1[j] + bv);
                  if (epilogue & DLI_EPI_GELU_NEW) v = round_bf16(gelu_new_f(v));
                  v += rr[j];
                }
                y[(size_t)m * n_out + n] = __float2bfloat16_rn(v);
              }
            }
          }
        }
      }
      if (finish) {
        if (splits > 1 && warp == 2 && lane == 0) splitk_counters[tile] = 0;
      }
    }
    tcgen05_fence_before();
  }
  __syncthreads();
  if (warp == 1) {
    tcgen05_fence_after();
    tmem_dealloc(tmem_base, Cfg::kTmemCols);
  }
}


// ======================================================================================
// Persistent prefill GEMM (M > 128): one CTA per SM walks the output tiles, the accumulator
// is double-buffered in TMEM (2 x BN columns) so the epilogue of tile i overlaps the MMAs
// of tile i+1, and the TMA -> MMA ring keeps running across tile boundaries.
//   A = x tile [128 m, 64 k], B = w tile [BN n, 64 k]; TMEM lane = m, column = n.
//   warp 0: TMA producer   warp 1: MMA issuer + TMEM owner   warps 2-5: epilogue
// Weight tiles do not depend on the previous kernel: the producer issues the first ring
// of weight loads before it waits for the previous kernel (programmatic dependent launch).
// ======================================================================================

// MT = 128-row halves of the activation tile (1 or 2).  MT = 2: the CTA multiplies a 256-row activation tile by the
// weight tile with two MMAs per k-step (rows 0-127 and 128-255, one accumulator each) that read the SAME weight
// tile from shared memory.  A 128 x 256 tile asks L2 for 48 KB per 64-wide k-block, i.e. per four MMAs — more than
// one SM's share of the L2 bandwidth delivers in the time the tensor core needs for them, so the 128-row kernel
// is bound by the operand feed (measured 0.85-0.95 of cuBLAS); 256 x 256 asks for 64 KB per EIGHT MMAs.  Its two
// accumulators fill the 512 TMEM columns, so the epilogue is not overlapped with the next tile.
template <int BN, bool kGated, int MT = 1>
struct PrefillCfg {
  static constexpr int kABytes = MT * 128 * kBlockK * 2;
  static constexpr int kBBytes = BN * kBlockK * 2;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kStagesRaw = (216 * 1024) / kStageBytes;
  static constexpr int kStages = kStagesRaw > 8 ? 8 : kStagesRaw;
  static constexpr int kAccCols = MT * BN;                          // TMEM columns of one tile's accumulators
  static constexpr int kAccStages = 2 * kAccCols <= 512 ? 2 : 1;    // double-buffered when it fits
  static constexpr int kTmemWant = kAccStages * kAccCols;
  static constexpr int kTmemCols = kTmemWant <= 32 ? 32 : kTmemWant <= 64 ? 64 : kTmemWant <= 128 ? 128 : kTmemWant <= 256 ? 256 : 512;
  static constexpr int kOutCols = kGated ? 128 : BN;  // output columns per tile
  static constexpr int kSmemBytes = kStages * kStageBytes + 1024 + 256;
  static_assert(!kGated || BN == 256, "gated tile is gate128|up128");
  static_assert(MT == 1 || MT == 2, "one or two 128-row halves");
  static_assert(kAccCols <= 512 && kStages >= 2, "tile does not fit TMEM / shared memory");
};

// Tile walk of the persistent kernel: m-tiles in groups of 16, inside a group m fastest, then n.  The
// 148 tiles in flight at any time then cover ~16 activation tiles x ~9 weight tiles, so both operands are
// re-used out of L2 (m fastest over ALL m-tiles would re-stream the whole activation matrix from HBM
// once per weight tile when M is tens of thousands of rows).  M <= 2048 is a single group.
__device__ __forceinline__ void prefill_tile_coords(int tile, int m_tiles, int num_tiles, int& m_blk, int& n_blk) {
  constexpr int kGroup = 16;
  const int n_tiles = num_tiles / m_tiles;
  const int per_group = kGroup * n_tiles;
  const int grp = tile / per_group;
  const int m0 = grp * kGroup;
  const int gsize = min(kGroup, m_tiles - m0);
  const int r = tile - grp * per_group;
  n_blk = r / gsize;
  m_blk = m0 + r - n_blk * gsize;
}

// kRope (the qkv projection of a prompt pass, head_dim 128, 128 x 256 tiles = two whole heads per tile): the epilogue
// rotates q and k (rotate-half RoPE with the rounding points of apply_rotary_pos_emb on bf16 tensors,
// TF:models/llama/modeling_llama.py:138-168, applied to the bf16-rounded projection exactly as the separate
// rope_kv_append_kernel applies it to the stored one) and appends k and v to the paged cache — the work of
// dli_rope_kv_append_bf16 without a second pass over the qkv matrix.
struct RopeArgs {
  const bf16* cos_t;           // [max_pos, 64]
  const bf16* sin_t;
  const int32_t* positions;    // [M]
  const int32_t* seq_ids;      // [M]
  const int32_t* page_table;   // [n_seqs, max_pages]
  bf16* k_cache;               // [num_pages, KV, page_size, 128]
  bf16* v_cache;
  int max_pages, H, KV, page_shift;
};

// threads: TMA warp + MMA warp + four epilogue warps per 128-row half
template <int BN, bool kGated, int MT, bool kRope = false>
__global__ void __launch_bounds__(64 + 128 * MT, 1)
gemm_prefill_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                    const bf16* __restrict__ bias, const bf16* residual, bf16* y, int M, int N, int K,
                    int epilogue, int m_tiles, int num_tiles, const __grid_constant__ RopeArgs rope) {
  static_assert(!kRope || (BN == 256 && !kGated && MT == 1), "the RoPE epilogue works on two whole 128-wide heads");
  using Cfg = PrefillCfg<BN, kGated, MT>;
  constexpr int S = Cfg::kStages;
  constexpr int kMRows = MT * 128;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* bar_mem = smem + S * Cfg::kStageBytes;
  const uint32_t full_bar = smem_addr(bar_mem);
  const uint32_t empty_bar = full_bar + 8 * S;
  const uint32_t tfull_bar = empty_bar + 8 * S;   // 2 accumulator stages
  const uint32_t tempty_bar = tfull_bar + 16;
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(bar_mem + 8 * (2 * S + 4));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int num_kb = K / kBlockK;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    for (int s = 0; s < S; ++s) {
      mbar_init(full_bar + 8 * s, 1);
      mbar_init(empty_bar + 8 * s, 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(tfull_bar + 8 * a, 1);
      mbar_init(tempty_bar + 8 * a, 4 * MT);  // one arrive per epilogue warp
    }
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(smem_addr(tmem_ptr_smem), Cfg::kTmemCols);
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;

  if (warp == 0) {
    if (lane == 0) {
      // ring of weight (B) loads for the first tile goes out before the dependency wait
      const int first = blockIdx.x;
      const int pre = (first < num_tiles) ? min(S, num_kb) : 0;
      int first_m = 0, first_n = 0;
      if (pre) prefill_tile_coords(first, m_tiles, num_tiles, first_m, first_n);
      for (int kb = 0; kb < pre; ++kb) {
        const uint32_t sA = smem_addr(smem + kb * Cfg::kStageBytes);
        mbar_expect_tx(full_bar + 8 * kb, Cfg::kStageBytes);
        tma_load_2d(sA + Cfg::kABytes, &tmB, full_bar + 8 * kb, kb * kBlockK, first_n * BN);
      }
      pdl_wait();
      uint32_t kc = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        int m_blk, n_blk;
        prefill_tile_coords(tile, m_tiles, num_tiles, m_blk, n_blk);
        for (int kb = 0; kb < num_kb; ++kb, ++kc) {
          const int s = kc % S;
          const uint32_t sA = smem_addr(smem + s * Cfg::kStageBytes);
          if ((int)kc < pre) {  // B already in flight, expect_tx already armed
            tma_load_2d(sA, &tmA, full_bar + 8 * s, kb * kBlockK, m_blk * kMRows);
            continue;
          }
          mbar_wait(empty_bar + 8 * s, ((kc / S) & 1) ^ 1);
          mbar_expect_tx(full_bar + 8 * s, Cfg::kStageBytes);
          tma_load_2d(sA, &tmA, full_bar + 8 * s, kb * kBlockK, m_blk * kMRows);
          tma_load_2d(sA + Cfg::kABytes, &tmB, full_bar + 8 * s, kb * kBlockK, n_blk * BN);
        }
      }
    }
  } else if (warp == 1) {
    // MMA issuer: the whole warp walks the loop with warp-uniform values (descriptors stay in uniform registers:
    // no per-instruction R2UR chain as under a `lane == 0` branch), one elected lane issues
    constexpr uint32_t idesc = make_idesc(BN);
    const uint64_t adesc0 = make_smem_desc(smem_addr(smem));
    const uint64_t bdesc0 = make_smem_desc(smem_addr(smem) + Cfg::kABytes);
    uint32_t kc = 0, tl = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++tl) {
      const uint32_t as = Cfg::kAccStages == 2 ? (tl & 1) : 0;
      const uint32_t aph = Cfg::kAccStages == 2 ? ((tl >> 1) & 1) : (tl & 1);
      mbar_wait(tempty_bar + 8 * as, aph ^ 1);  // epilogue has drained this accumulator
      tcgen05_fence_after();
      const uint32_t tacc = tmem_base + as * Cfg::kAccCols;
      for (int kb = 0; kb < num_kb; ++kb, ++kc) {
        const int s = kc % S;
        mbar_wait(full_bar + 8 * s, (kc / S) & 1);
        tcgen05_fence_after();
        const uint64_t adesc = adesc0 + (uint64_t)(s * (Cfg::kStageBytes >> 4));  // start-address field: bytes >> 4
        const uint64_t bdesc = bdesc0 + (uint64_t)(s * (Cfg::kStageBytes >> 4));
        if (elect_one()) {
#pragma unroll
          for (int k = 0; k < kBlockK / kUmmaK; ++k) {
#pragma unroll
            for (int h = 0; h < MT; ++h)  // rows 128 h .. 128 h + 127 of the activation tile: 128 rows x 128 B further
              umma_bf16(tacc + h * BN, adesc + (uint64_t)(h * (128 * 128 >> 4)) + 2 * k, bdesc + 2 * k, idesc,
                        (kb > 0 || k > 0) ? 1u : 0u);
          }
          umma_commit(empty_bar + 8 * s);
          if (kb == num_kb - 1) umma_commit(tfull_bar + 8 * as);
        }
        __syncwarp();
      }
    }
  } else {
    const int quad = warp & 3;
    pdl_wait();  // residual / bias may come from the previous kernel
    uint32_t tl = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++tl) {
      int m_blk, n_blk;
      prefill_tile_coords(tile, m_tiles, num_tiles, m_blk, n_blk);
      const uint32_t as = Cfg::kAccStages == 2 ? (tl & 1) : 0;
      mbar_wait(tfull_bar + 8 * as, Cfg::kAccStages == 2 ? ((tl >> 1) & 1) : (tl & 1));
      tcgen05_fence_after();
      const int n_tile0 = n_blk * Cfg::kOutCols;
      const int n_out = kGated ? N / 2 : N;
      if constexpr (kRope) {
        const uint32_t lane_base = tmem_base + as * Cfg::kAccCols + ((uint32_t)(quad * 32) << 16);
        const int m = m_blk * 128 + quad * 32 + lane;
        const bool live = m < M;
        int pos = 0;
        size_t cache_row = 0;  // ((page * KV) * page_size + slot): + kvh * page_size gives the token's row of a kv head
        if (live) {
          pos = rope.positions[m];
          const int page = rope.page_table[(size_t)rope.seq_ids[m] * rope.max_pages + (pos >> rope.page_shift)];
          cache_row = ((size_t)page * rope.KV << rope.page_shift) + (pos & ((1 << rope.page_shift) - 1));
        }
#pragma unroll 1
        for (int hh = 0; hh < 2; ++hh) {
          const int head = (n_tile0 >> 7) + hh;  // q heads, then k heads, then v heads
          if (head >= rope.H + 2 * rope.KV) break;
          bf16* yrow = y + (size_t)m * N + (size_t)head * 128;
          if (head < rope.H + rope.KV) {
            bf16* crow = head >= rope.H ? rope.k_cache + (cache_row + ((size_t)(head - rope.H) << rope.page_shift)) * 128 : nullptr;
#pragma unroll 1
            for (int j = 0; j < 4; ++j) {  // 16 rotation pairs (i, i + 64) per step
              uint32_t r1[16], r2[16];
              tmem_ld_32x16(lane_base + hh * 128 + j * 16, r1);
              tmem_ld_32x16(lane_base + hh * 128 + 64 + j * 16, r2);
              tmem_ld_wait();
              if (live) {
                float c[16], sn[16], o1[16], o2[16];
                unpack8(ld_16B(rope.cos_t + (size_t)pos * 64 + j * 16), c);
                unpack8(ld_16B(rope.cos_t + (size_t)pos * 64 + j * 16 + 8), c + 8);
                unpack8(ld_16B(rope.sin_t + (size_t)pos * 64 + j * 16), sn);
                unpack8(ld_16B(rope.sin_t + (size_t)pos * 64 + j * 16 + 8), sn + 8);
#pragma unroll
                for (int e = 0; e < 16; ++e) {
                  const float x1 = round_bf16(__uint_as_float(r1[e])), x2 = round_bf16(__uint_as_float(r2[e]));
                  o1[e] = round_bf16(x1 * c[e]) + round_bf16(-x2 * sn[e]);
                  o2[e] = round_bf16(x2 * c[e]) + round_bf16(x1 * sn[e]);
                }
#pragma unroll
                for (int v8 = 0; v8 < 2; ++v8) {
                  const uint4 u1 = pack8(o1 + v8 * 8), u2 = pack8(o2 + v8 * 8);
                  st_16B(yrow + j * 16 + v8 * 8, u1);
                  st_16B(yrow + 64 + j * 16 + v8 * 8, u2);
                  if (crow) {
                    st_16B(crow + j * 16 + v8 * 8, u1);
                    st_16B(crow + 64 + j * 16 + v8 * 8, u2);
                  }
                }
              }
            }
          } else {
            bf16* crow = rope.v_cache + (cache_row + ((size_t)(head - rope.H - rope.KV) << rope.page_shift)) * 128;
#pragma unroll 1
            for (int j = 0; j < 8; ++j) {
              uint32_t r1[16];
              tmem_ld_32x16(lane_base + hh * 128 + j * 16, r1);
              tmem_ld_wait();
              if (live) {
                float v[16];
#pragma unroll
                for (int e = 0; e < 16; ++e) v[e] = __uint_as_float(r1[e]);
#pragma unroll
                for (int v8 = 0; v8 < 2; ++v8) {
                  const uint4 u = pack8(v + v8 * 8);
                  st_16B(yrow + j * 16 + v8 * 8, u);
                  st_16B(crow + j * 16 + v8 * 8, u);
                }
              }
            }
          }
        }
      } else
#pragma unroll 1
      for (int c0 = 0; c0 < Cfg::kOutCols; c0 += 16) {
        const int h = (warp - 2) >> 2;  // warps 2-5: rows 0-127 of the tile, warps 6-9 (MT = 2): rows 128-255
        const uint32_t lane_base = tmem_base + as * Cfg::kAccCols + h * BN + ((uint32_t)(quad * 32) << 16);
        const int m = m_blk * kMRows + h * 128 + quad * 32 + lane;
        uint32_t r[16], r2[16];
        tmem_ld_32x16(lane_base + c0, r);
        if (kGated) tmem_ld_32x16(lane_base + 128 + c0, r2);
        tmem_ld_wait();
        const int n0 = n_tile0 + c0;
        if (m < M && n0 < n_out) {
          float v[16];
#pragma unroll
          for (int j = 0; j < 16; ++j) v[j] = __uint_as_float(r[j]);
          if (kGated) {
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              const float gte = round_bf16(v[j]), up = round_bf16(__uint_as_float(r2[j]));
              v[j] = round_bf16(silu_f(gte)) * up;
            }
          } else {
            if (epilogue & DLI_EPI_BIAS) {
#pragma unroll
              for (int j = 0; j < 16; ++j) v[j] += (n0 + j < N) ? __bfloat162float(bias[n0 + j]) : 0.f;
            }
#pragma unroll
            for (int j = 0; j < 16; ++j) v[j] = round_bf16(v[j]);
            if (epilogue & DLI_EPI_GELU_NEW) {
#pragma unroll
              for (int j = 0; j < 16; ++j) v[j] = round_bf16(gelu_new_f(v[j]));
            }
          }
          bf16* dst = y + (size_t)m * n_out + n0;
#pragma unroll
          for (int h = 0; h < 2; ++h) {
            if (n0 + h * 8 < n_out) {
              if (!kGated && (epilogue & DLI_EPI_RESIDUAL)) {
                float rr[8];
                unpack8(ld_16B(residual + (size_t)m * n_out + n0 + h * 8), rr);
#pragma unroll
                for (int j = 0; j < 8; ++j) v[h * 8 + j] += rr[j];
              }
              st_16B(dst + h * 8, pack8(v + h * 8));
            }
          }
        }
      }
      // this warp is done reading the accumulator stage: hand it back to the MMA warp
      tcgen05_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(tempty_bar + 8 * as);
    }
  }
  tcgen05_fence_before();
  __syncthreads();
  if (warp == 1) {
    tcgen05_fence_after();
    tmem_dealloc(tmem_base, Cfg::kTmemCols);
  }
}

// ---- host side ---------------------------------------------------------------------------
template <int BN, bool kGated, int MT = 1, bool kRope = false>
static int launch_prefill_gemm(const void* x, const void* w, const void* bias, const void* residual, void* y, int M,
                               int N /* rows of w */, int K, int epilogue, cudaStream_t s,
                               const RopeArgs& rope = RopeArgs{}) {
  using Cfg = PrefillCfg<BN, kGated, MT>;
  static PerDeviceOnce once;
  DLI_ONCE_PER_DEVICE(once, dev, DLI_CUDA_CALL(cudaFuncSetAttribute(gemm_prefill_kernel<BN, kGated, MT, kRope>,
                                                                    cudaFuncAttributeMaxDynamicSharedMemorySize,
                                                                    Cfg::kSmemBytes)));
  CUtensorMap tmA, tmB;
  int rc;
  if ((rc = make_tmap(&tmA, x, M, K, MT * 128)) != DLI_OK) return rc;
  if ((rc = make_tmap(&tmB, w, N, K, BN)) != DLI_OK) return rc;
  const int m_tiles = (M + MT * 128 - 1) / (MT * 128), n_tiles = (N + BN - 1) / BN;
  const int num_tiles = m_tiles * n_tiles;
  const int sms = sm_count_cached();
  const int grid = num_tiles < sms ? num_tiles : sms;
  DLI_LAUNCH("gemm_prefill_kernel", (gemm_prefill_kernel<BN, kGated, MT, kRope>), dim3(grid), dim3(64 + 128 * MT),
             (size_t)Cfg::kSmemBytes, s, tmA, tmB, (const bf16*)bias, (const bf16*)residual, (bf16*)y, M, N, K, epilogue,
             m_tiles, num_tiles, rope);
  return DLI_OK;
}

template <int BN, bool kSwap, bool kGated>
static int launch_gemm(const void* x, const void* w, const void* bias, const void* residual, void* y, int M,
                       int N /* rows of w */, int K, int epilogue, cudaStream_t s, void* scratch = nullptr,
                       size_t scratch_bytes = 0) {
  if constexpr (!kSwap) return launch_prefill_gemm<BN, kGated>(x, w, bias, residual, y, M, N, K, epilogue, s);
  using Cfg = GemmCfg<BN, kSwap, kGated>;
  static PerDeviceOnce once;
  DLI_ONCE_PER_DEVICE(once, dev, DLI_CUDA_CALL(cudaFuncSetAttribute(gemm_tcgen05_kernel<BN, kSwap, kGated>,
                                                                    cudaFuncAttributeMaxDynamicSharedMemorySize,
                                                                    Cfg::kSmemBytes)));
  CUtensorMap tmA, tmB;
  int rc;
  dim3 grid;
  if (!kSwap) {
    if ((rc = make_tmap(&tmA, x, M, K, 128)) != DLI_OK) return rc;
    if ((rc = make_tmap(&tmB, w, N, K, BN)) != DLI_OK) return rc;
    grid = dim3((M + 127) / 128, (N + BN - 1) / BN);
  } else {
    if ((rc = make_tmap(&tmA, w, N, K, Cfg::kARows)) != DLI_OK) return rc;
    if ((rc = make_tmap(&tmB, x, M, K, BN)) != DLI_OK) return rc;
    grid = dim3((M + BN - 1) / BN, (N + Cfg::kARows - 1) / Cfg::kARows);
  }
  float* partial = nullptr;
  int* counters = nullptr;
  if (kSwap && scratch) {
    // split K so that about one CTA per SM streams weights (each needs ~200 KB of smem)
    const int tiles = grid.x * grid.y, total_kb = K / kBlockK;
    int splits = sm_count_cached() / tiles;
    if (splits > total_kb) splits = total_kb;
    if (splits > 1) {
      const int kb_per = (total_kb + splits - 1) / splits;
      splits = (total_kb + kb_per - 1) / kb_per;  // no empty split
    }
    if (splits > 1) {
      const size_t cbytes = (((size_t)tiles * sizeof(int)) + 255) & ~(size_t)255;
      const size_t need = cbytes + (size_t)tiles * splits * Cfg::kAccCols * 128 * sizeof(float);
      if (need <= scratch_bytes) {
        counters = reinterpret_cast<int*>(scratch);
        partial = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(scratch) + cbytes);
        grid.z = splits;
      }
    }
  }
  DLI_LAUNCH("gemm_tcgen05_kernel", gemm_tcgen05_kernel<BN, kSwap, kGated>, grid, dim3(kGemmThreads),
             (size_t)Cfg::kSmemBytes, s, tmA, tmB, (const bf16*)bias, (const bf16*)residual, (bf16*)y, M, N, K, epilogue,
             partial, counters);
  return DLI_OK;
}

}  // namespace dli

extern "C" {

int dli_gemm_bf16(const void* x, const void* w, const void* bias, const void* residual, void* y, int M, int N,
                  int K, int epilogue, dli_stream_t stream) {
  using namespace dli;
  DLI_CHECK_ARG(x && w && y, "dli_gemm_bf16: null pointer");
  DLI_CHECK_ARG(M >= 0 && N > 0 && K > 0 && K % 64 == 0 && N % 16 == 0,
                "dli_gemm_bf16: need K %% 64 == 0 and N %% 16 == 0 (got M=%d N=%d K=%d)", M, N, K);
  DLI_CHECK_ARG(!(epilogue & DLI_EPI_BIAS) || bias, "dli_gemm_bf16: DLI_EPI_BIAS without bias");
  DLI_CHECK_ARG(!(epilogue & DLI_EPI_RESIDUAL) || residual, "dli_gemm_bf16: DLI_EPI_RESIDUAL without residual");
  DLI_CHECK_ARG(aligned16(x) && aligned16(w) && aligned16(y) && aligned16(residual),
                "dli_gemm_bf16: pointers must be 16-byte aligned");
  if (M == 0) return DLI_OK;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (M <= 16) return launch_gemm<16, true, false>(x, w, bias, residual, y, M, N, K, epilogue, s);
  if (M <= 32) return launch_gemm<32, true, false>(x, w, bias, residual, y, M, N, K, epilogue, s);
  if (M <= 64) return launch_gemm<64, true, false>(x, w, bias, residual, y, M, N, K, epilogue, s);
  if (M <= 128) return launch_gemm<128, true, false>(x, w, bias, residual, y, M, N, K, epilogue, s);
  if (N % 256 == 0 || N >= 2048) {
    // Tile choice for the persistent kernel (one CTA per SM walks rounds of tiles).  Two effects, both measured on
    // B200 (scratch/exp_gemm_vs_cublas.py, profiles/r2_notes.md): (1) wave quantisation — a 2k-token prefill has
    // few tiles, so the tile WIDTH is chosen to fill whole rounds (UMMA N may be any multiple of 16: 240-column
    // tiles turn 256 / 384 tiles = 1.73 / 2.59 waves into 288 / 416 = 1.95 / 2.81); (2) operand feed — a 256-row
    // tile (two MMAs per weight sub-tile) needs less L2 bandwidth per MMA than a 128-row tile, but its two
    // accumulators fill TMEM, so its epilogue (about 3 us) is not overlapped: it wins where it makes ONE round of
    // what were two (o / down projection at 2 048 rows: 56.6 -> 52.0 us, 179.1 -> 163.0 us; cuBLAS 51 / 157).
    // cost (us) = rounds x (k-blocks x rows/128 x width x c + epilogue), c fitted per tile shape.
    const int sms = sm_count_cached() > 0 ? sm_count_cached() : 148;
    const double kbs = K / 64;
    struct Shape { int mt, bn; double c, epi; };
    static const Shape shapes[] = {{1, 256, 1.60e-3, 0.5}, {1, 240, 1.72e-3, 0.5}, {2, 256, 1.48e-3, 3.0}, {2, 240, 1.55e-3, 3.0}};
    static const int force = getenv("DLI_GEMM_TILE") ? atoi(getenv("DLI_GEMM_TILE")) : 0;  // measurement knob: mt * 1000 + bn
    int best_mt = 1, best_bn = 256;
    double best = 1e30;
    for (const Shape& c : shapes) {
      if (c.mt == 2 && M <= 128) continue;
      const long long tiles = (long long)((M + c.mt * 128 - 1) / (c.mt * 128)) * ((N + c.bn - 1) / c.bn);
      const double v = (double)((tiles + sms - 1) / sms) * (kbs * c.mt * c.bn * c.c + c.epi);
      if (v < best) best = v, best_mt = c.mt, best_bn = c.bn;
    }
    if (force) best_mt = force / 1000, best_bn = force % 1000;
    if (best_mt == 2) {
      if (best_bn == 240) return launch_prefill_gemm<240, false, 2>(x, w, bias, residual, y, M, N, K, epilogue, s);
      return launch_prefill_gemm<256, false, 2>(x, w, bias, residual, y, M, N, K, epilogue, s);
    }
    if (best_bn == 240) return launch_gemm<240, false, false>(x, w, bias, residual, y, M, N, K, epilogue, s);
    return launch_gemm<256, false, false>(x, w, bias, residual, y, M, N, K, epilogue, s);
  }
  if (N % 128 == 0 || N > 128) return launch_gemm<128, false, false>(x, w, bias, residual, y, M, N, K, epilogue, s);
  if (N > 32) return launch_gemm<64, false, false>(x, w, bias, residual, y, M, N, K, epilogue, s);
  return launch_gemm<32, false, false>(x, w, bias, residual, y, M, N, K, epilogue, s);
}

int dli_gemm_qkv_rope_append_bf16(const void* x, const void* w_qkv, void* qkv, int M, int K, const void* cos,
                                  const void* sin, const int32_t* positions, const int32_t* seq_ids,
                                  const int32_t* page_table, int max_pages, void* k_cache, void* v_cache, int H, int KV,
                                  int hd, int page_size, dli_stream_t stream) {
  using namespace dli;
  DLI_CHECK_ARG(x && w_qkv && qkv && cos && sin && positions && seq_ids && page_table && k_cache && v_cache,
                "dli_gemm_qkv_rope_append_bf16: null pointer");
  DLI_CHECK_ARG(hd == 128 && H > 0 && KV > 0 && H % 2 == 0 && KV % 2 == 0,
                "dli_gemm_qkv_rope_append_bf16: needs head_dim 128 and even head counts (got hd=%d H=%d KV=%d); use "
                "dli_gemm_bf16 + dli_rope_kv_append_bf16", hd, H, KV);
  DLI_CHECK_ARG(M > 128 && K > 0 && K % 64 == 0,
                "dli_gemm_qkv_rope_append_bf16: a prompt pass of more than 128 rows, K %% 64 == 0 (got M=%d K=%d)", M, K);
  DLI_CHECK_ARG(page_size >= 8 && (page_size & (page_size - 1)) == 0 && max_pages > 0,
                "dli_gemm_qkv_rope_append_bf16: page_size must be a power of two >= 8 (got %d)", page_size);
  DLI_CHECK_ARG(aligned16(x) && aligned16(w_qkv) && aligned16(qkv) && aligned16(cos) && aligned16(sin) &&
                    aligned16(k_cache) && aligned16(v_cache),
                "dli_gemm_qkv_rope_append_bf16: pointers must be 16-byte aligned");
  int shift = 0;
  while ((1 << shift) < page_size) ++shift;
  RopeArgs ra;
  ra.cos_t = (const bf16*)cos;
  ra.sin_t = (const bf16*)sin;
  ra.positions = positions;
  ra.seq_ids = seq_ids;
  ra.page_table = page_table;
  ra.k_cache = (bf16*)k_cache;
  ra.v_cache = (bf16*)v_cache;
  ra.max_pages = max_pages;
  ra.H = H;
  ra.KV = KV;
  ra.page_shift = shift;
  return launch_prefill_gemm<256, false, 1, true>(x, w_qkv, nullptr, nullptr, qkv, M, (H + 2 * KV) * hd, K, 0,
                                                  static_cast<cudaStream_t>(stream), ra);
}

int dli_gemm_gated_silu_bf16(const void* x, const void* w_gu, void* y, int M, int F, int K, dli_stream_t stream) {
  using namespace dli;
  DLI_CHECK_ARG(x && w_gu && y, "dli_gemm_gated_silu_bf16: null pointer");
  DLI_CHECK_ARG(M >= 0 && F > 0 && F % 128 == 0 && K > 0 && K % 64 == 0,
                "dli_gemm_gated_silu_bf16: need F %% 128 == 0, K %% 64 == 0 (got M=%d F=%d K=%d)", M, F, K);
  DLI_CHECK_ARG(aligned16(x) && aligned16(w_gu) && aligned16(y),
                "dli_gemm_gated_silu_bf16: pointers must be 16-byte aligned");
  if (M == 0) return DLI_OK;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const int N2 = 2 * F;  // rows of w_gu
  if (M <= 16) return launch_gemm<16, true, true>(x, w_gu, nullptr, nullptr, y, M, N2, K, 0, s);
  if (M <= 32) return launch_gemm<32, true, true>(x, w_gu, nullptr, nullptr, y, M, N2, K, 0, s);
  if (M <= 64) return launch_gemm<64, true, true>(x, w_gu, nullptr, nullptr, y, M, N2, K, 0, s);
  if (M <= 128) return launch_gemm<128, true, true>(x, w_gu, nullptr, nullptr, y, M, N2, K, 0, s);
  // (256-row tiles measured slower here: 420 against 332 us at 2 048 rows — the gated epilogue is long and would
  // not be overlapped, and 896 tiles make 7 rounds where 1 792 make 13 half as long)
  return launch_gemm<256, false, true>(x, w_gu, nullptr, nullptr, y, M, N2, K, 0, s);
}

size_t dli_gemm_splitk_scratch_bytes(int M, int N) {
  // one int32 counter per 128-row weight tile + at most one fp32 partial tile per SM
  // (<= 256 CTAs) of 2 accumulators x BN columns x 128 rows, BN = M rounded up to 16/32/64/128
  if (M <= 0 || N <= 0 || M > 128) return 0;
  const size_t bn = M <= 16 ? 16 : M <= 32 ? 32 : M <= 64 ? 64 : 128;
  const size_t tiles = ((size_t)N + 127) / 128;
  return ((tiles * sizeof(int) + 255) & ~(size_t)255) + (size_t)256 * 2 * bn * 128 * sizeof(float);
}

int dli_gemm_splitk_bf16(const void* x, const void* w, const void* bias, const void* residual, void* y, int M, int N,
                         int K, int epilogue, void* scratch, size_t scratch_bytes, dli_stream_t stream) {
  using namespace dli;
  if (M > 128 || scratch == nullptr) return dli_gemm_bf16(x, w, bias, residual, y, M, N, K, epilogue, stream);
  DLI_CHECK_ARG(x && w && y, "dli_gemm_splitk_bf16: null pointer");
  DLI_CHECK_ARG(M >= 0 && N > 0 && K > 0 && K % 64 == 0 && N % 16 == 0,
                "dli_gemm_splitk_bf16: need K %% 64 == 0 and N %% 16 == 0 (got M=%d N=%d K=%d)", M, N, K);
  DLI_CHECK_ARG(!(epilogue & DLI_EPI_BIAS) || bias, "dli_gemm_splitk_bf16: DLI_EPI_BIAS without bias");
  DLI_CHECK_ARG(!(epilogue & DLI_EPI_RESIDUAL) || residual, "dli_gemm_splitk_bf16: DLI_EPI_RESIDUAL without residual");
  DLI_CHECK_ARG(aligned16(x) && aligned16(w) && aligned16(y) && aligned16(residual) && aligned16(scratch),
                "dli_gemm_splitk_bf16: pointers must be 16-byte aligned");
  if (M == 0) return DLI_OK;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (M <= 16) return launch_gemm<16, true, false>(x, w, bias, residual, y, M, N, K, epilogue, s, scratch, scratch_bytes);
  if (M <= 32) return launch_gemm<32, true, false>(x, w, bias, residual, y, M, N, K, epilogue, s, scratch, scratch_bytes);
  if (M <= 64) return launch_gemm<64, true, false>(x, w, bias, residual, y, M, N, K, epilogue, s, scratch, scratch_bytes);
  return launch_gemm<128, true, false>(x, w, bias, residual, y, M, N, K, epilogue, s, scratch, scratch_bytes);
}

int dli_gemm_splitk_gated_silu_bf16(const void* x, const void* w_gu, void* y, int M, int F, int K, void* scratch,
                                    size_t scratch_bytes, dli_stream_t stream) {
  using namespace dli;
  if (M > 128 || scratch == nullptr) return dli_gemm_gated_silu_bf16(x, w_gu, y, M, F, K, stream);
  DLI_CHECK_ARG(x && w_gu && y, "dli_gemm_splitk_gated_silu_bf16: null pointer");
  DLI_CHECK_ARG(M >= 0 && F > 0 && F % 128 == 0 && K > 0 && K % 64 == 0,
                "dli_gemm_splitk_gated_silu_bf16: need F %% 128 == 0, K %% 64 == 0 (got M=%d F=%d K=%d)", M, F, K);
  DLI_CHECK_ARG(aligned16(x) && aligned16(w_gu) && aligned16(y) && aligned16(scratch),
                "dli_gemm_splitk_gated_silu_bf16: pointers must be 16-byte aligned");
  if (M == 0) return DLI_OK;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const int N2 = 2 * F;
  if (M <= 16) return launch_gemm<16, true, true>(x, w_gu, nullptr, nullptr, y, M, N2, K, 0, s, scratch, scratch_bytes);
  if (M <= 32) return launch_gemm<32, true, true>(x, w_gu, nullptr, nullptr, y, M, N2, K, 0, s, scratch, scratch_bytes);
  if (M <= 64) return launch_gemm<64, true, true>(x, w_gu, nullptr, nullptr, y, M, N2, K, 0, s, scratch, scratch_bytes);
  return launch_gemm<128, true, true>(x, w_gu, nullptr, nullptr, y, M, N2, K, 0, s, scratch, scratch_bytes);
}

int dli_linear_bf16(const void* x, const void* w, const void* bias, const void* residual, void* y, int M, int N,
                    int K, int epilogue, dli_stream_t stream) {
  // measured on B200 (profiles/r1_notes.md): the CUDA-core GEMV streams at 6.0-7.3 TB/s for M <= 2 but
  // drops to 4.1-5.0 TB/s at M = 3-4 (FMA issue bound); the swap-AB tensor-core stream holds 5.7-6.4
  if (M <= 2 || K % 64 != 0 || N % 16 != 0)
    return dli_gemv_bf16(x, w, bias, residual, y, M, N, K, epilogue, stream);
  // no split-K scratch here: a matrix of few 128-row tiles would leave most SMs idle in the swap-AB
  // kernel (1.8 TB/s on 4096 x 14336), the mma.sync stream holds 4-5 TB/s on those
  if (M <= 8 && N < 8192) return dli_gemv_mma_bf16(x, w, bias, residual, y, M, N, K, epilogue, stream);
  return dli_gemm_bf16(x, w, bias, residual, y, M, N, K, epilogue, stream);
}

int dli_linear_gated_silu_bf16(const void* x, const void* w_gu, void* y, int M, int F, int K, dli_stream_t stream) {
  if (M <= 2 || K % 64 != 0) return dli_gemv_gated_silu_bf16(x, w_gu, y, M, F, K, stream);
  return dli_gemm_gated_silu_bf16(x, w_gu, y, M, F, K, stream);
}

}  // extern "C"
