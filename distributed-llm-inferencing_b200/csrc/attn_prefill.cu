// Causal prefill attention over the packed, already-rotated qkv buffer.
//
// Grid = (ceil(T/64) query tiles, H, B); 4 warps x 16 query rows.  K/V tiles of
// 64 tokens are staged with 16-byte cp.async (double-buffered, XOR-swizzled) and
// consumed by warp-level m16n8k16 MMAs with an fp32 online softmax (flash-style:
// the T x T score matrix is never written).  Heavier (later) query tiles are
// scheduled first.  Tensor-bound in principle (4*T^2/2*hd FLOPs per head) but
// only ~4 % of prefill FLOPs at T=2048; K/V re-reads by the
// G query heads of a KV head are served by L2.
//
// kPaged (chunked prefill): the T rows are a CHUNK of each prompt at positions prefix .. prefix+T-1.
// Queries still come from the packed qkv buffer of the chunk, keys and values of the WHOLE context
// (the earlier chunks and this one, appended by dli_rope_kv_append_bf16 just before) come from the
// paged KV cache, page by page like the decode kernel does it.  Same arithmetic otherwise.
#include <stdlib.h>

#include "mma_utils.cuh"

namespace dli {

constexpr int kPfThreads = 128;
constexpr int kPfTile = 64;

template <int HD, bool kPaged>
__global__ void __launch_bounds__(kPfThreads)
attn_prefill_kernel(const bf16* __restrict__ qkv, bf16* __restrict__ out, int T, int H, int KV,
                    float scale_log2, int window, const bf16* __restrict__ k_cache,
                    const bf16* __restrict__ v_cache, const int32_t* __restrict__ page_table, int max_pages,
                    int page_shift, int prefix) {
  constexpr int RB = HD * 2;
  constexpr int CH = HD / 8;
  constexpr int kTileBytes = kPfTile * RB;
  extern __shared__ __align__(128) uint8_t smem[];
  const uint32_t sQ = smem_u32(smem);
  const uint32_t sK = sQ + kTileBytes;
  const uint32_t sV = sK + 2 * kTileBytes;

  pdl_wait();
  // z (dispatched slowest) walks the query tiles from the longest to the shortest, x and y the heads and sequences:
  // the longest CTAs of all heads start first (see attn_prefill_tc.cu)
  const int qtile = gridDim.z - 1 - blockIdx.z;
  const int h = blockIdx.x, b = blockIdx.y;
  const int kvh = h / (H / KV);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const size_t ld = (size_t)(H + 2 * KV) * HD;
  const bf16* base = qkv + (size_t)b * T * ld;
  const int q0 = qtile * kPfTile;

  for (int id = tid; id < kPfTile * CH; id += kPfThreads) {
    const int row = id / CH, c = id - row * CH;
    const bool valid = q0 + row < T;
    const bf16* src = base + (size_t)(valid ? q0 + row : 0) * ld + (size_t)h * HD + c * 8;
    cp_async_16B(sQ + swz<RB>(row, c), src, valid ? 16 : 0);
  }
  auto load_tile = [&](int tile, int stage) {
    const uint32_t dK = sK + stage * kTileBytes, dV = sV + stage * kTileBytes;
#pragma unroll
    for (int i = 0; i < kPfTile * CH / kPfThreads; ++i) {
      const int id = tid + i * kPfThreads;
      const int row = id / CH, c = id - row * CH;
      const int tok = tile * kPfTile + row;  // kPaged: absolute position in the sequence
      if (kPaged) {
        const bool valid = tok < prefix + T;
        size_t off = 0;
        if (valid) {
          const int page = page_table[(size_t)b * max_pages + (tok >> page_shift)];
          off = ((((size_t)page * KV + kvh) << page_shift) + (tok & ((1 << page_shift) - 1))) * HD + c * 8;
        }
        cp_async_16B(dK + swz<RB>(row, c), k_cache + off, valid ? 16 : 0);
        cp_async_16B(dV + swz<RB>(row, c), v_cache + off, valid ? 16 : 0);
      } else {
        const bool valid = tok < T;
        const bf16* src = base + (size_t)(valid ? tok : 0) * ld + (size_t)(H + kvh) * HD + c * 8;
        cp_async_16B(dK + swz<RB>(row, c), src, valid ? 16 : 0);
        cp_async_16B(dV + swz<RB>(row, c), src + (size_t)KV * HD, valid ? 16 : 0);
      }
    }
  };

  // absolute position of the tile's first query row (prefix == 0 when not paged)
  const int qpos0 = (kPaged ? prefix : 0) + q0;
  int t_begin = 0;
  if (window > 0 && qpos0 - window + 1 > 0) t_begin = (qpos0 - window + 1) / kPfTile;
  // last key tile: the one holding the last query row's own position (the diagonal tile when prefix == 0)
  const int t_end = kPaged ? min((prefix + T + kPfTile - 1) / kPfTile, (qpos0 + kPfTile - 1) / kPfTile + 1) : qtile + 1;

  float m_run[2] = {-INFINITY, -INFINITY};
  float l_run[2] = {0.f, 0.f};
  float o[HD / 8][4];
#pragma unroll
  for (int i = 0; i < HD / 8; ++i) o[i][0] = o[i][1] = o[i][2] = o[i][3] = 0.f;
  uint32_t qf[HD / 16][4];

  load_tile(t_begin, 0);
  cp_async_commit();
  const int g = lane >> 2, t4 = lane & 3;
  const int qrow0 = q0 + warp * 16 + g;  // this thread's rows (chunk-relative): qrow0 and qrow0 + 8
  const int qabs0 = qpos0 + warp * 16 + g;  // ... and their absolute positions

  for (int tile = t_begin; tile < t_end; ++tile) {
    const int stage = (tile - t_begin) & 1;
    if (tile + 1 < t_end) {
      load_tile(tile + 1, stage ^ 1);
      cp_async_commit();
      cp_async_wait<1>();
    } else {
      cp_async_wait<0>();
    }
    __syncthreads();
    if (tile == t_begin) {
#pragma unroll
      for (int kk = 0; kk < HD / 16; ++kk) {
        const int mi = lane >> 3;
        ldmatrix_x4(qf[kk], sQ + swz<RB>(warp * 16 + (mi & 1) * 8 + (lane & 7), kk * 2 + (mi >> 1)));
      }
    }
    const uint32_t cK = sK + stage * kTileBytes, cV = sV + stage * kTileBytes;

    float s[8][4];
#pragma unroll
    for (int j = 0; j < 8; ++j) s[j][0] = s[j][1] = s[j][2] = s[j][3] = 0.f;
#pragma unroll
    for (int kk = 0; kk < HD / 16; ++kk) {
#pragma unroll
      for (int jn = 0; jn < 4; ++jn) {
        uint32_t kf[4];
        const int mi = lane >> 3;
        ldmatrix_x4(kf, cK + swz<RB>(jn * 16 + (mi >> 1) * 8 + (lane & 7), kk * 2 + (mi & 1)));
        mma_16816(s[2 * jn], qf[kk], kf[0], kf[1]);
        mma_16816(s[2 * jn + 1], qf[kk], kf[2], kf[3]);
      }
    }
    // causal (+ sliding window) mask; only tiles that touch a boundary need it
    const bool need_mask = (tile * kPfTile + kPfTile - 1 > qpos0) ||
                           (window > 0 && tile * kPfTile < qpos0 + kPfTile - window + 1);
    if (need_mask) {
#pragma unroll
      for (int j = 0; j < 8; ++j)
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const int tok = tile * kPfTile + j * 8 + 2 * t4 + (e & 1);
          const int qr = qabs0 + (e >> 1) * 8;
          if (tok > qr || (window > 0 && tok <= qr - window)) s[j][e] = -INFINITY;
        }
    }
    uint32_t pa[4][4];  // [k-step of 16 tokens][a0..a3]
#pragma unroll
    for (int r = 0; r < 2; ++r) {
      float mx = -INFINITY;
#pragma unroll
      for (int j = 0; j < 8; ++j) mx = fmaxf(mx, fmaxf(s[j][2 * r], s[j][2 * r + 1]));
      mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 1));
      mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 2));
      const float m_new = fmaxf(m_run[r], mx);
      const float m_use = (m_new == -INFINITY) ? 0.f : m_new;
      const float corr = ex2_ftz((m_run[r] - m_use) * scale_log2);
      m_run[r] = m_new;
      float rs = 0.f;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float p0 = ex2_ftz((s[j][2 * r] - m_use) * scale_log2);
        const float p1 = ex2_ftz((s[j][2 * r + 1] - m_use) * scale_log2);
        rs += p0 + p1;
        // n-tile j = tokens 8j..8j+7 -> k-step j/2, low (a0/a1) or high (a2/a3) half
        pa[j >> 1][(j & 1) * 2 + r] = pack_bf16(p0, p1);
      }
      l_run[r] = l_run[r] * corr + rs;
#pragma unroll
      for (int i = 0; i < HD / 8; ++i) {
        o[i][2 * r] *= corr;
        o[i][2 * r + 1] *= corr;
      }
    }
#pragma unroll
    for (int ks = 0; ks < 4; ++ks) {
#pragma unroll
      for (int jp = 0; jp < HD / 16; ++jp) {
        uint32_t vf[4];
        const int mi = lane >> 3;
        ldmatrix_x4_trans(vf, cV + swz<RB>(ks * 16 + (mi & 1) * 8 + (lane & 7), jp * 2 + (mi >> 1)));
        mma_16816(o[2 * jp], pa[ks], vf[0], vf[1]);
        mma_16816(o[2 * jp + 1], pa[ks], vf[2], vf[3]);
      }
    }
    __syncthreads();
  }

#pragma unroll
  for (int r = 0; r < 2; ++r) {
    l_run[r] += __shfl_xor_sync(0xffffffffu, l_run[r], 1);
    l_run[r] += __shfl_xor_sync(0xffffffffu, l_run[r], 2);
    const int qr = qrow0 + r * 8;
    if (qr < T) {
      const float inv = l_run[r] > 0.f ? 1.f / l_run[r] : 0.f;
      bf16* dst = out + ((size_t)b * T + qr) * ((size_t)H * HD) + (size_t)h * HD + 2 * t4;
#pragma unroll
      for (int i = 0; i < HD / 8; ++i)
        *reinterpret_cast<uint32_t*>(dst + i * 8) = pack_bf16(o[i][2 * r] * inv, o[i][2 * r + 1] * inv);
    }
  }
}

template <int HD, bool kPaged>
static int launch_prefill(const void* qkv, void* out, int B, int T, int H, int KV, float scale, int window,
                          cudaStream_t s, const void* k_cache = nullptr, const void* v_cache = nullptr,
                          const int32_t* page_table = nullptr, int max_pages = 0, int page_shift = 0, int prefix = 0) {
  constexpr int kSmem = 5 * kPfTile * HD * 2;
  static PerDeviceOnce once;
  DLI_ONCE_PER_DEVICE(once, dev, DLI_CUDA_CALL(cudaFuncSetAttribute(
                                     attn_prefill_kernel<HD, kPaged>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem)));
  dim3 grid(H, B, (T + kPfTile - 1) / kPfTile);
  DLI_LAUNCH("attn_prefill_kernel", (attn_prefill_kernel<HD, kPaged>), grid, dim3(kPfThreads), (size_t)kSmem, s,
             (const bf16*)qkv, (bf16*)out, T, H, KV, scale * 1.4426950408889634f, window, (const bf16*)k_cache,
             (const bf16*)v_cache, page_table, max_pages, page_shift, prefix);
  return DLI_OK;
}

}  // namespace dli

extern "C" int dli_attn_prefill_bf16(const void* qkv, void* out, int B, int T, int H, int KV, int hd,
                                     float scale, int window, dli_stream_t stream) {
  using namespace dli;
  DLI_CHECK_ARG(qkv && out, "dli_attn_prefill_bf16: null pointer");
  DLI_CHECK_ARG(hd == 64 || hd == 128, "dli_attn_prefill_bf16: hd must be 64 or 128 (got %d)", hd);
  DLI_CHECK_ARG(B >= 0 && T >= 0 && H > 0 && KV > 0 && H % KV == 0 && window >= 0,
                "dli_attn_prefill_bf16: bad sizes B=%d T=%d H=%d KV=%d", B, T, H, KV);
  DLI_CHECK_ARG(B <= 65535 && T <= 65535 * 64, "dli_attn_prefill_bf16: B/T out of grid range");
  DLI_CHECK_ARG(aligned16(qkv) && aligned16(out), "dli_attn_prefill_bf16: pointers must be 16-byte aligned");
  if (B == 0 || T == 0) return DLI_OK;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (hd == 128) return launch_prefill<128, false>(qkv, out, B, T, H, KV, scale, window, s);
  return launch_prefill<64, false>(qkv, out, B, T, H, KV, scale, window, s);
}

namespace dli {
int launch_attn_prefill_tc_paged(const void* qkv, void* out, const void* k_cache, const void* v_cache,
                                 const int32_t* page_table, int max_pages, int num_pages, int B, int T, int prefix,
                                 int H, int KV, int page_size, float scale, int window, cudaStream_t s);
}

extern "C" int dli_attn_prefill_paged_bf16(const void* qkv, void* out, const void* k_cache, const void* v_cache,
                                           const int32_t* page_table, int max_pages, int num_pages, int B, int T,
                                           int prefix, int H, int KV, int hd, int page_size, float scale, int window,
                                           dli_stream_t stream) {
  using namespace dli;
  DLI_CHECK_ARG(qkv && out && k_cache && v_cache && page_table, "dli_attn_prefill_paged_bf16: null pointer");
  DLI_CHECK_ARG(hd == 64 || hd == 128, "dli_attn_prefill_paged_bf16: hd must be 64 or 128 (got %d)", hd);
  DLI_CHECK_ARG(B >= 0 && T >= 0 && prefix >= 0 && H > 0 && KV > 0 && H % KV == 0 && window >= 0 && max_pages > 0,
                "dli_attn_prefill_paged_bf16: bad sizes B=%d T=%d prefix=%d H=%d KV=%d", B, T, prefix, H, KV);
  DLI_CHECK_ARG(page_size >= 8 && (page_size & (page_size - 1)) == 0,
                "dli_attn_prefill_paged_bf16: page_size must be a power of two >= 8 (got %d)", page_size);
  DLI_CHECK_ARG((long long)prefix + T <= (long long)max_pages * page_size,
                "dli_attn_prefill_paged_bf16: prefix + T = %d exceeds the page table (%d pages of %d)", prefix + T,
                max_pages, page_size);
  DLI_CHECK_ARG(B <= 65535 && T <= 65535 * 64, "dli_attn_prefill_paged_bf16: B/T out of grid range");
  DLI_CHECK_ARG(aligned16(qkv) && aligned16(out) && aligned16(k_cache) && aligned16(v_cache),
                "dli_attn_prefill_paged_bf16: pointers must be 16-byte aligned");
  if (B == 0 || T == 0) return DLI_OK;
  int shift = 0;
  while ((1 << shift) < page_size) ++shift;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  // head_dim 128: both attention GEMMs on tcgen05, K/V pages brought in by TMA (num_pages > 0 says how large the
  // pools are); DLI_PREFILL_ATTN=mma keeps the mma.sync kernel
  static const bool use_tc = !(getenv("DLI_PREFILL_ATTN") && getenv("DLI_PREFILL_ATTN")[0] == 'm');
  if (hd == 128 && use_tc && num_pages > 0 && page_size <= 128)
    return launch_attn_prefill_tc_paged(qkv, out, k_cache, v_cache, page_table, max_pages, num_pages, B, T, prefix, H,
                                        KV, page_size, scale, window, s);
  if (hd == 128)
    return launch_prefill<128, true>(qkv, out, B, T, H, KV, scale, window, s, k_cache, v_cache, page_table, max_pages,
                                     shift, prefix);
  return launch_prefill<64, true>(qkv, out, B, T, H, KV, scale, window, s, k_cache, v_cache, page_table, max_pages,
                                  shift, prefix);
}
