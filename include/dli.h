/*
 * dli.h — C-ABI of the B200-native sharded decoder forward.
 *
 * This is the drop-in boundary for ONE hot path of
 * MihirPanpatil/Distributed-LLM-Inferencing: the worker's decoder forward,
 * i.e. what `model.generate(...)` executes at worker/app.py:297-305 and
 * worker/app.py:359-367 (reference paths are relative to the reference repo;
 * `TF:` paths are relative to the `transformers` package the reference calls,
 * which is where the arithmetic lives — SURVEY.md §8c).
 *
 * Conventions
 *   - plain pointers and sizes only; every pointer is CALLER-OWNED DEVICE memory
 *     unless the parameter comment says "host";
 *   - bf16 tensors are `void*` to 2-byte bfloat16, row-major, innermost dim
 *     contiguous, 16-byte aligned base and row strides;
 *   - `stream` is a `cudaStream_t` passed as `void*` (NULL = legacy default);
 *   - every entry point returns DLI_OK or a negative error code and never
 *     throws; `dli_last_error()` gives the thread-local message of the last
 *     failure;
 *   - entry points only ENQUEUE work on `stream`; none synchronises, and none
 *     allocates device memory (scratch is passed in by the caller).
 *   - there is no CPU fallback: without a CUDA device the compute entry
 *     points fail with DLI_ERR_CUDA.
 */
#ifndef DLI_H_
#define DLI_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DLI_OK 0
#define DLI_ERR_INVALID (-1)     /* bad argument (shape, alignment, null) */
#define DLI_ERR_CUDA (-2)        /* a CUDA runtime / driver call failed   */
#define DLI_ERR_UNSUPPORTED (-3) /* shape outside what the kernels cover  */

/* epilogue selectors for the linear (projection) entry points */
#define DLI_EPI_NONE 0      /* y = bf16(acc)                                   */
#define DLI_EPI_BIAS 1      /* y = bf16(acc + bias)          (GPT-2 Conv1D)    */
#define DLI_EPI_RESIDUAL 2  /* y = bf16(res + bf16(acc[+bias]))                */
#define DLI_EPI_GELU_NEW 4  /* y = bf16(gelu_new(bf16(acc[+bias])))  (GPT-2)   */

typedef void* dli_stream_t;

/* ---- library ----------------------------------------------------------- */

/* Thread-local message of the last failing call ("" if none). */
const char* dli_last_error(void);
/* ABI version: major*1000 + minor. */
int dli_version(void);
/* Number of SMs of the current device (used by host code to size grids /
 * split counts); negative error code without a device. */
int dli_sm_count(void);

/* Programmatic dependent launch for every kernel the library enqueues from now
 * on (process-wide; returns the previous setting).  When on, a kernel may begin
 * (and the weight-streaming kernels prefetch their first weight chunk) while its
 * predecessor in the stream drains; each kernel waits for the predecessor's
 * results before reading activations, so results do not change.  Off by
 * default; the host engine turns it on for the decode step. */
int dli_set_pdl(int enable);

/* ---- embedding ---------------------------------------------------------
 * out[r,:] = table[ids[r],:] (+ pos_table[positions[r],:] when pos_table != NULL)
 * Replaces: TF:models/llama/modeling_llama.py:389 (embed_tokens) and
 * TF:models/gpt2/modeling_gpt2.py:577-585 (wte + wpe), reached from
 * worker/app.py:297,359.  ids/positions: int32 [rows].  d % 8 == 0. */
int dli_embedding_bf16(const int32_t* ids, const int32_t* positions,
                       const void* table, const void* pos_table, void* out,
                       int rows, int d, int vocab, dli_stream_t stream);

/* ---- norms with fused residual add --------------------------------------
 * h = residual + x (bf16 add; residual may be NULL => h = x)
 * y = weight * bf16(h * rsqrt(mean(h^2) + eps))           (RMSNorm)
 * If resid_out != NULL, h is written there (may alias `residual` or `x`).
 * Replaces: LlamaRMSNorm TF:models/llama/modeling_llama.py:62-67 plus the
 * residual adds at :325,:331.  d % 8 == 0, d <= 16384. */
int dli_add_rmsnorm_bf16(const void* x, const void* residual, const void* weight,
                         void* resid_out, void* y, int rows, int d, float eps,
                         dli_stream_t stream);

/* Same with LayerNorm (weight and bias), replacing
 * TF:models/gpt2/modeling_gpt2.py:252-254,:505 and the adds at :282,:307. */
int dli_add_layernorm_bf16(const void* x, const void* residual, const void* weight,
                           const void* bias, void* resid_out, void* y, int rows,
                           int d, float eps, dli_stream_t stream);

/* ---- RoPE + paged KV-cache append ---------------------------------------
 * qkv: [rows, (H + 2*KV) * hd] bf16, packed q | k | v per row.
 * Rotates q and k in place with the rotate-half convention
 *   (TF:models/llama/modeling_llama.py:138-168; cos/sin are the bf16 tables
 *   HF computes at :124-135, shape [max_pos, hd/2], passed by the host), then
 * writes k and v of each row into the paged cache, replacing the torch.cat
 * append of TF:cache_utils.py:102-121.
 * cos == NULL => no rotation (GPT-2), append only.
 * positions[r]: absolute position of row r; seq_ids[r]: its sequence (row of
 * page_table).  page_table: int32 [n_seqs, max_pages].
 * k_cache / v_cache: [num_pages, KV, page_size, hd] bf16.
 * hd in {64, 128}; page_size a power of two >= 8. */
int dli_rope_kv_append_bf16(void* qkv, const void* cos, const void* sin,
                            const int32_t* positions, const int32_t* seq_ids,
                            const int32_t* page_table, int max_pages,
                            void* k_cache, void* v_cache, int rows, int H, int KV,
                            int hd, int page_size, dli_stream_t stream);

/* ---- paged decode attention (split-KV, GQA-aware) ------------------------
 * One query token per sequence against its paged KV cache:
 *   out[b, h, :] = softmax(q[b,h,:] . K[b, kv(h), :S_b, :]^T * scale) V
 * Replaces: TF:integrations/sdpa_attention.py:40-104 (default) /
 * eager TF:models/llama/modeling_llama.py:199-221 at T=1, incl. the GQA
 * head grouping (repeat_kv) without materialising it.
 * q: row b at q + b*q_stride elements, [H*hd] (so it can point into qkv).
 * seq_lens[b]: number of cached tokens INCLUDING the one just appended.
 * window > 0: only the last `window` tokens are attended (Mistral sliding
 * window, TF:models/mistral/modeling_mistral.py:172,:372); 0 = full.
 * num_splits >= 1 KV splits per (sequence, kv head); scratch must hold
 * dli_paged_attn_decode_scratch_bytes(...) bytes when num_splits > 1
 * (ignored, may be NULL, when num_splits == 1).
 * out: [B, H*hd] bf16.  hd in {64,128}; H/KV <= 16. */
size_t dli_paged_attn_decode_scratch_bytes(int B, int H, int hd, int num_splits);
int dli_paged_attn_decode_bf16(const void* q, int q_stride, const void* k_cache,
                               const void* v_cache, const int32_t* page_table,
                               int max_pages, const int32_t* seq_lens, void* out,
                               void* scratch, size_t scratch_bytes, int B, int H,
                               int KV, int hd, int page_size, int num_splits,
                               float scale, int window, dli_stream_t stream);

/* Decode attention fused with RoPE and the KV-cache append of the new token
 * (replaces dli_rope_kv_append_bf16 + dli_paged_attn_decode_bf16 in a decode
 * step; same reference lines, same bf16 rounding points, bit-identical result).
 * qkv: packed rows [B, row_stride] holding q | k | v of the NEW token of each
 * sequence, NOT yet rotated; the new token's position is seq_lens[b] - 1.
 * q is rotated on the fly, the new k is rotated and written with v into the
 * sequence's cache page (k_cache / v_cache are written), then attended with
 * the rest of the context.  cos == NULL: no rotation (GPT-2).
 * l2_prefetch_a / _b (optional, may be NULL): device ranges — normally the
 * weights of the projections that follow — pulled into L2 with
 * cp.async.bulk.prefetch by spare CTAs while the attention kernel (a) and the
 * split merge (b) run; at small batch these two kernels are latency-bound and
 * use few SMs, so HBM would otherwise idle.  Ignored when the grid already
 * fills the GPU.  They change no result. */
int dli_paged_attn_decode_fused_bf16(void* qkv, int row_stride, const void* cos,
                                     const void* sin, void* k_cache, void* v_cache,
                                     const int32_t* page_table, int max_pages,
                                     const int32_t* seq_lens, void* out, void* scratch,
                                     size_t scratch_bytes, int B, int H, int KV, int hd,
                                     int page_size, int num_splits, float scale,
                                     int window, const void* l2_prefetch_a,
                                     size_t l2_prefetch_a_bytes, const void* l2_prefetch_b,
                                     size_t l2_prefetch_b_bytes, dli_stream_t stream);

/* Same computation as dli_paged_attn_decode_bf16 for callers that cannot
 * allocate device memory where they call it: a server that handed all of HBM
 * to its weight/KV pools at start-up, a foreign-language binding that owns no
 * device allocator, or a per-token loop that cannot take the device-wide
 * synchronisation of cudaMalloc/cudaFree.  (No counterpart in the reference:
 * it never manages scratch memory itself.)  The split-KV partials go to a
 * fixed array inside the library instead of a caller buffer; q, caches, page
 * table, seq_lens and out are still the caller's.  This entry point calls no
 * cudaMalloc/cudaFree or other allocating routine.
 *   - Limit: dli_paged_attn_decode_scratch_bytes(B, H, hd, num_splits) must not
 *     exceed dli_paged_attn_decode_static_max_bytes() (8 MiB); larger requests
 *     return DLI_ERR_INVALID with the two sizes in dli_last_error().  Every GPU
 *     that loads the library holds those 8 MiB until the process exits.
 *   - The first call on each device looks up the array's device address and
 *     sets the kernel's shared-memory attribute; make that first call (on that
 *     device, with it current) before starting any stream capture.
 *   - All calls on a device share the one array, which keeps the previous
 *     call's contents; nothing needs clearing because each call writes every
 *     slot before reading it.  Two calls that overlap on the GPU (different
 *     streams, or different host threads without ordering) race on it: keep
 *     calls to this entry point ordered on one stream per device. */
size_t dli_paged_attn_decode_static_max_bytes(void);
int dli_paged_attn_decode_static_bf16(const void* q, int q_stride, const void* k_cache,
                                      const void* v_cache, const int32_t* page_table,
                                      int max_pages, const int32_t* seq_lens, void* out,
                                      int B, int H, int KV, int hd, int page_size,
                                      int num_splits, float scale, int window,
                                      dli_stream_t stream);

/* ---- causal prefill attention --------------------------------------------
 * B sequences of T new tokens each, no earlier context: q, k, v are read from
 * the packed (already rotated) qkv buffer [B*T, (H+2KV)*hd].
 * Replaces the same reference code as above at T>1 with a causal mask
 * (TF:masking_utils create_causal_mask, called at
 * TF:models/llama/modeling_llama.py:398-405).  window as above.
 * out: [B*T, H*hd] bf16. */
int dli_attn_prefill_bf16(const void* qkv, void* out, int B, int T, int H, int KV,
                          int hd, float scale, int window, dli_stream_t stream);

/* Same computation on the tensor cores (head_dim 128 only): both GEMMs of flash attention are
 * tcgen05.mma with TMEM accumulators (S = Q K^T, O += P V with V as an MN-major operand), Q/K/V
 * tiles of 128 rows arrive by TMA, the softmax runs one query row per thread out of TMEM. */
int dli_attn_prefill_tc_bf16(const void* qkv, void* out, int B, int T, int H, int KV, int hd,
                             float scale, int window, dli_stream_t stream);

/* Causal attention of a CHUNK of each prompt against the whole context so far
 * (chunked prefill: the prompt is fed T tokens at a time so that pipeline
 * stages can work on different chunks at once).  Same reference
 * lines and arithmetic as dli_attn_prefill_bf16.  qkv holds the chunk's rows
 * (already rotated) at positions prefix .. prefix+T-1 of each sequence; keys
 * and values of positions 0 .. prefix+T-1 are read from the paged cache, so
 * dli_rope_kv_append_bf16 must have appended the chunk first.  page_table:
 * int32 [B, max_pages], row b = sequence b of this call.  prefix == 0 gives
 * the result of dli_attn_prefill_bf16.  num_pages = pages in the k_cache /
 * v_cache pools: with head_dim 128 the tcgen05 kernel reads whole pages by TMA
 * and needs the pool extent (0 = use the mma.sync kernel); cache rows beyond a
 * sequence's end inside its last page must hold finite values (a pool that was
 * zero-filled once is fine: they are multiplied by a probability of exactly 0). */
int dli_attn_prefill_paged_bf16(const void* qkv, void* out, const void* k_cache,
                                const void* v_cache, const int32_t* page_table,
                                int max_pages, int num_pages, int B, int T, int prefix,
                                int H, int KV, int hd, int page_size, float scale,
                                int window, dli_stream_t stream);

/* ---- projections -----------------------------------------------------------
 * y[M,N] = epi(x[M,K] . w[N,K]^T)   — w is the HF nn.Linear layout [out,in].
 * Replaces q/k/v/o_proj TF:models/llama/modeling_llama.py:262-264,:288, the
 * MLP projections :183, lm_head :486-487 and GPT-2's Conv1D
 * (TF:pytorch_utils.py:97-123, weight transposed once at load).
 * `epilogue` is an OR of DLI_EPI_*; bias [N] / residual [M,N] may be NULL when
 * their flag is clear.  y may alias residual.  K % 8 == 0.
 *
 * dli_gemv_bf16   : CUDA-core weight-streaming kernel for M <= 8 (decode).
 * dli_gemm_bf16   : tcgen05/TMEM tensor-core GEMM fed by TMA; any M.
 *                   K % 64 == 0, N % 16 == 0.
 * dli_linear_bf16 : picks between them by M. */
int dli_gemv_bf16(const void* x, const void* w, const void* bias,
                  const void* residual, void* y, int M, int N, int K, int epilogue,
                  dli_stream_t stream);
int dli_gemm_bf16(const void* x, const void* w, const void* bias,
                  const void* residual, void* y, int M, int N, int K, int epilogue,
                  dli_stream_t stream);
int dli_linear_bf16(const void* x, const void* w, const void* bias,
                    const void* residual, void* y, int M, int N, int K,
                    int epilogue, dli_stream_t stream);

/* Gated MLP up-projection: y[M,F] = bf16(silu(bf16(x.Wg^T))) * bf16(x.Wu^T)
 * (TF:models/llama/modeling_llama.py:183, act = SiLU).
 * w_gu: [2F, K] with gate and up rows interleaved in blocks of 128 rows:
 * rows [256j, 256j+128) = gate rows [128j, 128j+128), next 128 = up rows
 * (the host packs this once at load).  F % 128 == 0, K % 64 == 0. */
int dli_gemv_gated_silu_bf16(const void* x, const void* w_gu, void* y, int M, int F,
                             int K, dli_stream_t stream);

/* The same projections for 3..16 activation rows: the weight-streaming loop of
 * the GEMV with its multiply-adds on mma.sync.m16n8k16 (x = A operand, 8 weight
 * rows = B operand, both loaded straight from global memory into fragments), so
 * the kernel stays an HBM stream where the CUDA-core GEMV turns FMA-issue bound.
 * Same reference lines, layouts, epilogues and rounding points as dli_gemv_bf16 /
 * dli_gemv_gated_silu_bf16; M <= 16, K % 8 == 0 (gated: F % 128 == 0).  The
 * summation order differs from the GEMV's (fixed, deterministic). */
int dli_gemv_mma_bf16(const void* x, const void* w, const void* bias,
                      const void* residual, void* y, int M, int N, int K,
                      int epilogue, dli_stream_t stream);
int dli_gemv_mma_gated_silu_bf16(const void* x, const void* w_gu, void* y, int M,
                                 int F, int K, dli_stream_t stream);
int dli_gemm_gated_silu_bf16(const void* x, const void* w_gu, void* y, int M, int F,
                             int K, dli_stream_t stream);
int dli_linear_gated_silu_bf16(const void* x, const void* w_gu, void* y, int M,
                               int F, int K, dli_stream_t stream);

/* The qkv projection of a PROMPT pass fused with RoPE and the KV-cache append:
 *   qkv = x @ w_qkv^T  (rounded to bf16), then exactly what dli_rope_kv_append_bf16
 *   does to that matrix — q and k rotated in place (rotate-half, each product and
 *   the sum rounded to bf16 as apply_rotary_pos_emb does on bf16 tensors), k and v
 *   of every row written to the sequence's cache page — in the GEMM's epilogue,
 *   without a second pass over the [M, (H+2KV)*hd] matrix.  Bit-identical to
 *   dli_gemm_bf16 followed by dli_rope_kv_append_bf16.
 * Replaces: q_proj/k_proj/v_proj + apply_rotary_pos_emb + DynamicCache.update
 *   (TF:models/llama/modeling_llama.py:262-264,138-168; TF:cache_utils.py:102-121).
 * x [M, K], w_qkv [(H+2KV)*hd, K], qkv [M, (H+2KV)*hd]; cos / sin bf16 [max_pos, hd/2];
 * positions / seq_ids int32 [M]; page_table int32 [n_seqs, max_pages];
 * k_cache / v_cache [num_pages, KV, page_size, hd].  Needs hd == 128, H and KV even,
 * M > 128, K % 64 == 0, no bias (anything else: the two separate calls). */
int dli_gemm_qkv_rope_append_bf16(const void* x, const void* w_qkv, void* qkv, int M, int K,
                                  const void* cos, const void* sin, const int32_t* positions,
                                  const int32_t* seq_ids, const int32_t* page_table,
                                  int max_pages, void* k_cache, void* v_cache, int H, int KV,
                                  int hd, int page_size, dli_stream_t stream);

/* Split-K variants of the tensor-core GEMM for batched decode (4 < M <= 128),
 * where the kernel is a weight stream and a matrix with few 128-row tiles would
 * leave most SMs (and most of HBM's bandwidth) idle: K is split over about
 * SMs/tiles CTAs per tile, each parks its fp32 partial in `scratch`, and the CTA
 * that finishes a tile last adds the partials in split order (deterministic) and
 * runs the epilogue.  scratch: dli_gemm_splitk_scratch_bytes(M, N) bytes (N = rows
 * of w; 2F for the gated form), ZERO-FILLED before its first use; every call
 * leaves its counters zero again.  Calls sharing a scratch buffer must be ordered
 * on one stream.  M > 128 or scratch == NULL falls through to dli_gemm_bf16. */
size_t dli_gemm_splitk_scratch_bytes(int M, int N);
int dli_gemm_splitk_bf16(const void* x, const void* w, const void* bias,
                         const void* residual, void* y, int M, int N, int K,
                         int epilogue, void* scratch, size_t scratch_bytes,
                         dli_stream_t stream);
int dli_gemm_splitk_gated_silu_bf16(const void* x, const void* w_gu, void* y, int M,
                                    int F, int K, void* scratch, size_t scratch_bytes,
                                    dli_stream_t stream);

/* RMSNorm fused into the decode projection (M <= 4 rows per pass):
 *   y = epi( (ln_w * bf16(h * rsqrt(mean(h^2) + eps))) . w^T )
 * i.e. LlamaRMSNorm (TF:models/llama/modeling_llama.py:62-67) followed by the
 * q/k/v or gate/up projection (:262-264, :183) in one kernel; the normalised
 * row is built in shared memory with HF's rounding points, so the result is
 * bit-identical to dli_add_rmsnorm_bf16 + dli_gemv_bf16.  K <= 16384. */
int dli_rmsnorm_gemv_bf16(const void* h, const void* ln_w, float eps, const void* w,
                          const void* bias, const void* residual, void* y, int M, int N,
                          int K, int epilogue, dli_stream_t stream);
int dli_rmsnorm_gemv_gated_silu_bf16(const void* h, const void* ln_w, float eps,
                                     const void* w_gu, void* y, int M, int F, int K,
                                     dli_stream_t stream);

/* ---- greedy sampling epilogue ----------------------------------------------
 * out[b] = argmax_v logits[b, v]  (lowest index wins ties, like torch.argmax),
 * replacing the logits -> argmax step TF:generation/utils.py:2762-2793 with
 * do_sample=False.  logits: bf16 [B, V] with row stride `ld` elements. */
int dli_argmax_bf16(const void* logits, int ld, int32_t* out, int B, int V,
                    dli_stream_t stream);

/* Vocabulary-parallel greedy head (pipeline stages each hold a row slice of lm_head):
 * dli_argmax_candidate_bf16 writes, per row b, cand[b] = (best logit as fp32, idx_offset +
 * its index, stored as int32 bits in the second float) for a logits SLICE [B, V_slice];
 * dli_argmax_merge picks, from n such candidate arrays [n, B, 2], the highest logit with
 * the lowest token id on ties — exactly what dli_argmax_bf16 returns on the whole row. */
int dli_argmax_candidate_bf16(const void* logits, int ld, int idx_offset, float* cand, int B,
                              int V, dli_stream_t stream);
int dli_argmax_merge(const float* cands, int n, int B, int32_t* out, dli_stream_t stream);
/* Same, for candidate sets that do not lie back to back: set s starts n_stride floats after set s-1
 * (n_stride >= 2*B) — the per-stage candidates inside the pipeline's all-gathered mailboxes
 * (pipeline.py), read where the exchange put them. */
int dli_argmax_merge_strided(const float* cands, int n, int B, int n_stride, int32_t* out,
                             dli_stream_t stream);

/* lm_head with the greedy choice fused into the weight stream (final norm + lm_head +
 * argmax of TF:models/llama/modeling_llama.py:421,486-487 and TF:generation/utils.py:2762-2793
 * with do_sample=False, as reached from worker/app.py:297-305) for M <= 4 rows:
 *   logits[m, n] = bf16( sum_k xn[m,k] * w[n,k] ),  xn = rmsnorm(h) * ln_w (ln_w NULL: xn = h),
 * exactly as dli_rmsnorm_gemv_bf16 / dli_gemv_bf16 write them, and while a CTA produces its
 * share of the outputs it keeps the best (bf16 logit, n) it has seen; a second, one-warp-per-row
 * kernel reduces the per-CTA candidates (highest logit, lowest n on ties = torch.argmax on the
 * row).  The [M, N] logits are never read back.
 *   tokens   != NULL: tokens[m] = idx_offset + argmax_n logits[m, n]
 *   cand_out != NULL: cand_out[m] = (best logit, idx_offset + n as int32 bits) — one vocabulary
 *                     slice of a pipeline's head, merged across stages by dli_argmax_merge
 * (exactly one of the two).  scratch: dli_gemv_argmax_scratch_bytes(M) bytes of device memory
 * for the per-CTA candidates; every slot read is written by the same call. */
size_t dli_gemv_argmax_scratch_bytes(int M);
int dli_gemv_argmax_bf16(const void* h, const void* ln_w, float eps, const void* w, void* logits,
                         int M, int N, int K, int idx_offset, int32_t* tokens, float* cand_out,
                         void* scratch, size_t scratch_bytes, dli_stream_t stream);

/* The reference's default sampling epilogue (worker/app.py:297-305: do_sample=True,
 * top_k=50, top_p=0.95, temperature=0.8): TemperatureLogitsWarper -> TopKLogitsWarper
 * -> TopPLogitsWarper -> softmax -> multinomial (TF:generation/logits_process.py,
 * TF:generation/utils.py:2762-2793) on fp32 copies of the bf16 logits.  Top-k keeps
 * ties at the threshold, top-p keeps at least one token, exactly as the HF warpers do.
 * The draw is an inverse-CDF lookup with a Philox4x32-10 uniform keyed by
 * (*seed, *step, row): reproducible for a given seed, equal to torch.multinomial in
 * distribution (not in bit pattern: torch uses another stream and algorithm).
 * seed/step are DEVICE pointers so a captured graph can be replayed (NULL = 0).
 * top_k = 0 disables top-k (then V <= 2048); top_k <= 2048.
 * dbg_probs/dbg_ids [B, dbg_cap] (optional): the kept tokens in descending
 * probability and their renormalised probabilities, -1/0 padded (parity tests). */
int dli_sample_bf16(const void* logits, int ld, int32_t* out, int B, int V,
                    float temperature, int top_k, float top_p, const uint64_t* seed,
                    const int32_t* step, float* dbg_probs, int32_t* dbg_ids, int dbg_cap,
                    dli_stream_t stream);

/* The same epilogue for a batch whose rows belong to DIFFERENT requests
 * (continuous batching): seeds[B] / steps[B] give every row its own key — the
 * row index stays out of it, so a request's draws do not depend on the batch
 * row it occupies — and rows with row_mask[b] == 0 (may be NULL = all rows)
 * are left untouched, so greedy requests keep the argmax token already in out. */
int dli_sample_rows_bf16(const void* logits, int ld, int32_t* out, int B, int V,
                         float temperature, int top_k, float top_p,
                         const uint64_t* seeds, const int32_t* steps,
                         const int32_t* row_mask, dli_stream_t stream);

/* Device-side bookkeeping between two decode steps, so a whole step can live in
 * a replayed CUDA graph with no host round trip (the reference syncs the host
 * every token: stopping criteria at TF:generation/utils.py:2804 and the
 * torch.cat of input_ids at :2800): token_log[step*B + b] = tokens[b] (when
 * token_log != NULL and step < log_capacity), positions[b] += 1, seq_lens[b] += 1
 * (each optional; rows whose seq_lens is 0 are idle batch slots and are left alone),
 * then *step += 1.  All int32. */
int dli_decode_advance(const int32_t* tokens, int32_t* token_log, int32_t* positions,
                       int32_t* seq_lens, int32_t* step, int B, int log_capacity,
                       dli_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* DLI_H_ */
