"""Thin torch-tensor front end over the C-ABI (include/dli.h).

torch is used only for device memory and streams; every function passes raw
device pointers to the hand-written kernels and enqueues on torch's current
stream.  Tensors must be CUDA, contiguous where noted, bf16 unless stated.
"""
from __future__ import annotations

import os as _os

import torch

from . import _lib
from ._lib import EPI_BIAS, EPI_GELU_NEW, EPI_NONE, EPI_RESIDUAL  # noqa: F401

# number of kernels this process has launched through the C-ABI (bench.py's gpu_launches)
launch_count = 0


def _p(t):
    return None if t is None else t.data_ptr()


def _stream():
    return torch.cuda.current_stream().cuda_stream


def _req(t, dtype=torch.bfloat16):
    if t is None:
        return
    if not t.is_cuda:
        raise ValueError("dli ops need CUDA tensors (there is no CPU path)")
    if t.dtype != dtype:
        raise ValueError(f"expected {dtype}, got {t.dtype}")
    if not t.is_contiguous():
        raise ValueError("tensor must be contiguous")


def sm_count() -> int:
    n = _lib.load().dli_sm_count()
    if n <= 0:
        _lib.check("dli_sm_count", n)
    return n


def embedding(ids, table, out=None, positions=None, pos_table=None):
    """ids int32 [rows]; table [V, d] -> [rows, d]."""
    global launch_count
    _req(ids, torch.int32), _req(table), _req(pos_table), _req(positions, torch.int32)
    rows, (vocab, d) = ids.numel(), table.shape
    if out is None:
        out = torch.empty(rows, d, dtype=torch.bfloat16, device=table.device)
    _lib.check("dli_embedding_bf16", _lib.load().dli_embedding_bf16(
        _p(ids), _p(positions), _p(table), _p(pos_table), _p(out), rows, d, vocab, _stream()))
    launch_count += 1
    return out


def add_rmsnorm(x, weight, eps, residual=None, resid_out=None, out=None):
    """y = rmsnorm(residual + x) * weight; optionally writes h = residual + x."""
    global launch_count
    _req(x), _req(weight), _req(residual), _req(resid_out)
    rows, d = x.shape
    if out is None:
        out = torch.empty_like(x)
    _lib.check("dli_add_rmsnorm_bf16", _lib.load().dli_add_rmsnorm_bf16(
        _p(x), _p(residual), _p(weight), _p(resid_out), _p(out), rows, d, float(eps), _stream()))
    launch_count += 1
    return out


def add_layernorm(x, weight, bias, eps, residual=None, resid_out=None, out=None):
    global launch_count
    _req(x), _req(weight), _req(bias), _req(residual), _req(resid_out)
    rows, d = x.shape
    if out is None:
        out = torch.empty_like(x)
    _lib.check("dli_add_layernorm_bf16", _lib.load().dli_add_layernorm_bf16(
        _p(x), _p(residual), _p(weight), _p(bias), _p(resid_out), _p(out), rows, d, float(eps), _stream()))
    launch_count += 1
    return out


def rope_kv_append(qkv, cos, sin, positions, seq_ids, page_table, k_cache, v_cache, H, KV, hd):
    """Rotate q,k of packed qkv [rows,(H+2KV)*hd] in place and append k,v to the paged cache."""
    global launch_count
    _req(qkv), _req(cos), _req(sin), _req(k_cache), _req(v_cache)
    _req(positions, torch.int32), _req(seq_ids, torch.int32), _req(page_table, torch.int32)
    rows = qkv.shape[0]
    page_size = k_cache.shape[2]
    _lib.check("dli_rope_kv_append_bf16", _lib.load().dli_rope_kv_append_bf16(
        _p(qkv), _p(cos), _p(sin), _p(positions), _p(seq_ids), _p(page_table), page_table.shape[1],
        _p(k_cache), _p(v_cache), rows, H, KV, hd, page_size, _stream()))
    launch_count += 1


def qkv_rope_append_supported(rows, K, H, KV, hd, bias) -> bool:
    """Can the qkv projection of this prompt pass run as ONE kernel with RoPE and the cache append in its epilogue?"""
    return rows > 128 and hd == 128 and H % 2 == 0 and KV % 2 == 0 and K % 64 == 0 and bias is None


def linear_qkv_rope_append(x, w_qkv, qkv, cos, sin, positions, seq_ids, page_table, k_cache, v_cache, H, KV, hd):
    """qkv = x @ w_qkv.T with q, k rotated and k, v appended to the paged cache in the GEMM's epilogue (a prompt pass;
    the same bits as linear() followed by rope_kv_append())."""
    global launch_count
    _req(x), _req(w_qkv), _req(qkv), _req(cos), _req(sin), _req(k_cache), _req(v_cache)
    _req(positions, torch.int32), _req(seq_ids, torch.int32), _req(page_table, torch.int32)
    M, K = x.shape
    if w_qkv.shape != ((H + 2 * KV) * hd, K) or qkv.shape != (M, (H + 2 * KV) * hd):
        raise ValueError("linear_qkv_rope_append: shapes do not match H, KV, hd")
    _lib.check("dli_gemm_qkv_rope_append_bf16", _lib.load().dli_gemm_qkv_rope_append_bf16(
        _p(x), _p(w_qkv), _p(qkv), M, K, _p(cos), _p(sin), _p(positions), _p(seq_ids), _p(page_table),
        page_table.shape[1], _p(k_cache), _p(v_cache), H, KV, hd, k_cache.shape[2], _stream()))
    launch_count += 1
    return qkv


def attn_decode_scratch_bytes(B, H, hd, num_splits) -> int:
    return int(_lib.load().dli_paged_attn_decode_scratch_bytes(B, H, hd, num_splits))


def paged_attn_decode(q, q_stride, k_cache, v_cache, page_table, seq_lens, out, H, KV, hd,
                      scale, num_splits=1, scratch=None, window=0):
    """q: tensor whose row b starts at q.data_ptr() + b*q_stride elements."""
    global launch_count
    _req(k_cache), _req(v_cache), _req(out)
    _req(page_table, torch.int32), _req(seq_lens, torch.int32)
    B = seq_lens.numel()
    page_size = k_cache.shape[2]
    sbytes = 0 if scratch is None else scratch.numel() * scratch.element_size()
    _lib.check("dli_paged_attn_decode_bf16", _lib.load().dli_paged_attn_decode_bf16(
        _p(q), q_stride, _p(k_cache), _p(v_cache), _p(page_table), page_table.shape[1], _p(seq_lens),
        _p(out), _p(scratch), sbytes, B, H, KV, hd, page_size, num_splits, float(scale), window,
        _stream()))
    launch_count += 2 if num_splits > 1 else 1
    return out


def attn_prefill(qkv, out, B, T, H, KV, hd, scale, window=0, tensor_core=None):
    """Causal prefill attention.  head_dim 128 runs on tcgen05 (dli_attn_prefill_tc_bf16) unless
    tensor_core=False or DLI_PREFILL_ATTN=mma; head_dim 64 uses the mma.sync kernel."""
    global launch_count
    _req(qkv), _req(out)
    if tensor_core is None:
        tensor_core = hd == 128 and _os.environ.get("DLI_PREFILL_ATTN", "tc") != "mma"
    if tensor_core:
        _lib.check("dli_attn_prefill_tc_bf16", _lib.load().dli_attn_prefill_tc_bf16(
            _p(qkv), _p(out), B, T, H, KV, hd, float(scale), window, _stream()))
        launch_count += 1
        return out
    _lib.check("dli_attn_prefill_bf16", _lib.load().dli_attn_prefill_bf16(
        _p(qkv), _p(out), B, T, H, KV, hd, float(scale), window, _stream()))
    launch_count += 1
    return out


def attn_prefill_paged(qkv, out, k_cache, v_cache, page_table, B, T, prefix, H, KV, hd, scale, window=0):
    """Causal attention of a prompt chunk (positions prefix..prefix+T-1) against the paged cache
    (chunked prefill); page_table holds the rows of the B sequences of this call."""
    global launch_count
    _req(qkv), _req(out), _req(k_cache), _req(v_cache), _req(page_table, torch.int32)
    _lib.check("dli_attn_prefill_paged_bf16", _lib.load().dli_attn_prefill_paged_bf16(
        _p(qkv), _p(out), _p(k_cache), _p(v_cache), _p(page_table), page_table.shape[1], k_cache.shape[0], B, T, prefix,
        H, KV, hd, k_cache.shape[2], float(scale), window, _stream()))
    launch_count += 1
    return out


def _linear(fn_name, x, w, bias, residual, out, epilogue):
    global launch_count
    _req(x), _req(w), _req(bias), _req(residual), _req(out)
    M, K = x.shape
    N = w.shape[0]
    if w.shape[1] != K:
        raise ValueError(f"linear: x is [{M},{K}] but w is {tuple(w.shape)}")
    if out is None:
        out = torch.empty(M, N, dtype=torch.bfloat16, device=x.device)
    epi = epilogue | (EPI_BIAS if bias is not None else 0) | (EPI_RESIDUAL if residual is not None else 0)
    fn = getattr(_lib.load(), fn_name)
    _lib.check(fn_name, fn(_p(x), _p(w), _p(bias), _p(residual), _p(out), M, N, K, epi, _stream()))
    launch_count += (M + 3) // 4 if fn_name == "dli_gemv_bf16" else 1
    return out


def linear(x, w, bias=None, residual=None, out=None, epilogue=EPI_NONE):
    """y = epi(x @ w.T); CUDA-core GEMV for M <= 2, tcgen05 GEMM otherwise."""
    return _linear("dli_linear_bf16", x, w, bias, residual, out, epilogue)


def gemv(x, w, bias=None, residual=None, out=None, epilogue=EPI_NONE):
    return _linear("dli_gemv_bf16", x, w, bias, residual, out, epilogue)


def gemv_mma(x, w, bias=None, residual=None, out=None, epilogue=EPI_NONE):
    """Weight stream with mma.sync for 3..16 rows (dli_gemv_mma_bf16)."""
    return _linear("dli_gemv_mma_bf16", x, w, bias, residual, out, epilogue)


def gemm(x, w, bias=None, residual=None, out=None, epilogue=EPI_NONE):
    return _linear("dli_gemm_bf16", x, w, bias, residual, out, epilogue)


def _gated(fn_name, x, w_gu, out):
    global launch_count
    _req(x), _req(w_gu), _req(out)
    M, K = x.shape
    F = w_gu.shape[0] // 2
    if out is None:
        out = torch.empty(M, F, dtype=torch.bfloat16, device=x.device)
    fn = getattr(_lib.load(), fn_name)
    _lib.check(fn_name, fn(_p(x), _p(w_gu), _p(out), M, F, K, _stream()))
    launch_count += (M + 3) // 4 if fn_name == "dli_gemv_gated_silu_bf16" else 1
    return out


def linear_gated_silu(x, w_gu, out=None):
    """silu(x @ Wg.T) * (x @ Wu.T) with w_gu = pack_gate_up(Wg, Wu)."""
    return _gated("dli_linear_gated_silu_bf16", x, w_gu, out)


def gemv_gated_silu(x, w_gu, out=None):
    return _gated("dli_gemv_gated_silu_bf16", x, w_gu, out)


def gemv_mma_gated_silu(x, w_gu, out=None):
    return _gated("dli_gemv_mma_gated_silu_bf16", x, w_gu, out)


def gemm_gated_silu(x, w_gu, out=None):
    return _gated("dli_gemm_gated_silu_bf16", x, w_gu, out)


def pack_gate_up(w_gate, w_up):
    """Interleave gate and up rows in blocks of 128 (layout dli.h documents)."""
    F, K = w_gate.shape
    if F % 128 != 0:
        raise ValueError("intermediate size must be a multiple of 128")
    g = w_gate.reshape(F // 128, 128, K)
    u = w_up.reshape(F // 128, 128, K)
    return torch.stack((g, u), dim=1).reshape(2 * F, K).contiguous()


def argmax(logits, out=None):
    """logits bf16 [B, V] (row stride may exceed V) -> int32 [B]."""
    global launch_count
    if logits.dtype != torch.bfloat16 or not logits.is_cuda or logits.stride(1) != 1:
        raise ValueError("argmax needs a CUDA bf16 tensor with unit inner stride")
    B, V = logits.shape
    if out is None:
        out = torch.empty(B, dtype=torch.int32, device=logits.device)
    _lib.check("dli_argmax_bf16", _lib.load().dli_argmax_bf16(
        _p(logits), logits.stride(0), _p(out), B, V, _stream()))
    launch_count += 1
    return out


def decode_advance(tokens, token_log, positions, seq_lens, step):
    """Log sampled tokens and advance positions / lengths / step counter on device."""
    global launch_count
    for t in (tokens, token_log, positions, seq_lens, step):
        _req(t, torch.int32)
    B = tokens.numel()
    cap = 0 if token_log is None else token_log.numel() // max(B, 1)
    _lib.check("dli_decode_advance", _lib.load().dli_decode_advance(
        _p(tokens), _p(token_log), _p(positions), _p(seq_lens), _p(step), B, cap, _stream()))
    launch_count += 1


def set_pdl(enable: bool) -> bool:
    """Programmatic dependent launch for all subsequent kernel launches; returns the old setting."""
    return bool(_lib.load().dli_set_pdl(1 if enable else 0))


def _bytes(t):
    return 0 if t is None else t.numel() * t.element_size()


def paged_attn_decode_fused(qkv, cos, sin, k_cache, v_cache, page_table, seq_lens, out, H, KV, hd, scale,
                            num_splits=1, scratch=None, window=0, prefetch_a=None, prefetch_b=None):
    """RoPE + KV append of the new token + split-KV decode attention in one kernel (+ merge)."""
    global launch_count
    _req(qkv), _req(cos), _req(sin), _req(k_cache), _req(v_cache), _req(out)
    _req(page_table, torch.int32), _req(seq_lens, torch.int32)
    B = seq_lens.numel()
    page_size = k_cache.shape[2]
    sbytes = 0 if scratch is None else scratch.numel() * scratch.element_size()
    _lib.check("dli_paged_attn_decode_fused_bf16", _lib.load().dli_paged_attn_decode_fused_bf16(
        _p(qkv), qkv.shape[1], _p(cos), _p(sin), _p(k_cache), _p(v_cache), _p(page_table), page_table.shape[1],
        _p(seq_lens), _p(out), _p(scratch), sbytes, B, H, KV, hd, page_size, num_splits, float(scale), window,
        _p(prefetch_a), _bytes(prefetch_a), _p(prefetch_b), _bytes(prefetch_b), _stream()))
    launch_count += 2 if (num_splits > 1 and scratch is not None) else 1  # cluster merge: one kernel
    return out


def rmsnorm_linear(h, ln_w, eps, w, bias=None, residual=None, out=None, epilogue=EPI_NONE):
    """epi(rmsnorm(h) * ln_w @ w.T) in one weight-streaming kernel (decode, M <= 4 per pass)."""
    global launch_count
    _req(h), _req(ln_w), _req(w), _req(bias), _req(residual), _req(out)
    M, K = h.shape
    N = w.shape[0]
    if out is None:
        out = torch.empty(M, N, dtype=torch.bfloat16, device=h.device)
    epi = epilogue | (EPI_BIAS if bias is not None else 0) | (EPI_RESIDUAL if residual is not None else 0)
    _lib.check("dli_rmsnorm_gemv_bf16", _lib.load().dli_rmsnorm_gemv_bf16(
        _p(h), _p(ln_w), float(eps), _p(w), _p(bias), _p(residual), _p(out), M, N, K, epi, _stream()))
    launch_count += (M + 3) // 4
    return out


def rmsnorm_linear_gated_silu(h, ln_w, eps, w_gu, out=None):
    global launch_count
    _req(h), _req(ln_w), _req(w_gu), _req(out)
    M, K = h.shape
    F = w_gu.shape[0] // 2
    if out is None:
        out = torch.empty(M, F, dtype=torch.bfloat16, device=h.device)
    _lib.check("dli_rmsnorm_gemv_gated_silu_bf16", _lib.load().dli_rmsnorm_gemv_gated_silu_bf16(
        _p(h), _p(ln_w), float(eps), _p(w_gu), _p(out), M, F, K, _stream()))
    launch_count += (M + 3) // 4
    return out


def gemm_splitk_scratch_bytes(M, N) -> int:
    return int(_lib.load().dli_gemm_splitk_scratch_bytes(M, N))


def linear_splitk(x, w, scratch, bias=None, residual=None, out=None, epilogue=EPI_NONE):
    """Batched-decode projection (4 < M <= 128): split-K tcgen05 weight stream.  `scratch` is a
    zero-initialised uint8/float tensor of gemm_splitk_scratch_bytes(M, N) bytes."""
    global launch_count
    _req(x), _req(w), _req(bias), _req(residual), _req(out)
    M, K = x.shape
    N = w.shape[0]
    if out is None:
        out = torch.empty(M, N, dtype=torch.bfloat16, device=x.device)
    epi = epilogue | (EPI_BIAS if bias is not None else 0) | (EPI_RESIDUAL if residual is not None else 0)
    _lib.check("dli_gemm_splitk_bf16", _lib.load().dli_gemm_splitk_bf16(
        _p(x), _p(w), _p(bias), _p(residual), _p(out), M, N, K, epi, _p(scratch), _bytes(scratch), _stream()))
    launch_count += 1
    return out


def linear_splitk_gated_silu(x, w_gu, scratch, out=None):
    global launch_count
    _req(x), _req(w_gu), _req(out)
    M, K = x.shape
    F = w_gu.shape[0] // 2
    if out is None:
        out = torch.empty(M, F, dtype=torch.bfloat16, device=x.device)
    _lib.check("dli_gemm_splitk_gated_silu_bf16", _lib.load().dli_gemm_splitk_gated_silu_bf16(
        _p(x), _p(w_gu), _p(out), M, F, K, _p(scratch), _bytes(scratch), _stream()))
    launch_count += 1
    return out


def sample(logits, out=None, temperature=0.8, top_k=50, top_p=0.95, seed=None, step=None, debug_cap=0):
    """Temperature -> top-k -> top-p -> multinomial on bf16 logits [B, V].
    seed: int64 device tensor [1]; step: int32 device tensor [1] (both optional).
    debug_cap > 0 additionally returns (probs [B, cap] fp32, ids [B, cap] int32) of the kept set."""
    global launch_count
    if logits.dtype != torch.bfloat16 or not logits.is_cuda or logits.stride(1) != 1:
        raise ValueError("sample needs a CUDA bf16 tensor with unit inner stride")
    _req(seed, torch.int64), _req(step, torch.int32)
    B, V = logits.shape
    if out is None:
        out = torch.empty(B, dtype=torch.int32, device=logits.device)
    probs = ids = None
    if debug_cap:
        probs = torch.empty(B, debug_cap, dtype=torch.float32, device=logits.device)
        ids = torch.empty(B, debug_cap, dtype=torch.int32, device=logits.device)
    _lib.check("dli_sample_bf16", _lib.load().dli_sample_bf16(
        _p(logits), logits.stride(0), _p(out), B, V, float(temperature), int(top_k), float(top_p), _p(seed), _p(step),
        _p(probs), _p(ids), int(debug_cap), _stream()))
    launch_count += 1
    return (out, probs, ids) if debug_cap else out


def sample_rows(logits, out, seeds, steps, row_mask=None, temperature=0.8, top_k=50, top_p=0.95):
    """sample() for rows of different requests: per-row seeds (int64 [B]) and steps (int32 [B]); rows whose
    row_mask (int32 [B]) is 0 keep the token already in `out`."""
    global launch_count
    if logits.dtype != torch.bfloat16 or not logits.is_cuda or logits.stride(1) != 1:
        raise ValueError("sample_rows needs a CUDA bf16 tensor with unit inner stride")
    _req(seeds, torch.int64), _req(steps, torch.int32), _req(row_mask, torch.int32), _req(out, torch.int32)
    B, V = logits.shape
    _lib.check("dli_sample_rows_bf16", _lib.load().dli_sample_rows_bf16(
        _p(logits), logits.stride(0), _p(out), B, V, float(temperature), int(top_k), float(top_p), _p(seeds), _p(steps),
        _p(row_mask), _stream()))
    launch_count += 1
    return out


def argmax_candidate(logits, idx_offset, cand):
    """Best (logit, idx_offset + index) of each row of a logits slice -> cand [B, 2] float32."""
    global launch_count
    _req(cand, torch.float32)
    B, V = logits.shape
    _lib.check("dli_argmax_candidate_bf16", _lib.load().dli_argmax_candidate_bf16(
        _p(logits), logits.stride(0), int(idx_offset), _p(cand), B, V, _stream()))
    launch_count += 1
    return cand


def argmax_merge(cands, out):
    """cands [n, B, 2] float32 (per-slice candidates) -> out int32 [B]: the global greedy token."""
    global launch_count
    _req(cands, torch.float32), _req(out, torch.int32)
    n, B = cands.shape[0], cands.shape[1]
    _lib.check("dli_argmax_merge", _lib.load().dli_argmax_merge(_p(cands), n, B, _p(out), _stream()))
    launch_count += 1
    return out


def argmax_merge_strided(cands_ptr: int, n: int, B: int, n_stride: int, out):
    """argmax_merge over n candidate sets [B, 2] float32 that start n_stride floats apart (raw device pointer:
    the sets sit inside the pipeline's mailboxes)."""
    global launch_count
    _req(out, torch.int32)
    _lib.check("dli_argmax_merge_strided", _lib.load().dli_argmax_merge_strided(cands_ptr, n, B, n_stride, _p(out), _stream()))
    launch_count += 1
    return out


def gemv_argmax_scratch_bytes(M) -> int:
    return int(_lib.load().dli_gemv_argmax_scratch_bytes(M))


def linear_argmax(h, w, logits, scratch, tokens=None, cand=None, ln_w=None, eps=0.0, idx_offset=0):
    """lm_head for <= 4 rows with the greedy choice fused into the weight stream: logits = (rmsnorm(h) * ln_w
    if ln_w is given else h) @ w.T, and tokens[m] = idx_offset + argmax (or `cand` [M, 2] float32 = (best logit,
    idx_offset + index) for one vocabulary slice).  `scratch`: gemv_argmax_scratch_bytes(M) bytes."""
    global launch_count
    _req(h), _req(w), _req(logits), _req(ln_w), _req(tokens, torch.int32), _req(cand, torch.float32)
    M, K = h.shape
    N = w.shape[0]
    _lib.check("dli_gemv_argmax_bf16", _lib.load().dli_gemv_argmax_bf16(
        _p(h), _p(ln_w), float(eps), _p(w), _p(logits), M, N, K, int(idx_offset), _p(tokens), _p(cand), _p(scratch),
        _bytes(scratch), _stream()))
    launch_count += 2
    return tokens if tokens is not None else cand
