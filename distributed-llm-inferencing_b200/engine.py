"""ShardEngine — one contiguous layer range of a causal LM on one B200.

This is the host-side mirror of what the reference worker does with a shard
(reference worker/app.py:332-372 `run_shard_inference`, intended semantics per
SURVEY.md §0: run layers [start_layer, end_layer] on hidden states and hand the
result to the next shard) and, with the full range, of the unsharded path
(worker/app.py:289-308).  Stage 0 additionally owns the embedding, the last
stage the final norm, lm_head and the greedy argmax.

PyTorch is plumbing here: device memory, streams, CUDA-graph capture.  All
arithmetic goes through the C-ABI kernels (ops.py -> include/dli.h).
"""
from __future__ import annotations

import math
import os
import time
from dataclasses import dataclass
from typing import Dict, List, Optional

import torch

from . import ops

BF = torch.bfloat16
_GEMV_MAX_ROWS = int(os.environ.get("DLI_GEMV_MAX_ROWS", "2"))  # rows served by the CUDA-core GEMV (<= 4)


@dataclass
class ModelSpec:
    """Architecture numbers (same fields as the oracle's spec)."""

    arch: str  # "llama" (Llama / TinyLlama / Mistral) or "gpt2"
    n_layers: int
    d_model: int
    n_heads: int
    n_kv_heads: int
    head_dim: int
    d_ff: int
    vocab: int
    eps: float = 1e-5
    rope_theta: float = 10000.0
    sliding_window: int = 0
    max_pos: int = 8192
    tie_embeddings: bool = False
    # None = plain RoPE; else the checkpoint's scaling dict, rope_type "linear" or "llama3" (rope_inv_freq)
    rope_scaling: Optional[dict] = None

    @staticmethod
    def from_hf_config(cfg) -> "ModelSpec":
        """From a transformers config object or a config.json dict."""
        g = (lambda k, d=None: cfg.get(k, d)) if isinstance(cfg, dict) else (lambda k, d=None: getattr(cfg, k, d))
        mt = g("model_type")
        if mt == "gpt2":
            d, h = g("n_embd"), g("n_head")
            return ModelSpec("gpt2", g("n_layer"), d, h, h, d // h, g("n_inner") or 4 * d, g("vocab_size"),
                             g("layer_norm_epsilon", 1e-5), 0.0, 0, g("n_positions", 1024), True)
        if mt in ("llama", "mistral"):
            d, h = g("hidden_size"), g("num_attention_heads")
            # transformers 5 keeps theta and the scaling keys in `rope_parameters`; older checkpoints
            # (Llama-3.1's config.json) have `rope_theta` + `rope_scaling` with `rope_type` or `type`
            rp = g("rope_parameters") or {}
            rp = dict(rp) if isinstance(rp, dict) else {}
            legacy = g("rope_scaling")
            if isinstance(legacy, dict):
                rp = {**legacy, **rp}
            theta = rp.get("rope_theta") or g("rope_theta", 10000.0)
            kind = rp.get("rope_type") or rp.get("type") or "default"
            scaling = None
            if kind != "default":
                if kind not in ("linear", "llama3"):
                    # a scaled checkpoint run with plain RoPE generates wrong tokens without any error: refuse
                    raise ValueError(f"unsupported rope_type {kind!r}: this engine implements default, linear and "
                                     "llama3 RoPE (TF:modeling_rope_utils.py)")
                scaling = {k: rp[k] for k in ("factor", "low_freq_factor", "high_freq_factor",
                                              "original_max_position_embeddings") if k in rp}
                scaling["rope_type"] = kind
                if kind == "llama3":
                    scaling.setdefault("original_max_position_embeddings", g("max_position_embeddings", 8192))
                    missing = [k for k in ("factor", "low_freq_factor", "high_freq_factor") if k not in scaling]
                    if missing:
                        raise ValueError(f"rope_type llama3 needs {missing} in the checkpoint's rope parameters")
                elif "factor" not in scaling:
                    raise ValueError("rope_type linear needs `factor` in the checkpoint's rope parameters")
            if float(rp.get("partial_rotary_factor", 1.0)) != 1.0:
                raise ValueError("partial_rotary_factor != 1 is not supported")
            win = (g("sliding_window") or 0) if mt == "mistral" else 0
            return ModelSpec("llama", g("num_hidden_layers"), d, h, g("num_key_value_heads") or h,
                             g("head_dim") or d // h, g("intermediate_size"), g("vocab_size"),
                             g("rms_norm_eps", 1e-6), float(theta), int(win), g("max_position_embeddings", 8192),
                             bool(g("tie_word_embeddings", False)), scaling)
        raise ValueError(f"unsupported model_type {mt!r} (llama, mistral, gpt2)")

    def layer_weight_bytes(self) -> int:
        """bf16 bytes of one decoder layer's matrices (SURVEY.md §8d W_layer)."""
        d, hd = self.d_model, self.head_dim
        qkv_o = (self.n_heads + 2 * self.n_kv_heads) * hd * d + self.n_heads * hd * d
        mlp = (3 if self.arch == "llama" else 2) * d * self.d_ff
        return 2 * (qkv_o + mlp)

    def head_weight_bytes(self) -> int:
        return 2 * self.d_model * self.vocab

    def kv_bytes_per_token(self, n_layers: Optional[int] = None) -> int:
        return 2 * 2 * self.n_kv_heads * self.head_dim * (self.n_layers if n_layers is None else n_layers)


# SURVEY.md §A.2 shapes (random-init; there are no checkpoints offline)
NAMED_SPECS = {
    "gpt2": ModelSpec("gpt2", 12, 768, 12, 12, 64, 3072, 50257, 1e-5, 0.0, 0, 1024, True),
    "tinyllama-1.1b": ModelSpec("llama", 22, 2048, 32, 4, 64, 5632, 32000, 1e-5, 10000.0, 0, 2048),
    "llama-3-8b": ModelSpec("llama", 32, 4096, 32, 8, 128, 14336, 128256, 1e-5, 500000.0, 0, 8192),
    "mistral-7b": ModelSpec("llama", 32, 4096, 32, 8, 128, 14336, 32000, 1e-5, 10000.0, 4096, 32768),
    "llama-3-70b": ModelSpec("llama", 80, 8192, 64, 8, 128, 28672, 128256, 1e-5, 500000.0, 0, 8192),
}


def rope_inv_freq(spec: ModelSpec) -> torch.Tensor:
    """fp32 inverse frequencies [hd/2] on the host, op for op what transformers computes:
    default RoPE (TF:models/llama/modeling_llama.py:88-103), `linear` (TF:modeling_rope_utils.py:133-184:
    every frequency divided by `factor`) and `llama3` (TF:modeling_rope_utils.py:550-626: wavelengths above
    original_len/low_freq_factor divided by `factor`, those below original_len/high_freq_factor kept, a
    linear blend between).  None of the three rescales cos/sin (attention_factor 1)."""
    hd = spec.head_dim
    inv_freq = 1.0 / (spec.rope_theta ** (torch.arange(0, hd, 2, dtype=torch.int64).to(torch.float32) / hd))
    sc = spec.rope_scaling
    if not sc:
        return inv_freq
    kind = sc.get("rope_type", "default")
    if kind == "linear":
        return inv_freq / sc["factor"]
    if kind == "llama3":
        factor, lo, hi = sc["factor"], sc["low_freq_factor"], sc["high_freq_factor"]
        old_len = sc["original_max_position_embeddings"]
        lo_wavelen, hi_wavelen = old_len / lo, old_len / hi
        wavelen = 2 * math.pi / inv_freq
        scaled = torch.where(wavelen > lo_wavelen, inv_freq / factor, inv_freq)
        smooth = (old_len / wavelen - lo) / (hi - lo)
        blended = (1 - smooth) * scaled / factor + smooth * scaled
        medium = ~(wavelen < hi_wavelen) * ~(wavelen > lo_wavelen)
        return torch.where(medium, blended, scaled)
    raise ValueError(f"unsupported rope_type {kind!r}")


def rope_tables(spec: ModelSpec, max_pos: int, device) -> tuple:
    """bf16 cos/sin half tables [max_pos, hd/2], computed on the host with the
    exact op sequence of LlamaRotaryEmbedding (TF:models/llama/modeling_llama.py:73-135)
    so they are bit-identical to what the reference's forward multiplies by."""
    inv_freq = rope_inv_freq(spec)
    freqs = torch.arange(max_pos, dtype=torch.float32)[:, None] * inv_freq[None, :]
    return freqs.cos().to(BF).to(device), freqs.sin().to(BF).to(device)


class PagedKVCache:
    """Paged KV cache of one stage: per layer K and V pools
    [num_pages, KV, page_size, hd] plus one page table shared by the stage's
    layers.  Replaces the grow-by-torch.cat cache of TF:cache_utils.py:102-121
    (an O(S) copy per layer per token) with in-place page writes."""

    def __init__(self, n_layers, n_kv, hd, page_size, max_seqs, max_seq_len, device, pool_tokens=None):
        """pool_tokens: size of the page pool in tokens; default max_seqs * max_seq_len (every sequence can reach the
        maximum at once).  A smaller pool is shared on demand: sequences reserve pages as they grow (the continuous
        batcher does) and a reservation that finds the pool empty raises."""
        self.page_size = page_size
        self.max_pages = (max_seq_len + page_size - 1) // page_size
        self.num_pages = max_seqs * self.max_pages if not pool_tokens else \
            max(self.max_pages, (int(pool_tokens) + page_size - 1) // page_size)
        self.k = [torch.zeros(self.num_pages, n_kv, page_size, hd, dtype=BF, device=device) for _ in range(n_layers)]
        self.v = [torch.zeros(self.num_pages, n_kv, page_size, hd, dtype=BF, device=device) for _ in range(n_layers)]
        self.free: List[int] = list(range(self.num_pages - 1, -1, -1))
        self.owned: Dict[int, List[int]] = {}
        self._host_table = torch.zeros(max_seqs, self.max_pages, dtype=torch.int32)
        self.page_table = self._host_table.to(device)

    def capacity(self, seq: int) -> int:
        """Tokens sequence `seq` has pages for."""
        return len(self.owned.get(seq, ())) * self.page_size

    def reserve(self, seq: int, n_tokens: int) -> bool:
        """Make sure sequence `seq` owns pages for `n_tokens` tokens.  Returns True if pages were added (the caller
        then commit()s the page table before the next kernel reads it)."""
        need = (n_tokens + self.page_size - 1) // self.page_size
        if need > self.max_pages:
            raise ValueError(f"sequence of {n_tokens} tokens exceeds max_seq_len of the cache")
        have = self.owned.setdefault(seq, [])
        if len(have) >= need:
            return False
        if need - len(have) > len(self.free):
            raise RuntimeError("KV cache out of pages")
        while len(have) < need:
            if not self.free:
                raise RuntimeError("KV cache out of pages")
            p = self.free.pop()
            self._host_table[seq, len(have)] = p
            have.append(p)
        return True

    def commit(self) -> None:
        self.page_table.copy_(self._host_table, non_blocking=True)

    def release(self, seq: int) -> None:
        for p in self.owned.pop(seq, []):
            self.free.append(p)

    def release_all(self) -> None:
        for s in list(self.owned):
            self.release(s)

    def bytes(self) -> int:
        return sum(t.numel() * 2 for t in self.k) * 2


class ShardEngine:
    def __init__(self, spec: ModelSpec, start_layer: int = 0, end_layer: Optional[int] = None,
                 device: str = "cuda:0", max_batch: int = 1, max_seq_len: int = 2048, page_size: int = 16,
                 use_cuda_graph: bool = True, head_shard: Optional[tuple] = None, kv_pool_tokens: Optional[int] = None):
        """head_shard=(rank, world): vocabulary-parallel lm_head for a pipeline — every stage holds
        rows [rank*ceil(V/world), ...) of the head (the reference's sharder copies the head into
        every shard anyway, shard_model.py:76-78), so the last stage does not carry the whole
        [V, d] matrix on top of its layers; see pipeline.py."""
        if not torch.cuda.is_available():
            raise RuntimeError("ShardEngine needs a CUDA device: this package has no CPU path")
        self.spec = spec
        self.start = start_layer
        self.end = spec.n_layers - 1 if end_layer is None else end_layer
        if not (0 <= self.start <= self.end < spec.n_layers):
            raise ValueError(f"bad layer range [{self.start}, {self.end}] for {spec.n_layers} layers")
        self.is_first = self.start == 0
        self.is_last = self.end == spec.n_layers - 1
        self.head_shard = head_shard
        if head_shard is not None:
            r, w = head_shard
            per = (spec.vocab + w - 1) // w
            self.v0, self.v1 = min(r * per, spec.vocab), min((r + 1) * per, spec.vocab)
        self.has_head = self.is_last and head_shard is None  # whole head + token choice on this stage
        self.n_local = self.end - self.start + 1
        self.device = torch.device(device)
        self.max_batch, self.max_seq_len = max_batch, max_seq_len
        self.use_cuda_graph = use_cuda_graph
        torch.cuda.set_device(self.device)
        self.sms = ops.sm_count()
        self.l2_prefetch = os.environ.get("DLI_L2_PREFETCH", "0") == "1"
        # prompt passes: qkv projection + RoPE + cache append as one kernel (DLI_FUSE_QKV_ROPE=0: three steps as before)
        self.fuse_qkv_rope = os.environ.get("DLI_FUSE_QKV_ROPE", "1") == "1"
        # kernels may start (and prefetch weights) while their predecessor drains
        ops.set_pdl(os.environ.get("DLI_PDL", "1") == "1")
        self.w: Dict[str, torch.Tensor] = {}
        self.layers: List[Dict[str, torch.Tensor]] = []
        self.kv = PagedKVCache(self.n_local, spec.n_kv_heads, spec.head_dim, page_size, max_batch, max_seq_len,
                               self.device, pool_tokens=kv_pool_tokens)
        if spec.arch == "llama":
            self.cos, self.sin = rope_tables(spec, max_seq_len, self.device)
        else:
            self.cos = self.sin = None
        self.scale = spec.head_dim ** -0.5
        self._ws_rows = 0
        self.sampling = None  # None = greedy; else dict(temperature=, top_k=, top_p=) for ops.sample
        # stopping rule of generate(): ids that end a sequence and the filler fed after them (set by the worker
        # from config.json / generation_config.json; empty = run to max_length, what the benchmarks do)
        self.eos_token_ids: tuple = ()
        self.pad_token_id: Optional[int] = None
        # generate(): prompts longer than this are prefilled in chunks of this many tokens (bounded activation
        # workspace; None = whole prompt at once)
        self.prefill_chunk: Optional[int] = None
        self._sk = None  # split-K workspace of the batched-decode projections (zeroed once, self-resetting)
        self._graphs: Dict[tuple, tuple] = {}
        self._dec: Dict[tuple, dict] = {}

    # ------------------------------------------------------------------ weights
    def load_state_dict(self, sd: Dict[str, torch.Tensor]) -> None:
        """Take HF-named tensors (any dtype, any device), keep ONLY this stage's
        layer range (+ embedding on stage 0, final norm/lm_head on the last
        stage): out-of-range tensors of a reference shard directory are random
         and are never read."""
        s, dev = self.spec, self.device

        def t(name, transpose=False):
            x = sd[name]
            if transpose:
                x = x.t()
            return x.to(dtype=BF).contiguous().to(dev)

        self.layers = []
        if s.arch == "llama":
            if self.is_first:
                self.w["embed"] = t("model.embed_tokens.weight")
            for i in range(self.start, self.end + 1):
                p = f"model.layers.{i}."
                qkv = torch.cat([sd[p + f"self_attn.{n}_proj.weight"] for n in "qkv"], dim=0)
                self.layers.append({
                    "ln1": t(p + "input_layernorm.weight"),
                    "wqkv": qkv.to(BF).contiguous().to(dev),
                    "wo": t(p + "self_attn.o_proj.weight"),
                    "ln2": t(p + "post_attention_layernorm.weight"),
                    "wgu": ops.pack_gate_up(sd[p + "mlp.gate_proj.weight"].to(BF), sd[p + "mlp.up_proj.weight"].to(BF)).to(dev),
                    "wdown": t(p + "mlp.down_proj.weight"),
                })
            if self.head_shard is not None:
                full = sd["lm_head.weight"] if ("lm_head.weight" in sd and not s.tie_embeddings) \
                    else sd["model.embed_tokens.weight"]
                self.w["norm"] = t("model.norm.weight")
                self.w["head"] = full[self.v0:self.v1].to(BF).contiguous().to(dev)
            elif self.is_last:
                self.w["norm"] = t("model.norm.weight")
                if "lm_head.weight" in sd and not s.tie_embeddings:
                    self.w["head"] = t("lm_head.weight")
                else:
                    self.w["head"] = self.w["embed"] if self.is_first else t("model.embed_tokens.weight")
        else:
            if self.is_first:
                self.w["embed"] = t("transformer.wte.weight")
                self.w["wpe"] = t("transformer.wpe.weight")
            for i in range(self.start, self.end + 1):
                p = f"transformer.h.{i}."
                # Conv1D stores [in, out] (TF:pytorch_utils.py:112): transpose once to [out, in]
                self.layers.append({
                    "ln1": t(p + "ln_1.weight"), "ln1b": t(p + "ln_1.bias"),
                    "wqkv": t(p + "attn.c_attn.weight", True), "bqkv": t(p + "attn.c_attn.bias"),
                    "wo": t(p + "attn.c_proj.weight", True), "bo": t(p + "attn.c_proj.bias"),
                    "ln2": t(p + "ln_2.weight"), "ln2b": t(p + "ln_2.bias"),
                    "wfc": t(p + "mlp.c_fc.weight", True), "bfc": t(p + "mlp.c_fc.bias"),
                    "wdown": t(p + "mlp.c_proj.weight", True), "bdown": t(p + "mlp.c_proj.bias"),
                })
            if self.is_last:
                self.w["norm"] = t("transformer.ln_f.weight")
                self.w["normb"] = t("transformer.ln_f.bias")
                self.w["head"] = self.w["embed"] if self.is_first else t("transformer.wte.weight")
        self._finalize_layers()

    def load_random(self, seed: int = 0, std: float = 0.02) -> None:
        """Random-init weights generated directly on the device (synthetic benchmark weights of the named
        architecture; no checkpoints offline).  Every tensor group has its own seed derived from (seed,
        layer index) — not from the stage — so a stage of an N-way split holds exactly the weights the
        full-range engine holds for the same layers (bench.py checks pipeline tokens against it)."""
        s, dev = self.spec, self.device

        def gen(tag: int) -> torch.Generator:
            return torch.Generator(device=dev).manual_seed(seed * 1_000_003 + tag)

        def n(g, *shape, mean=0.0):
            out = torch.empty(*shape, dtype=BF, device=dev)
            # fill in slices to bound the fp32 temporary
            flat = out.view(-1)
            step = 1 << 26
            for o in range(0, flat.numel(), step):
                m = min(step, flat.numel() - o)
                flat[o:o + m] = (mean + std * torch.randn(m, device=dev, generator=g)).to(BF)
            return out

        d, hd, ff = s.d_model, s.head_dim, s.d_ff
        self.layers = []
        if s.arch != "llama":
            raise NotImplementedError("load_random covers the llama family; use load_state_dict for gpt2")
        EMBED, NORM, HEAD = 900_001, 900_002, 900_003  # tags beyond any layer index
        if self.is_first:
            self.w["embed"] = n(gen(EMBED), s.vocab, d)
        for i in range(self.start, self.end + 1):
            g = gen(i)
            self.layers.append({
                "ln1": n(g, d, mean=1.0), "wqkv": n(g, (s.n_heads + 2 * s.n_kv_heads) * hd, d),
                "wo": n(g, d, s.n_heads * hd), "ln2": n(g, d, mean=1.0), "wgu": n(g, 2 * ff, d), "wdown": n(g, d, ff),
            })

        def head():
            if s.tie_embeddings:
                return self.w["embed"] if self.is_first else n(gen(EMBED), s.vocab, d)
            return n(gen(HEAD), s.vocab, d)

        if self.head_shard is not None:
            self.w["norm"] = n(gen(NORM), d, mean=1.0)
            full = head()  # the rows [v0, v1) of the SAME matrix the undivided head holds
            self.w["head"] = full[self.v0:self.v1].clone()
            del full
        elif self.is_last:
            self.w["norm"] = n(gen(NORM), d, mean=1.0)
            self.w["head"] = head()
        self._finalize_layers()

    def _finalize_layers(self) -> None:
        """Views used as L2-prefetch ranges: the first ~24 MB of the gate/up matrix is pulled
        into L2 while the (latency-bound) split merge of the attention runs."""
        for L in self.layers:
            if "wgu" in L:
                rows = min(L["wgu"].shape[0], (24 << 20) // (2 * L["wgu"].shape[1]))
                L["wgu_head"] = L["wgu"][:rows]
        self._graphs.clear()

    def weight_bytes(self) -> int:
        seen, total = set(), 0
        for t in list(self.w.values()) + [x for l in self.layers for x in l.values()]:
            if t.data_ptr() not in seen:
                seen.add(t.data_ptr())
                total += t.numel() * t.element_size()
        return total

    def unload(self) -> None:
        """Free weights and KV pages (reference /unload_model, worker/app.py:208-250)."""
        self.w.clear()
        self.layers.clear()
        self._graphs.clear()
        self._dec.clear()
        self.kv.release_all()
        self.kv.k.clear()
        self.kv.v.clear()
        self._ws_rows = 0
        for name in ("_h", "_x", "_qkv", "_attn", "_act"):
            if hasattr(self, name):
                delattr(self, name)
        torch.cuda.empty_cache()

    # ------------------------------------------------------------------ workspace
    def _workspace(self, rows: int) -> None:
        if rows <= self._ws_rows:
            return
        s, dev = self.spec, self.device
        self._h = torch.empty(rows, s.d_model, dtype=BF, device=dev)
        self._x = torch.empty(rows, s.d_model, dtype=BF, device=dev)
        self._qkv = torch.empty(rows, (s.n_heads + 2 * s.n_kv_heads) * s.head_dim, dtype=BF, device=dev)
        self._attn = torch.empty(rows, s.n_heads * s.head_dim, dtype=BF, device=dev)
        self._act = torch.empty(rows, s.d_ff, dtype=BF, device=dev)
        self._ws_rows = rows
        self._graphs.clear()  # captured graphs hold the old workspace pointers

    def _num_splits(self, B: int) -> int:
        """KV splits per (sequence, kv head).  About two CTAs per SM (measured sweet spot of the
        3-stage kernel: B = 4 / 8 / 16 at 2k context are 21 / 15 / 21 % faster with 8 / 4 / 2 splits
        than with one CTA per SM; 32+ sequences need no split), but no finer than 4 tiles (256
        tokens) per split: the merge cost grows with the split count, and at batch 1 the 9 splits
        of a 2k context (72 CTAs, whole split in flight) measured best in the full step."""
        if os.environ.get("DLI_ATTN_SPLITS"):
            return int(os.environ["DLI_ATTN_SPLITS"])
        tiles = (self.max_seq_len + 63) // 64
        want = max(1, 2 * self.sms // (B * self.spec.n_kv_heads))
        return max(1, min((tiles + 3) // 4, want, 64))

    # ------------------------------------------------------------------ one layer range
    def _layers_forward(self, h, rows, positions, seq_ids, seq_lens, prefill_shape, st, h_in=None, h_out=None):
        """h: [rows, d] residual stream (updated in place).  prefill_shape=(B,T) or None for decode.
        h_in: where the stage's INPUT lives when that is not h (read only: the first norm and the first
        residual add read it, e.g. a neighbour's hidden state inside a pipeline mailbox); h_out: where the
        LAST projection writes the stage's output when that is not h."""
        s = self.spec
        res = h if h_in is None else h_in  # residual source until the first projection has written h
        last_li = len(self.layers) - 1
        x, qkv, attn, act = self._x[:rows], self._qkv[:rows], self._attn[:rows], self._act[:rows]
        H, KV, hd = s.n_heads, s.n_kv_heads, s.head_dim
        decode = prefill_shape is None
        # rows <= 2: CUDA-core GEMV with the RMSNorm fused in (6.0-7.3 TB/s); 3..128 rows: split-K tcgen05
        # weight stream on every SM (the GEMV's FMA work grows with the rows: 4.1-5.0 TB/s at 3-4 rows)
        fuse_norm = decode and rows <= _GEMV_MAX_ROWS and s.arch == "llama"
        split_k = decode and _GEMV_MAX_ROWS < rows <= 128
        if split_k and self._sk is None:
            n_max = max(s.vocab, 2 * s.d_ff, (H + 2 * KV) * hd, s.d_model)
            self._sk = torch.zeros(ops.gemm_splitk_scratch_bytes(128, n_max), dtype=torch.uint8, device=self.device)

        def lin(x_, w_, **kw):
            return ops.linear_splitk(x_, w_, self._sk, **kw) if split_k else ops.linear(x_, w_, **kw)

        def gated(x_, w_, out):
            if split_k:
                return ops.linear_splitk_gated_silu(x_, w_, self._sk, out=out)
            return ops.linear_gated_silu(x_, w_, out=out)

        for li, L in enumerate(self.layers):
            dst = h_out if (li == last_li and h_out is not None) else h
            if fuse_norm:
                ops.rmsnorm_linear(res, L["ln1"], s.eps, L["wqkv"], out=qkv)
            elif s.arch == "llama":
                ops.add_rmsnorm(res, L["ln1"], s.eps, out=x)
                # a prompt pass: RoPE and the cache append ride in the projection's epilogue (same bits)
                fused_rope = (not decode and self.fuse_qkv_rope
                              and ops.qkv_rope_append_supported(rows, x.shape[1], H, KV, hd, None))
                if fused_rope:
                    ops.linear_qkv_rope_append(x, L["wqkv"], qkv, self.cos, self.sin, positions, seq_ids,
                                               self.kv.page_table, self.kv.k[li], self.kv.v[li], H, KV, hd)
                else:
                    lin(x, L["wqkv"], out=qkv)
            else:
                ops.add_layernorm(res, L["ln1"], L["ln1b"], s.eps, out=x)
                lin(x, L["wqkv"], bias=L["bqkv"], out=qkv)
            if decode:
                # RoPE + KV append of the new token + split-KV attention in one kernel; the
                # kernel indexes page-table rows by batch row, so it gets the slot's rows
                ops.paged_attn_decode_fused(qkv, self.cos, self.sin, self.kv.k[li], self.kv.v[li], st["page_rows"],
                                            seq_lens, attn, H, KV, hd, self.scale, num_splits=st["splits"],
                                            scratch=st["scratch"], window=s.sliding_window,
                                            # spare SMs pull the next projections' weights into L2 meanwhile
                                            prefetch_a=L["wo"] if self.l2_prefetch else None,
                                            prefetch_b=L.get("wgu_head") if self.l2_prefetch else None)
            else:
                if not (s.arch == "llama" and fused_rope):
                    ops.rope_kv_append(qkv, self.cos, self.sin, positions, seq_ids, self.kv.page_table,
                                       self.kv.k[li], self.kv.v[li], H, KV, hd)
                if len(prefill_shape) > 2:  # a chunk at positions prefix..: keys/values from the paged cache
                    ops.attn_prefill_paged(qkv, attn, self.kv.k[li], self.kv.v[li], st["page_rows"], prefill_shape[0],
                                           prefill_shape[1], prefill_shape[2], H, KV, hd, self.scale,
                                           window=s.sliding_window)
                else:
                    ops.attn_prefill(qkv, attn, prefill_shape[0], prefill_shape[1], H, KV, hd, self.scale,
                                     window=s.sliding_window)
            if s.arch == "llama":
                lin(attn, L["wo"], residual=res, out=h)
                res = h
                if fuse_norm:
                    ops.rmsnorm_linear_gated_silu(h, L["ln2"], s.eps, L["wgu"], out=act)
                else:
                    ops.add_rmsnorm(h, L["ln2"], s.eps, out=x)
                    gated(x, L["wgu"], act)
                lin(act, L["wdown"], residual=h, out=dst)
            else:
                lin(attn, L["wo"], bias=L["bo"], residual=res, out=h)
                res = h
                ops.add_layernorm(h, L["ln2"], L["ln2b"], s.eps, out=x)
                lin(x, L["wfc"], bias=L["bfc"], out=act, epilogue=ops.EPI_GELU_NEW)
                lin(act, L["wdown"], bias=L["bdown"], residual=h, out=dst)
        return h if h_out is None else h_out

    def _head(self, h_last, logits, tokens, st=None):
        """final norm + lm_head + token choice on [B, d]: greedy argmax, or (self.sampling)
        the reference's temperature/top-k/top-p multinomial with the slot's seed and step."""
        s = self.spec
        B = h_last.shape[0]
        xn = self._x[:B]
        rows = st.get("row_sampling") if st is not None else None
        if s.arch == "llama" and B <= _GEMV_MAX_ROWS and rows is None and self.sampling is None:
            # greedy at <= 2 rows: the argmax rides in the lm_head GEMV's epilogue (per-CTA candidates + a
            # one-warp merge), the [B, V] logits are written but never read back
            ops.linear_argmax(h_last, self.w["head"], logits, self._head_scratch(B), tokens=tokens,
                              ln_w=self.w["norm"], eps=s.eps)
            return
        if s.arch == "llama" and B <= _GEMV_MAX_ROWS:
            ops.rmsnorm_linear(h_last, self.w["norm"], s.eps, self.w["head"], out=logits)
        else:
            if s.arch == "llama":
                ops.add_rmsnorm(h_last, self.w["norm"], s.eps, out=xn)
            else:
                ops.add_layernorm(h_last, self.w["norm"], self.w["normb"], s.eps, out=xn)
            ops.linear(xn, self.w["head"], out=logits)
        if rows is not None:
            # rows of different requests (continuous batching): greedy token for every row, then the rows whose
            # request samples are overwritten by a draw keyed with the request's own seed and token position
            ops.argmax(logits, out=tokens)
            ops.sample_rows(logits, tokens, rows["seeds"], rows["steps"], rows["mask"], **rows["params"])
        elif self.sampling is None:
            ops.argmax(logits, out=tokens)
        else:
            ops.sample(logits, out=tokens, seed=st["seed"], step=st["step"], **self.sampling)

    def _head_scratch(self, B: int) -> torch.Tensor:
        """Per-CTA (logit, index) candidates of the fused lm_head argmax (shared by all slots: calls are
        stream-ordered and every slot read is written by the same call)."""
        if getattr(self, "_hs", None) is None or self._hs_rows < B:
            self._hs = torch.empty(max(ops.gemv_argmax_scratch_bytes(B), 16) // 4, dtype=torch.float32,
                                   device=self.device)
            self._hs_rows = B
            self._graphs.clear()
        return self._hs

    # ------------------------------------------------------------------ prefill
    def _state(self, B: int, mb: int = 0) -> dict:
        return self._dec[(B, mb)]

    def start_batch(self, B: int, prompt_len: int, max_total_len: int, mb: int = 0, reserve_len: Optional[int] = None) -> None:
        """Reserve KV pages for the B sequences of micro-batch slot `mb` (sequence ids
        mb*B .. mb*B+B-1) and reset the slot's decode state.  Slot 0 with the full
        batch is the single-GPU case; a pipeline stage keeps one slot per micro-batch
        in flight."""
        if (mb + 1) * B > self.max_batch:
            raise ValueError(f"micro-batch {mb} of {B} sequences exceeds max_batch {self.max_batch}")
        if max_total_len > self.max_seq_len:
            raise ValueError(f"total length {max_total_len} > max_seq_len {self.max_seq_len}")
        for b in range(mb * B, (mb + 1) * B):
            self.kv.release(b)
        for b in range(mb * B, (mb + 1) * B):  # reserve_len < max_total_len: the caller grows the sequence on demand
            self.kv.reserve(b, max_total_len if reserve_len is None else min(max_total_len, reserve_len))
        self.kv.commit()
        self._make_state(B, mb)
        st = self._dec[(B, mb)]
        st["positions"].fill_(prompt_len)
        st["seq_lens"].fill_(prompt_len + 1)
        st["step"].zero_()

    def _make_state(self, B: int, mb: int = 0) -> dict:
        """Allocate (once) the decode state of slot (B, mb); does not touch KV reservations."""
        dev = self.device
        if (B, mb) not in self._dec:
            splits = self._num_splits(B)
            # few CTAs (small batch): the splits of a (sequence, kv head) form a thread-block cluster
            # and merge through distributed shared memory — no scratch buffer, no merge kernel
            cluster = (os.environ.get("DLI_ATTN_CLUSTER", "0") == "1" and splits > 1
                       and min(splits, 8) * self.spec.n_kv_heads * B <= self.sms)
            if cluster:
                splits = min(splits, 8)
            nbytes = ops.attn_decode_scratch_bytes(B, self.spec.n_heads, self.spec.head_dim, splits)
            self._dec[(B, mb)] = {
                "splits": splits,
                "scratch": None if cluster else torch.empty(max(nbytes // 4, 1), dtype=torch.float32, device=dev),
                "positions": torch.zeros(B, dtype=torch.int32, device=dev),
                "seq_lens": torch.zeros(B, dtype=torch.int32, device=dev),
                "seq_ids": torch.arange(mb * B, (mb + 1) * B, dtype=torch.int32, device=dev),
                "page_rows": self.kv.page_table[mb * B:(mb + 1) * B],  # view: rows of this slot's sequences
                "tokens": torch.zeros(B, dtype=torch.int32, device=dev),
                "step": torch.zeros(1, dtype=torch.int32, device=dev),
                "seed": torch.zeros(1, dtype=torch.int64, device=dev),
                # residual stream of the slot's decode step: stage input (recv target) and output (send source)
                "h": torch.empty(B, self.spec.d_model, dtype=BF, device=dev),
                "logits": torch.empty(B, self.spec.vocab, dtype=BF, device=dev) if self.has_head else None,
                # persistent: its address is baked into the captured decode graph
                # the stage that chooses the tokens logs them; stage 0 logs the tokens it consumes (a pipeline head
                # answers the requests, and with a vocabulary-parallel head only it ever sees the final choice)
                "token_log": torch.zeros(self.max_seq_len, B, dtype=torch.int32, device=dev)
                if (self.has_head or self.is_first) else None,
            }
            if self.head_shard is not None:
                st, w = self._dec[(B, mb)], self.head_shard[1]
                st["hf"] = torch.empty(B, self.spec.d_model, dtype=BF, device=dev)  # final hidden of the slot
                st["logits"] = torch.empty(B, self.v1 - self.v0, dtype=BF, device=dev)
                st["cand"] = torch.zeros(B, 2, dtype=torch.float32, device=dev)  # (best logit, global index bits)
                if self.is_first:
                    st["cands"] = torch.zeros(w, B, 2, dtype=torch.float32, device=dev)  # one row per stage
                    st["cand"] = st["cands"][0]  # stage 0 writes its own candidate in place
        return self._dec[(B, mb)]

    def prefill(self, B: int, T: int, ids: Optional[torch.Tensor] = None, hidden: Optional[torch.Tensor] = None,
                mb: int = 0, prefix: Optional[int] = None, final: bool = True):
        """Run the stage over B prompts of T tokens.
        Stage 0 takes `ids` (int32 [B*T] on device), later stages `hidden` [B*T, d].
        Returns hidden [B*T, d] (a view of the workspace), or on the last stage the
        int32 next tokens [B] (logits stay in the slot's state).
        Chunked prefill: `prefix` (not None) says the T tokens are a chunk at positions
        prefix..prefix+T-1 of prompts whose earlier chunks already went through this stage
        (attention then reads keys/values from the paged cache); `final=False` on every chunk
        but the last skips the head and returns the hidden states."""
        rows = B * T
        self._workspace(max(rows, B))
        st = self._dec[(B, mb)]
        dev = self.device
        positions = (torch.arange(T, dtype=torch.int32, device=dev) + (prefix or 0)).repeat(B)
        seq_ids = st["seq_ids"].repeat_interleave(T)
        h = self._h[:rows]
        if self.is_first:
            if self.spec.arch == "gpt2":
                ops.embedding(ids, self.w["embed"], out=h, positions=positions, pos_table=self.w["wpe"])
            else:
                ops.embedding(ids, self.w["embed"], out=h)
        elif hidden.data_ptr() != h.data_ptr():
            h.copy_(hidden)
        self._layers_forward(h, rows, positions, seq_ids, None, (B, T) if prefix is None else (B, T, prefix), st)
        if not self.is_last or not final:
            return h
        last_rows = (torch.arange(B, dtype=torch.int32, device=dev) + 1) * T - 1
        if self.head_shard is not None:
            return ops.embedding(last_rows, h, out=st["hf"])  # final hidden of the last positions -> all stages
        h_last = ops.embedding(last_rows, h, out=st["h"])  # row gather of the last positions
        self._head(h_last, st["logits"], st["tokens"], st)
        return st["tokens"]

    def prefill_chunked(self, B: int, T: int, chunk: int, ids: Optional[torch.Tensor] = None,
                        hidden: Optional[torch.Tensor] = None, mb: int = 0):
        """prefill() fed `chunk` tokens of every prompt at a time (bounded activation memory; in a
        pipeline the stages work on different chunks at once).  ids / hidden are laid out like
        prefill()'s: [B*T(, d)], sequence-major.  Returns what prefill() returns, except that a
        non-last stage returns the hidden states of the LAST chunk only."""
        out = None
        src = ids if self.is_first else hidden
        view = src.view(B, T, *src.shape[1:])
        for c0 in range(0, T, chunk):
            tc = min(chunk, T - c0)
            part = view[:, c0:c0 + tc].reshape(B * tc, *src.shape[1:]).contiguous()
            kw = {"ids": part} if self.is_first else {"hidden": part}
            out = self.prefill(B, tc, mb=mb, prefix=c0, final=c0 + tc >= T, **kw)
        return out

    # ------------------------------------------------------------------ decode
    def _decode_body(self, B: int, mb: int):
        """Stage input: st["tokens"] (stage 0) or st["h"] (later stages, written by the
        caller / the NCCL recv).  Stage output: st["h"] (updated in place) or, on the
        last stage, st["tokens"]."""
        st = self._dec[(B, mb)]
        h = st["h"]
        if self.is_first:
            if self.spec.arch == "gpt2":
                ops.embedding(st["tokens"], self.w["embed"], out=h, positions=st["positions"], pos_table=self.w["wpe"])
            else:
                ops.embedding(st["tokens"], self.w["embed"], out=h)
        self._layers_forward(h, B, st["positions"], st["seq_ids"], st["seq_lens"], None, st)
        if self.has_head:
            self._head(h, st["logits"], st["tokens"], st)
        # head-shard mode: stage 0 logs the token it just consumed (the previous step's choice)
        log = st["token_log"] if (self.has_head or (self.head_shard is not None and self.is_first)) else None
        ops.decode_advance(st["tokens"], log, st["positions"], st["seq_lens"], st["step"])
        return st["tokens"] if self.has_head else h

    def decode_step(self, B: int, hidden_in: Optional[torch.Tensor] = None, mb: int = 0):
        """One token for each of the B sequences of slot `mb`.  The stage's whole step
        (embedding of the previous token / received hidden state .. layers .. argmax ..
        position advance) is a CUDA graph captured once per (B, mb) and replayed, with
        no host synchronisation."""
        self._workspace(B)
        st = self._dec[(B, mb)]
        if hidden_in is not None and hidden_in.data_ptr() != st["h"].data_ptr():
            st["h"].copy_(hidden_in)
        if not self.use_cuda_graph:
            return self._decode_body(B, mb)
        key = (B, mb)
        if key not in self._graphs:
            saved = {k: st[k].clone() for k in ("positions", "seq_lens", "tokens", "step", "h")}
            side = torch.cuda.Stream(device=self.device)
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):
                self._decode_body(B, mb)  # warm-up outside capture (function attributes, occupancy caches)
            torch.cuda.current_stream().wait_stream(side)
            torch.cuda.synchronize()
            for k, v in saved.items():
                st[k].copy_(v)
            g = torch.cuda.CUDAGraph()
            n0 = ops.launch_count
            with torch.cuda.graph(g):
                self._decode_body(B, mb)
            launches = ops.launch_count - n0
            for k, v in saved.items():  # capture does not execute, but keep the state exact
                st[k].copy_(v)
            torch.cuda.synchronize()
            # the warm-up wrote the KV slot of the current position; the replay rewrites it identically
            self._graphs[key] = (g, launches)
        g, launches = self._graphs[key]
        g.replay()
        ops.launch_count += launches
        return st["tokens"] if self.has_head else st["h"]

    # ------------------------------------------------------------------ vocabulary-parallel head
    def head_slice_step(self, B: int, mb: int = 0):
        """This stage's share of the lm_head for slot mb: final norm + the rows [v0, v1) of the head on
        st["hf"], then the slice's best (logit, global token id) into st["cand"].  Graph-captured."""
        st = self._dec[(B, mb)]
        key = (B, mb, "head")

        def body():
            s = self.spec
            if B <= _GEMV_MAX_ROWS and s.arch == "llama":
                ops.linear_argmax(st["hf"], self.w["head"], st["logits"], self._head_scratch(B), cand=st["cand"],
                                  ln_w=self.w["norm"], eps=s.eps, idx_offset=self.v0)
                return
            else:
                xn = self._x[:B]
                if s.arch == "llama":
                    ops.add_rmsnorm(st["hf"], self.w["norm"], s.eps, out=xn)
                else:
                    ops.add_layernorm(st["hf"], self.w["norm"], self.w["normb"], s.eps, out=xn)
                ops.linear(xn, self.w["head"], out=st["logits"])
            ops.argmax_candidate(st["logits"], self.v0, st["cand"])

        self._workspace(B)
        if not self.use_cuda_graph:
            body()
            return st["cand"]
        if key not in self._graphs:
            side = torch.cuda.Stream(device=self.device)
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):
                body()
            torch.cuda.current_stream().wait_stream(side)
            torch.cuda.synchronize()
            g = torch.cuda.CUDAGraph()
            n0 = ops.launch_count
            with torch.cuda.graph(g):
                body()
            self._graphs[key] = (g, ops.launch_count - n0)
        g, launches = self._graphs[key]
        g.replay()
        ops.launch_count += launches
        return st["cand"]

    def combine_candidates(self, B: int, mb: int = 0):
        """Stage 0: global argmax over the stages' candidates -> st["tokens"] (lowest id wins ties)."""
        st = self._dec[(B, mb)]
        ops.argmax_merge(st["cands"], st["tokens"])
        return st["tokens"]

    # ------------------------------------------------------------------ pipeline mailboxes
    # Decode hand-off between pipeline stages without per-pair sends (pipeline.MailboxRunner): every tick each
    # stage contributes ONE fixed-size "mail" — [hidden tile b x d bf16 | head candidate b x (logit, id) | token
    # ids b x int32] — and one all-gather per tick delivers all N mails to everybody.  Tick t writes mail_out[t % P]
    # and the exchange fills mail_in[t % P][src]; the kernels of a slot read and write those buffers in place
    # (their addresses are fixed per slot, so they are baked into the slot's CUDA graphs): no copies either side.
    def attach_mailbox(self, b: int, n_slots: int, period: int, hop: int, rank: int, world: int,
                       with_logits: bool = False) -> None:
        """b sequences per slot, n_slots slots in flight, `period` ticks per round of a slot, `hop` ticks between
        a stage and the next.  The head lives on the last stage (has_head) or is split over the stages (head_shard).
        with_logits (vocabulary-parallel head, one sequence per slot): the mail also carries this stage's slice of
        the logits row, so that stage 0 can SAMPLE from the whole row (the reference's default call) and not only
        take the argmax."""
        d = self.spec.d_model
        off_c = b * d * 2
        off_t = off_c + b * 8
        off_l = (off_t + b * 4 + 15) // 16 * 16
        per = (self.spec.vocab + world - 1) // world  # vocabulary rows per stage (the last one may hold fewer)
        if with_logits and (b != 1 or self.head_shard is None):
            raise ValueError("logits travel in the mail only with a vocabulary-parallel head and one sequence per slot")
        nbytes = ((off_l + per * 2 if with_logits else off_t + b * 4) + 255) // 256 * 256
        dev = self.device
        out = torch.zeros(period, nbytes, dtype=torch.uint8, device=dev)
        inn = torch.zeros(period, world, nbytes, dtype=torch.uint8, device=dev)
        m = {"b": b, "M": n_slots, "P": period, "H": hop, "rank": rank, "N": world, "bytes": nbytes, "out": out, "in": inn,
             "vp": self.head_shard is not None}
        m["out_h"] = [out[p, :off_c].view(BF).view(b, d) for p in range(period)]
        m["out_cand"] = [out[p, off_c:off_t].view(torch.float32).view(b, 2) for p in range(period)]
        m["out_tok"] = [out[p, off_t:off_t + b * 4].view(torch.int32) for p in range(period)]
        m["in_h"] = [[inn[p, r, :off_c].view(BF).view(b, d) for r in range(world)] for p in range(period)]
        m["in_tok"] = [inn[p, world - 1, off_t:off_t + b * 4].view(torch.int32) for p in range(period)]
        m["in_cands_ptr"] = [inn[p].data_ptr() + off_c for p in range(period)]  # set r starts r * nbytes later
        m["with_logits"] = with_logits
        if with_logits:
            m["per"] = per
            m["out_logits"] = [out[p, off_l:off_l + (self.v1 - self.v0) * 2].view(BF).view(1, self.v1 - self.v0)
                               for p in range(period)]
            m["in_logits"] = [inn[p, :, off_l:off_l + per * 2].view(BF) for p in range(period)]  # [N, per], strided
            if self.is_first:
                m["logits_row"] = torch.zeros(1, world * per, dtype=BF, device=dev)  # the slices side by side
        self._mail = m
        self._graphs = {k: v for k, v in self._graphs.items()
                        if not any(isinstance(e, str) and e.startswith("mail") for e in k)}

    def _graphed(self, key, body, restore=()):
        """Run `body` as a CUDA graph captured on first use (after one eager warm-up run whose effect on the
        tensors in `restore` is undone, so the captured state is the state before the call)."""
        if not self.use_cuda_graph:
            body()
            return
        if key not in self._graphs:
            saved = [t.clone() for t in restore]
            side = torch.cuda.Stream(device=self.device)
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):
                body()  # warm-up outside capture (function attributes, occupancy caches, lazy allocations)
            torch.cuda.current_stream().wait_stream(side)
            torch.cuda.synchronize()
            for t, v in zip(restore, saved):
                t.copy_(v)
            g = torch.cuda.CUDAGraph()
            n0 = ops.launch_count
            # thread_local: other threads (NCCL's watchdog polling its events) must not invalidate the capture
            with torch.cuda.graph(g, capture_error_mode="thread_local"):
                body()
            launches = ops.launch_count - n0
            for t, v in zip(restore, saved):
                t.copy_(v)
            torch.cuda.synchronize()
            self._graphs[key] = (g, launches)
        g, launches = self._graphs[key]
        g.replay()
        ops.launch_count += launches

    def mail_phase(self, sigma: int, slot: int) -> int:
        """Tick phase (t mod P) at which the piece with stage offset `sigma` (a stage index; N for the head slices,
        N+1 / N for stage 0 taking the token in) works on `slot`."""
        m = self._mail
        return (slot + m["H"] * sigma) % m["P"]

    def mail_compute(self, mb: int, dry: bool = False):
        """This stage's decode step of slot mb: input from the mailbox the previous stage's mail arrived in (stage 0:
        the slot's token), output into this tick's outgoing mail (last stage with the whole head: the chosen tokens)."""
        m = self._mail
        B, H, P = m["b"], m["H"], m["P"]
        ph = self.mail_phase(m["rank"], mb)
        rd = (ph - H) % P
        st = self._dec[(B, mb)]
        self._workspace(B)

        def body():
            h = st["h"]
            if self.is_first:
                if self.spec.arch == "gpt2":
                    ops.embedding(st["tokens"], self.w["embed"], out=h, positions=st["positions"], pos_table=self.w["wpe"])
                else:
                    ops.embedding(st["tokens"], self.w["embed"], out=h)
                h_in = None
            else:
                h_in = m["in_h"][rd][m["rank"] - 1]
            h_out = None if self.has_head else m["out_h"][ph]
            self._layers_forward(h, B, st["positions"], st["seq_ids"], st["seq_lens"], None, st, h_in=h_in, h_out=h_out)
            tokens = st["tokens"]
            if self.has_head:
                tokens = m["out_tok"][ph]
                self._head(h, st["logits"], tokens, st)
            log = st["token_log"] if (self.has_head or self.is_first) else None
            ops.decode_advance(tokens, log, st["positions"], st["seq_lens"], st["step"])

        if dry:
            return body, [st[k] for k in ("positions", "seq_lens", "step")]
        self._graphed((B, mb, "mail_compute"), body, restore=[st[k] for k in ("positions", "seq_lens", "step")])

    def mail_head(self, mb: int, dry: bool = False):
        """Vocabulary-parallel head: final norm + this stage's rows of the head on slot mb's final hidden state (as it
        arrived from the last stage), the slice's best (logit, token id) into this tick's outgoing mail."""
        m = self._mail
        B, H, P, N = m["b"], m["H"], m["P"], m["N"]
        ph = self.mail_phase(N, mb)
        hf = m["in_h"][(ph - H) % P][N - 1]
        st = self._dec[(B, mb)]
        cand = m["out_cand"][ph]
        logits = m["out_logits"][ph] if m["with_logits"] else st["logits"]
        self._workspace(B)

        def body():
            s = self.spec
            if B <= _GEMV_MAX_ROWS and s.arch == "llama":
                ops.linear_argmax(hf, self.w["head"], logits, self._head_scratch(B), cand=cand,
                                  ln_w=self.w["norm"], eps=s.eps, idx_offset=self.v0)
                return
            xn = self._x[:B]
            if s.arch == "llama":
                ops.add_rmsnorm(hf, self.w["norm"], s.eps, out=xn)
            else:
                ops.add_layernorm(hf, self.w["norm"], self.w["normb"], s.eps, out=xn)
            ops.linear(xn, self.w["head"], out=logits)
            ops.argmax_candidate(logits, self.v0, cand)

        if dry:
            return body, []
        self._graphed((B, mb, "mail_head"), body)

    def mail_ingest(self, mb: int, dry: bool = False, sample: bool = False):
        """Stage 0: slot mb's next input token from the mail that just arrived — the global argmax over the stages'
        candidates (vocabulary-parallel head; lowest id wins ties) or the last stage's chosen token.  sample (mail
        with logits): the stages' logits slices are put side by side and the slot's token is drawn from the whole row
        with the slot's st["row_sampling"] (seed, mask, the reference's temperature / top-k / top-p), keyed by the
        position of the token that produced the logits — the draw the whole head on one stage makes."""
        m = self._mail
        B, H, P, N = m["b"], m["H"], m["P"], m["N"]
        ph = self.mail_phase(N + 1 if m["vp"] else N, mb)
        rd = (ph - H) % P
        st = self._dec[(B, mb)]

        def body():
            if m["vp"]:
                ops.argmax_merge_strided(m["in_cands_ptr"][rd], N, B, m["bytes"] // 4, st["tokens"])
                if sample:
                    rows, per = st["row_sampling"], m["per"]
                    m["logits_row"].view(N, per).copy_(m["in_logits"][rd])  # one strided copy: slice r -> columns r*per..
                    torch.sub(st["positions"], 1, out=rows["steps"])     # stage 0 has already advanced past that token
                    ops.sample_rows(m["logits_row"][:, :self.spec.vocab], st["tokens"], rows["seeds"], rows["steps"],
                                    rows["mask"], **rows["params"])
            else:
                st["tokens"].copy_(m["in_tok"][rd])  # 4*b bytes, device to device

        if dry:
            return body, [st["tokens"]]
        self._graphed((B, mb, "mail_ingest", sample), body, restore=[st["tokens"]])

    def mail_tick(self, ingest: Optional[int], compute: Optional[int], head: Optional[int], sample: bool = False) -> None:
        """Everything this stage does in one tick — token in (stage 0), the stage's decode step, the head slice —
        as ONE CUDA graph per combination of slots (in a full pipeline the combination is a function of the tick's
        phase, so there are P graphs): one submission per tick instead of three, no idle gaps between them."""
        bodies, restore = [], []
        for fn, slot in ((self.mail_ingest, ingest), (self.mail_compute, compute), (self.mail_head, head)):
            if slot is not None:
                b, r = fn(slot, dry=True, sample=sample) if fn == self.mail_ingest else fn(slot, dry=True)
                bodies.append(b)
                restore += r
        if not bodies:
            return

        def body():
            for b in bodies:
                b()

        self._graphed((self._mail["b"], "mail_tick", ingest, compute, head, sample and ingest is not None), body,
                      restore=restore)

    def decode_step_host(self, tokens_in: torch.Tensor, tokens_out: torch.Tensor) -> torch.Tensor:
        """One decode step with HOST buffers, the per-token call a streaming client
        makes: pinned int32 [B] tokens in -> device, one step, next tokens -> pinned
        int32 [B] out, then a stream synchronise (4*B bytes each way)."""
        B = tokens_in.numel()
        st = self._dec[(B, 0)]
        st["tokens"].copy_(tokens_in, non_blocking=True)
        self.decode_step(B)
        tokens_out.copy_(st["tokens"], non_blocking=True)
        torch.cuda.current_stream().synchronize()
        return tokens_out

    # ------------------------------------------------------------------ public API
    @torch.no_grad()
    def generate(self, input_ids: torch.Tensor, max_length: int, return_logits: bool = False,
                 forced: Optional[torch.Tensor] = None, deadline: Optional[float] = None,
                 do_sample: bool = False, temperature: float = 0.8, top_k: int = 50, top_p: float = 0.95,
                 seed: int = 0, on_tokens=None):
        """on_tokens(ids [B, T + k] int64 on the host): called every 8 decode steps with everything generated so far
        (a streaming client's progress; the steps stay on the CUDA-graph path).
        Greedy generate (the parity mode of worker/app.py:297-305: the reference
        samples).  input_ids: int [B, T] on any device;
        `max_length` is the TOTAL length like the reference's.  Returns int64
        [B, max_length] on the host, plus the per-step last-position logits
        (fp32, host) when return_logits.  `forced` [B, n_new] teacher-forces the
        token fed back at each step (parity tests); the returned tokens are still
        this engine's own argmax.  Either option runs the steps un-graphed."""
        if not (self.is_first and self.is_last):
            raise RuntimeError("generate() needs the full layer range; use PipelineEngine for shards")
        B, T = input_ids.shape
        check_max_length(T, max_length)
        n_new = max_length - T
        self.start_batch(B, T, max_length)
        st = self._dec[(B, 0)]
        st.pop("row_sampling", None)  # a per-request mode left by a ContinuousBatcher does not apply here
        # greedy is the parity mode; do_sample=True is the reference's default call
        # (worker/app.py:297-305).  The mode is part of the captured graph.
        mode = dict(temperature=float(temperature), top_k=int(top_k), top_p=float(top_p)) if do_sample else None
        if mode != self.sampling:
            self.sampling = mode
            self._graphs.clear()
        st["seed"].fill_(int(seed))
        st["step"].fill_(-1)  # prefill draws with step -1, decode steps with 0, 1, ...
        # narrowed on the host first: the host->device copy is 4 bytes per prompt token
        ids = input_ids.to(dtype=torch.int32).to(device=self.device, non_blocking=True).reshape(-1).contiguous()
        eager = return_logits or forced is not None
        if forced is not None:
            forced = forced.to(device=self.device, dtype=torch.int32)
        logits_out, toks = [], []
        if self.prefill_chunk and T > self.prefill_chunk:
            toks.append(self.prefill_chunked(B, T, self.prefill_chunk, ids=ids).clone())
        else:
            toks.append(self.prefill(B, T, ids=ids).clone())
        st["step"].zero_()
        if return_logits:
            logits_out.append(st["logits"].float().cpu())
        graph = self.use_cuda_graph
        if eager:
            self.use_cuda_graph = False
        eos = tuple(self.eos_token_ids) if forced is None else ()
        try:
            for i in range(n_new - 1):
                if (deadline is not None or eos or on_tokens is not None) and (i & 7) == 0:
                    # the reference can only check its timeout around generate() (worker/app.py:293,:311);
                    # here it is checked between decode steps, together with the EOS rule
                    torch.cuda.current_stream().synchronize()
                    if deadline is not None and time.time() > deadline:
                        raise TimeoutError("Inference generation timed out")
                    if on_tokens is not None:
                        sofar = torch.stack(toks) if eager else torch.cat((toks[0][None], st["token_log"][:i]), dim=0)
                        on_tokens(torch.cat((input_ids.cpu().long(), sofar.t().cpu().long()), dim=1))
                    if eos:
                        seen = torch.stack(toks) if eager else torch.cat((toks[0][None], st["token_log"][:i]), dim=0)
                        if all_finished(seen, eos):
                            n_new = i + 1  # tokens produced so far: the prefill's + i decode steps
                            break
                if forced is not None:
                    st["tokens"].copy_(forced[:, i])
                self.decode_step(B)
                if eager:
                    toks.append(st["tokens"].clone())
                if return_logits:
                    logits_out.append(st["logits"].float().cpu())
        finally:
            self.use_cuda_graph = graph
        if eager:
            new = torch.stack(toks, dim=1)
        else:
            new = torch.cat((toks[0][None], st["token_log"][: n_new - 1]), dim=0).t()  # [B, n_new]
        new = apply_eos(new.cpu().long()[:, :n_new], eos, self.pad_token_id)
        out = torch.cat((input_ids.cpu().long(), new), dim=1)
        return (out, logits_out) if return_logits else out

    @property
    def last_logits(self):
        return None if not self._dec else next(iter(self._dec.values()))["logits"]


def check_max_length(prompt_len: int, max_length: int) -> None:
    """transformers refuses to generate when the prompt already fills max_length
    (TF:generation/utils.py, _validate_generated_length); the reference worker turns that into a 500
    "Inference failed: Input length of input_ids is ..." — same exception type and message here."""
    if prompt_len >= max_length:
        raise ValueError(f"Input length of input_ids is {prompt_len}, but `max_length` is set to {max_length}. "
                         "This can lead to unexpected behavior. You should consider increasing `max_length` or, "
                         "better yet, setting `max_new_tokens`.")


def apply_eos(new: torch.Tensor, eos_ids, pad_id: Optional[int]) -> torch.Tensor:
    """transformers' stopping rule on the NEW tokens [B, n] (host, int64), as generate() applies it
    (TF:generation/utils.py:2789-2811 with EosTokenCriteria): a sequence is finished once it has produced an
    EOS id; from then on it is fed pad_token_id (eos if no pad id is configured); generation ends when every
    sequence is finished, so the result is cut after the LAST sequence's EOS."""
    if not eos_ids or new.numel() == 0:
        return new
    is_eos = torch.zeros_like(new, dtype=torch.bool)
    for e in eos_ids:
        is_eos |= new == int(e)
    B, n = new.shape
    first = torch.where(is_eos.any(dim=1), is_eos.float().argmax(dim=1), torch.full((B,), n, dtype=torch.long))
    pad = int(eos_ids[0]) if pad_id is None else int(pad_id)
    cols = torch.arange(n)[None, :]
    new = torch.where(cols > first[:, None], torch.full_like(new, pad), new)
    keep = n if bool((first >= n).any()) else int(first.max().item()) + 1
    return new[:, :keep]


def all_finished(tokens_so_far: torch.Tensor, eos_ids) -> bool:
    """tokens_so_far: [steps, B] (any device).  True when every sequence has produced an EOS id."""
    if not eos_ids or tokens_so_far.numel() == 0:
        return False
    hit = torch.zeros_like(tokens_so_far, dtype=torch.bool)
    for e in eos_ids:
        hit |= tokens_so_far == int(e)
    return bool(hit.any(dim=0).all().item())


class LocalChain:
    """Several ShardEngines resident in ONE process, run back to back on one GPU:
    the in-process form of the shard pipeline, used by the worker when the
    `shard_ids` of an /inference request (worker/app.py:260,:274) are all loaded
    here.  The stages must cover layers 0..L-1 without gaps."""

    def __init__(self, engines: List[ShardEngine]):
        engines = sorted(engines, key=lambda e: e.start)
        L = engines[0].spec.n_layers
        if engines[0].start != 0 or engines[-1].end != L - 1 or any(
                a.end + 1 != b.start for a, b in zip(engines, engines[1:])):
            have = [(e.start, e.end) for e in engines]
            raise ValueError(f"shards resident on this worker cover layers {have}, not all of 0..{L - 1}")
        self.engines = engines

    @torch.no_grad()
    def generate(self, input_ids: torch.Tensor, max_length: int, deadline: Optional[float] = None,
                 do_sample: bool = False, temperature: float = 0.8, top_k: int = 50, top_p: float = 0.95,
                 seed: int = 0, on_tokens=None) -> torch.Tensor:
        if len(self.engines) == 1:
            return self.engines[0].generate(input_ids, max_length, deadline=deadline, do_sample=do_sample,
                                            temperature=temperature, top_k=top_k, top_p=top_p, seed=seed,
                                            on_tokens=on_tokens)
        B, T = input_ids.shape
        check_max_length(T, max_length)
        n_new = max_length - T
        first, last = self.engines[0], self.engines[-1]
        for e in self.engines:
            e.start_batch(B, T, max_length)
        mode = dict(temperature=float(temperature), top_k=int(top_k), top_p=float(top_p)) if do_sample else None
        if mode != last.sampling:  # the token choice lives on the last stage
            last.sampling = mode
            last._graphs.clear()
        lst = last._state(B)
        lst["seed"].fill_(int(seed))
        lst["step"].fill_(-1)
        x = input_ids.to(device=first.device, dtype=torch.int32).reshape(-1).contiguous()
        x = first.prefill(B, T, ids=x)
        for e in self.engines[1:]:
            x = e.prefill(B, T, hidden=x)
        lst["step"].zero_()
        toks = [x.clone()]
        eos = tuple(last.eos_token_ids)
        for i in range(n_new - 1):
            if (deadline is not None or eos or on_tokens is not None) and (i & 7) == 0:
                torch.cuda.current_stream().synchronize()
                if deadline is not None and time.time() > deadline:
                    raise TimeoutError("Inference generation timed out")
                if on_tokens is not None:
                    on_tokens(torch.cat((input_ids.cpu().long(), torch.stack(toks, dim=1).cpu().long()), dim=1))
                if eos and all_finished(torch.stack(toks), eos):
                    break
            first._state(B)["tokens"].copy_(toks[-1])
            x = first.decode_step(B)
            for e in self.engines[1:]:
                x = e.decode_step(B, hidden_in=x)
            toks.append(x.clone())
        new = apply_eos(torch.stack(toks, dim=1).cpu().long(), eos, last.pad_token_id)
        return torch.cat((input_ids.cpu().long(), new), dim=1)
