"""ORACLE — test infrastructure only, never the product path.

CPU restatement (plain PyTorch ops, op by op) of the decoder forward that the
reference worker executes through `model.generate(...)` at
reference worker/app.py:297-305 and :359-367.  The arithmetic itself lives in the
un-vendored third-party dependency `transformers` (pinned 4.30.2 in
worker/requirements.txt:2; the pin predates GQA / rope_theta / Mistral and is not
installable offline, so — per SURVEY.md §8c — the restatement follows the
installed transformers 5.5.0, cited as `TF:` paths below).

Pinning: tests/test_oracle.py checks this file against (a) transformers itself,
imported in the test process, on the same seeded weights (logits, hidden states
at shard boundaries, greedy tokens) and (b) the committed fixtures in
tests/golden/ produced by oracle/gen_golden.py from transformers and from the
reference's own worker/app.py.  The reference repository contains no tests or
golden vectors of its own, so transformers is the only pin.

Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs may import
this module.

Rounding points reproduced (bf16 runs): RMSNorm in fp32 -> bf16 -> * weight;
cos/sin computed in fp32 then cast; every Linear output rounded to the model
dtype; residual stream in the model dtype; logits cast to fp32 for argmax.
Attention is evaluated in fp32 and rounded once (the reference's default SDPA
kernels keep fp32 accumulators internally and differ per backend, so no single
bit pattern exists to copy; the tolerance is stated in the tests).
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Tuple

import torch
import torch.nn.functional as F


@dataclass
class ModelSpec:
    """Architecture numbers the forward needs."""

    arch: str  # "llama" (Llama/TinyLlama/Mistral) or "gpt2"
    n_layers: int
    d_model: int
    n_heads: int
    n_kv_heads: int
    head_dim: int
    d_ff: int
    vocab: int
    eps: float = 1e-5
    rope_theta: float = 10000.0
    sliding_window: int = 0  # 0 = full attention
    max_pos: int = 8192
    tie_embeddings: bool = False
    rope_scaling: Optional[dict] = None  # {"rope_type": "linear" | "llama3", "factor": ..., ...}

    @staticmethod
    def from_hf(cfg) -> "ModelSpec":
        mt = cfg.model_type
        if mt == "gpt2":
            return ModelSpec("gpt2", cfg.n_layer, cfg.n_embd, cfg.n_head, cfg.n_head, cfg.n_embd // cfg.n_head,
                             cfg.n_inner or 4 * cfg.n_embd, cfg.vocab_size, cfg.layer_norm_epsilon, 0.0, 0,
                             cfg.n_positions, True)
        if mt in ("llama", "mistral"):
            hd = getattr(cfg, "head_dim", None) or cfg.hidden_size // cfg.num_attention_heads
            rp = getattr(cfg, "rope_parameters", None) or {}
            theta = rp.get("rope_theta", None) if isinstance(rp, dict) else None
            theta = theta or getattr(cfg, "rope_theta", 10000.0)
            win = getattr(cfg, "sliding_window", None) or 0
            scaling = None
            if isinstance(rp, dict) and rp.get("rope_type", "default") != "default":
                scaling = {k: v for k, v in rp.items() if k != "rope_theta"}
            return ModelSpec("llama", cfg.num_hidden_layers, cfg.hidden_size, cfg.num_attention_heads,
                             cfg.num_key_value_heads, hd, cfg.intermediate_size, cfg.vocab_size, cfg.rms_norm_eps,
                             float(theta), int(win) if mt == "mistral" else 0, cfg.max_position_embeddings,
                             bool(getattr(cfg, "tie_word_embeddings", False)), scaling)
        raise ValueError(f"unsupported model_type {mt}")


def shard_ranges(n_layers: int, n_shards: int) -> List[Tuple[int, int]]:
    """Inclusive (start_layer, end_layer) per shard — the reference's formula:
    L // N layers each, the whole remainder on the last shard
    (master/dashboard/management/commands/shard_model.py:55-66, :99-106)."""
    per, rem = n_layers // n_shards, n_layers % n_shards
    out = []
    for i in range(n_shards):
        start = i * per
        end = start + per + (rem if i == n_shards - 1 else 0)
        out.append((start, end - 1))
    return out


# --------------------------------------------------------------------------- pieces
def rmsnorm(h: torch.Tensor, w: torch.Tensor, eps: float) -> torch.Tensor:
    """LlamaRMSNorm.forward, TF:models/llama/modeling_llama.py:62-67."""
    dt = h.dtype
    x = h.to(torch.float32)
    var = x.pow(2).mean(-1, keepdim=True)
    x = x * torch.rsqrt(var + eps)
    return w * x.to(dt)


def rope_cos_sin(spec: ModelSpec, positions: torch.Tensor, dtype) -> Tuple[torch.Tensor, torch.Tensor]:
    """LlamaRotaryEmbedding (default rope), TF:models/llama/modeling_llama.py:73-135.
    Returns the half tables [T, hd/2] (HF concatenates the same half twice)."""
    hd = spec.head_dim
    inv_freq = 1.0 / (spec.rope_theta ** (torch.arange(0, hd, 2, dtype=torch.int64).to(torch.float32) / hd))
    sc = spec.rope_scaling or {}
    if sc.get("rope_type") == "linear":  # TF:modeling_rope_utils.py:133-184
        inv_freq = inv_freq / sc["factor"]
    elif sc.get("rope_type") == "llama3":  # TF:modeling_rope_utils.py:550-626
        n0 = sc["original_max_position_embeddings"]
        wl = 2 * math.pi / inv_freq
        base = torch.where(wl > n0 / sc["low_freq_factor"], inv_freq / sc["factor"], inv_freq)
        t = (n0 / wl - sc["low_freq_factor"]) / (sc["high_freq_factor"] - sc["low_freq_factor"])
        mid = ~(wl < n0 / sc["high_freq_factor"]) * ~(wl > n0 / sc["low_freq_factor"])
        inv_freq = torch.where(mid, (1 - t) * base / sc["factor"] + t * base, base)
    elif sc:
        raise ValueError(f"oracle: unsupported rope scaling {sc}")
    freqs = positions.to(torch.float32)[:, None] * inv_freq[None, :]
    return freqs.cos().to(dtype), freqs.sin().to(dtype)


def _rotate_half(x):
    h = x.shape[-1] // 2
    return torch.cat((-x[..., h:], x[..., :h]), dim=-1)


def apply_rope(x: torch.Tensor, cos: torch.Tensor, sin: torch.Tensor) -> torch.Tensor:
    """apply_rotary_pos_emb, TF:models/llama/modeling_llama.py:138-168.
    x: [B, T, heads, hd]; cos/sin half tables [T, hd/2]."""
    c = torch.cat((cos, cos), -1)[None, :, None, :]
    s = torch.cat((sin, sin), -1)[None, :, None, :]
    return x * c + _rotate_half(x) * s


def attention(q, k, v, q_pos0: int, scale: float, window: int) -> torch.Tensor:
    """Causal softmax(QK^T*scale)V with GQA head grouping; fp32 math, one rounding.
    Follows TF:integrations/sdpa_attention.py:40-104 / eager
    TF:models/llama/modeling_llama.py:199-221 (mask from create_causal_mask,
    sliding window per TF:models/mistral/modeling_mistral.py:172,:372).
    q: [B,T,H,hd]; k,v: [B,S,KV,hd] where S = q_pos0 + T."""
    B, T, H, hd = q.shape
    S, KV = k.shape[1], k.shape[2]
    G = H // KV
    qf = q.to(torch.float32).permute(0, 2, 1, 3).reshape(B, KV, G, T, hd)
    kf = k.to(torch.float32).permute(0, 2, 1, 3)  # [B,KV,S,hd]
    vf = v.to(torch.float32).permute(0, 2, 1, 3)
    s = torch.einsum("bkgtd,bksd->bkgts", qf, kf) * scale
    qi = torch.arange(q_pos0, q_pos0 + T)[:, None]
    kj = torch.arange(S)[None, :]
    mask = kj > qi
    if window:
        mask = mask | (kj <= qi - window)
    s = s.masked_fill(mask, float("-inf"))
    o = torch.einsum("bkgts,bksd->bkgtd", torch.softmax(s, dim=-1), vf)
    return o.reshape(B, H, T, hd).permute(0, 2, 1, 3).reshape(B, T, H * hd).to(q.dtype)


def gelu_new(x):
    """NewGELUActivation, TF:activations.py:59-66 (op by op in x.dtype)."""
    return 0.5 * x * (1.0 + torch.tanh(math.sqrt(2.0 / math.pi) * (x + 0.044715 * torch.pow(x, 3.0))))


@dataclass
class KVCache:
    """Grow-by-concatenation cache, TF:cache_utils.py:102-121 (per layer)."""

    k: Dict[int, torch.Tensor] = field(default_factory=dict)
    v: Dict[int, torch.Tensor] = field(default_factory=dict)

    def update(self, layer, k, v):
        if layer in self.k:
            self.k[layer] = torch.cat((self.k[layer], k), dim=1)
            self.v[layer] = torch.cat((self.v[layer], v), dim=1)
        else:
            self.k[layer], self.v[layer] = k, v
        return self.k[layer], self.v[layer]

    def length(self):
        return next(iter(self.k.values())).shape[1] if self.k else 0


# --------------------------------------------------------------------------- layers
def llama_layer(h, sd, i, spec: ModelSpec, cos, sin, cache: KVCache, pos0: int):
    """LlamaDecoderLayer.forward, TF:models/llama/modeling_llama.py:303-332 with
    LlamaAttention.forward :262-289 and LlamaMLP.forward :183."""
    p = f"model.layers.{i}."
    B, T, _ = h.shape
    H, KV, hd = spec.n_heads, spec.n_kv_heads, spec.head_dim
    x = rmsnorm(h, sd[p + "input_layernorm.weight"], spec.eps)
    q = F.linear(x, sd[p + "self_attn.q_proj.weight"]).view(B, T, H, hd)
    k = F.linear(x, sd[p + "self_attn.k_proj.weight"]).view(B, T, KV, hd)
    v = F.linear(x, sd[p + "self_attn.v_proj.weight"]).view(B, T, KV, hd)
    q, k = apply_rope(q, cos, sin), apply_rope(k, cos, sin)
    k_all, v_all = cache.update(i, k, v)
    a = attention(q, k_all, v_all, pos0, hd ** -0.5, spec.sliding_window)
    h = h + F.linear(a, sd[p + "self_attn.o_proj.weight"])
    x = rmsnorm(h, sd[p + "post_attention_layernorm.weight"], spec.eps)
    g = F.linear(x, sd[p + "mlp.gate_proj.weight"])
    u = F.linear(x, sd[p + "mlp.up_proj.weight"])
    h = h + F.linear(F.silu(g) * u, sd[p + "mlp.down_proj.weight"])
    return h


def gpt2_layer(h, sd, i, spec: ModelSpec, cache: KVCache, pos0: int):
    """GPT2Block.forward, TF:models/gpt2/modeling_gpt2.py:246-309; Conv1D
    (weights stored [in,out]) TF:pytorch_utils.py:97-123; MLP :229-243."""
    p = f"transformer.h.{i}."
    B, T, d = h.shape
    H, hd = spec.n_heads, spec.head_dim
    x = F.layer_norm(h, (d,), sd[p + "ln_1.weight"], sd[p + "ln_1.bias"], spec.eps)
    qkv = torch.addmm(sd[p + "attn.c_attn.bias"], x.view(-1, d), sd[p + "attn.c_attn.weight"]).view(B, T, 3 * d)
    q, k, v = (t.view(B, T, H, hd) for t in qkv.split(d, dim=2))
    k_all, v_all = cache.update(i, k, v)
    a = attention(q, k_all, v_all, pos0, hd ** -0.5, 0)
    a = torch.addmm(sd[p + "attn.c_proj.bias"], a.view(-1, d), sd[p + "attn.c_proj.weight"]).view(B, T, d)
    h = h + a
    x = F.layer_norm(h, (d,), sd[p + "ln_2.weight"], sd[p + "ln_2.bias"], spec.eps)
    m = torch.addmm(sd[p + "mlp.c_fc.bias"], x.view(-1, d), sd[p + "mlp.c_fc.weight"])
    m = gelu_new(m)
    m = torch.addmm(sd[p + "mlp.c_proj.bias"], m, sd[p + "mlp.c_proj.weight"]).view(B, T, d)
    return h + m


# --------------------------------------------------------------------------- shard forward
def embed(ids: torch.Tensor, sd, spec: ModelSpec, pos0: int) -> torch.Tensor:
    """TF:models/llama/modeling_llama.py:389 / TF:models/gpt2/modeling_gpt2.py:577-585."""
    if spec.arch == "gpt2":
        pos = torch.arange(pos0, pos0 + ids.shape[1])
        return F.embedding(ids, sd["transformer.wte.weight"]) + F.embedding(pos, sd["transformer.wpe.weight"])[None]
    return F.embedding(ids, sd["model.embed_tokens.weight"])


def run_layers(h, sd, spec: ModelSpec, start: int, end: int, cache: KVCache, pos0: int) -> torch.Tensor:
    """Layers start..end inclusive on hidden states [B,T,d] — the intended shard
    unit of reference worker/app.py:332-372."""
    T = h.shape[1]
    if spec.arch == "llama":
        cos, sin = rope_cos_sin(spec, torch.arange(pos0, pos0 + T), h.dtype)
        for i in range(start, end + 1):
            h = llama_layer(h, sd, i, spec, cos, sin, cache, pos0)
    else:
        for i in range(start, end + 1):
            h = gpt2_layer(h, sd, i, spec, cache, pos0)
    return h


def head(h_last: torch.Tensor, sd, spec: ModelSpec) -> torch.Tensor:
    """Final norm + lm_head on [B,d] -> logits [B,V] in the model dtype
    (TF:models/llama/modeling_llama.py:421,:486-487; GPT-2 :628,:705-706)."""
    if spec.arch == "gpt2":
        x = F.layer_norm(h_last, (spec.d_model,), sd["transformer.ln_f.weight"], sd["transformer.ln_f.bias"], spec.eps)
        return F.linear(x, sd["transformer.wte.weight"])
    x = rmsnorm(h_last, sd["model.norm.weight"], spec.eps)
    w = sd["model.embed_tokens.weight"] if spec.tie_embeddings and "lm_head.weight" not in sd else sd["lm_head.weight"]
    return F.linear(x, w)


@torch.no_grad()
def forward(ids: torch.Tensor, sd, spec: ModelSpec, cache: Optional[KVCache] = None,
            shards: Optional[List[Tuple[int, int]]] = None, boundaries: Optional[list] = None) -> torch.Tensor:
    """Whole-model forward of new tokens `ids` [B,T] -> last-position logits [B,V].
    `shards` walks the layers stage by stage; hidden states leaving each stage are
    appended to `boundaries` (what the pipeline hands to the next GPU)."""
    cache = cache if cache is not None else KVCache()
    pos0 = cache.length()
    h = embed(ids, sd, spec, pos0)
    for (a, b) in shards or [(0, spec.n_layers - 1)]:
        h = run_layers(h, sd, spec, a, b, cache, pos0)
        if boundaries is not None:
            boundaries.append(h.clone())
    return head(h[:, -1], sd, spec)


@torch.no_grad()
def greedy_generate(ids: torch.Tensor, sd, spec: ModelSpec, max_length: int,
                    shards: Optional[List[Tuple[int, int]]] = None, return_logits: bool = False):
    """generate(do_sample=False, max_length=<total length>): prefill then one token
    per step, argmax over fp32 logits (TF:generation/utils.py:2736-2811).
    `max_length` counts prompt + new tokens like worker/app.py:259,:299."""
    cache = KVCache()
    out, all_logits = ids.clone(), []
    logits = forward(ids, sd, spec, cache, shards)
    while out.shape[1] < max_length:
        lf = logits.to(torch.float32)
        if return_logits:
            all_logits.append(lf.clone())
        nxt = lf.argmax(dim=-1, keepdim=True)
        out = torch.cat((out, nxt), dim=1)
        if out.shape[1] >= max_length:
            break
        logits = forward(nxt, sd, spec, cache, shards)
    return (out, all_logits) if return_logits else out


# --------------------------------------------------------------------------- seeded weights
def seeded_state_dict(spec: ModelSpec, seed: int, std: float = 0.02, dtype=torch.bfloat16) -> Dict[str, torch.Tensor]:
    """Random-init weights with HF parameter names, independent of HF's own init
    code so the same tensors can be rebuilt on any machine from (spec, seed, std):
    N(0, std) matrices/embeddings, norm weights 1 (+ N(0, std) so they are not
    trivially the identity), biases N(0, std)."""
    g = torch.Generator().manual_seed(seed)

    def n(*shape, mean=0.0):
        return (mean + std * torch.randn(*shape, generator=g, dtype=torch.float32)).to(dtype)

    sd = {}
    d, ff, V, L = spec.d_model, spec.d_ff, spec.vocab, spec.n_layers
    if spec.arch == "llama":
        sd["model.embed_tokens.weight"] = n(V, d)
        for i in range(L):
            p = f"model.layers.{i}."
            sd[p + "input_layernorm.weight"] = n(d, mean=1.0)
            sd[p + "self_attn.q_proj.weight"] = n(spec.n_heads * spec.head_dim, d)
            sd[p + "self_attn.k_proj.weight"] = n(spec.n_kv_heads * spec.head_dim, d)
            sd[p + "self_attn.v_proj.weight"] = n(spec.n_kv_heads * spec.head_dim, d)
            sd[p + "self_attn.o_proj.weight"] = n(d, spec.n_heads * spec.head_dim)
            sd[p + "post_attention_layernorm.weight"] = n(d, mean=1.0)
            sd[p + "mlp.gate_proj.weight"] = n(ff, d)
            sd[p + "mlp.up_proj.weight"] = n(ff, d)
            sd[p + "mlp.down_proj.weight"] = n(d, ff)
        sd["model.norm.weight"] = n(d, mean=1.0)
        if not spec.tie_embeddings:
            sd["lm_head.weight"] = n(V, d)
    else:
        sd["transformer.wte.weight"] = n(V, d)
        sd["transformer.wpe.weight"] = n(spec.max_pos, d)
        for i in range(L):
            p = f"transformer.h.{i}."
            sd[p + "ln_1.weight"], sd[p + "ln_1.bias"] = n(d, mean=1.0), n(d)
            sd[p + "attn.c_attn.weight"], sd[p + "attn.c_attn.bias"] = n(d, 3 * d), n(3 * d)
            sd[p + "attn.c_proj.weight"], sd[p + "attn.c_proj.bias"] = n(d, d), n(d)
            sd[p + "ln_2.weight"], sd[p + "ln_2.bias"] = n(d, mean=1.0), n(d)
            sd[p + "mlp.c_fc.weight"], sd[p + "mlp.c_fc.bias"] = n(d, ff), n(ff)
            sd[p + "mlp.c_proj.weight"], sd[p + "mlp.c_proj.bias"] = n(ff, d), n(d)
        sd["transformer.ln_f.weight"], sd["transformer.ln_f.bias"] = n(d, mean=1.0), n(d)
    return sd


def seeded_prompts(vocab: int, batch: int, length: int, seed: int) -> torch.Tensor:
    g = torch.Generator().manual_seed(seed)
    return torch.randint(0, vocab, (batch, length), generator=g, dtype=torch.int64)
