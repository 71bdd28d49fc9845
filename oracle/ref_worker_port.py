"""ORACLE / CPU BASELINE — test and benchmark infrastructure only.

Port of the reference worker's CPU inference path (USE_GPU=0), i.e. the body of
`run_inference` / `run_shard_inference` in reference worker/app.py:289-308 and
:355-370: tokenize -> `model.generate(**inputs, max_length=..., do_sample=True,
top_p=0.95, top_k=50, temperature=0.8, num_return_sequences=1)` -> decode, on a
`transformers` causal LM living on the CPU.  The reference file itself cannot
travel to the GPU box (/root/reference exists only in the build container), and
its arithmetic is the un-vendored `transformers` package, which IS installed on
the box — so this port calls the same third-party entry points with the same
arguments.  oracle/gen_golden.py pins the port's HTTP-visible behaviour against
the real worker/app.py (tests/golden/worker_contract.json).

Used by bench.py (`cpu_baseline` leg and `--impl reference`) and by tests only.
"""
from __future__ import annotations

import time
from typing import Dict, Optional

import torch

# generate() arguments exactly as the reference passes them (worker/app.py:297-305, :359-367)
REFERENCE_GENERATE_KWARGS = dict(num_return_sequences=1, do_sample=True, top_p=0.95, top_k=50, temperature=0.8)


class IdTokenizer:
    """Stand-in for AutoTokenizer (no tokenizer files offline): the prompt is a
    whitespace-separated list of token ids; decode writes them back the same way."""

    def __call__(self, prompt: str, return_tensors: str = "pt") -> Dict[str, torch.Tensor]:
        ids = torch.tensor([[int(t) for t in prompt.split()]], dtype=torch.int64)
        return {"input_ids": ids, "attention_mask": torch.ones_like(ids)}

    def decode(self, ids, skip_special_tokens: bool = True) -> str:
        return " ".join(str(int(i)) for i in ids)


def hf_config_for(spec):
    """transformers config of a ModelSpec (engine.ModelSpec or oracle ModelSpec)."""
    import transformers

    if spec.arch == "gpt2":
        return transformers.GPT2Config(vocab_size=spec.vocab, n_embd=spec.d_model, n_layer=spec.n_layers,
                                       n_head=spec.n_heads, n_positions=spec.max_pos, n_inner=spec.d_ff,
                                       layer_norm_epsilon=spec.eps)
    kw = dict(vocab_size=spec.vocab, hidden_size=spec.d_model, intermediate_size=spec.d_ff,
              num_hidden_layers=spec.n_layers, num_attention_heads=spec.n_heads,
              num_key_value_heads=spec.n_kv_heads, head_dim=spec.head_dim, max_position_embeddings=spec.max_pos,
              rms_norm_eps=spec.eps, rope_theta=spec.rope_theta, tie_word_embeddings=spec.tie_embeddings)
    if spec.sliding_window:
        return transformers.MistralConfig(sliding_window=spec.sliding_window, **kw)
    return transformers.LlamaConfig(**kw)


def build_cpu_model(spec, dtype=torch.bfloat16, seed: int = 0, std: float = 0.02):
    """Random-init transformers model on the CPU without paying HF's per-tensor
    init: parameters are materialised empty and filled by tiling one seeded
    N(0, std) block (norm weights = 1), which is enough for timing."""
    import transformers

    cfg = hf_config_for(spec)
    with torch.device("meta"):
        m = transformers.AutoModelForCausalLM.from_config(cfg, dtype=dtype)
    m = m.to_empty(device="cpu")
    g = torch.Generator().manual_seed(seed)
    block = (std * torch.randn(1 << 22, generator=g)).to(dtype)
    with torch.no_grad():
        for name, p in m.named_parameters():
            flat = p.data.view(-1)
            if p.dim() == 1 and ("norm" in name or "ln_" in name) and "bias" not in name:
                flat.fill_(1.0)
                continue
            k = flat.numel() // block.numel()
            if k:
                flat[: k * block.numel()].view(k, -1).copy_(block)  # broadcast copy, memcpy speed
            flat[k * block.numel():] = block[: flat.numel() - k * block.numel()]
        for name, b in m.named_buffers():  # rotary inv_freq etc. were on meta
            if "inv_freq" in name:
                hd = spec.head_dim
                b.data = 1.0 / (spec.rope_theta ** (torch.arange(0, hd, 2, dtype=torch.int64).float() / hd))
    if hasattr(m, "tie_weights"):
        m.tie_weights()
    return m.eval()


def run_inference_port(model, tokenizer, prompt: str, max_length: int = 100, timeout: float = 60.0,
                       greedy: bool = False) -> Dict[str, object]:
    """The reference's unsharded /inference body (worker/app.py:288-318) on DEVICE='cpu'.
    `greedy=True` is the parity mode (do_sample=False); default is the reference's sampling."""
    start = time.time()
    inputs = tokenizer(prompt, return_tensors="pt")
    inputs = {k: v.to("cpu") for k, v in inputs.items()}
    if (time.time() - start) > timeout:
        raise TimeoutError("Inference preparation timed out")
    kw = dict(REFERENCE_GENERATE_KWARGS)
    if greedy:
        kw = dict(num_return_sequences=1, do_sample=False)
    with torch.no_grad():
        outputs = model.generate(**inputs, max_length=max_length, pad_token_id=0, **kw)
    result = tokenizer.decode(outputs[0], skip_special_tokens=True)
    if (time.time() - start) > timeout:
        raise TimeoutError("Inference generation timed out")
    return {"status": "success", "result": result, "execution_time": time.time() - start}


def time_cpu_decode(spec, prompt_len: int, new_tokens: int, dtype=torch.bfloat16, greedy: bool = True,
                    model=None, warmup_tokens: int = 1) -> Dict[str, object]:
    """Decode tokens/s of the port on this host's cores for a bounded sample:
    one prompt of `prompt_len` synthetic ids, `new_tokens` generated.  Decode rate is
    (new_tokens - 1) / (t_total - t_prefill) with t_prefill measured by a
    max_length = prompt_len + 1 call (the reference exposes only whole-request time)."""
    model = model if model is not None else build_cpu_model(spec, dtype)
    tok = IdTokenizer()
    g = torch.Generator().manual_seed(1)
    prompt = " ".join(str(int(i)) for i in torch.randint(0, spec.vocab, (prompt_len,), generator=g))
    for _ in range(max(0, warmup_tokens)):
        run_inference_port(model, tok, prompt, prompt_len + 1, timeout=1e9, greedy=greedy)
    t0 = time.time()
    run_inference_port(model, tok, prompt, prompt_len + 1, timeout=1e9, greedy=greedy)
    t_prefill = time.time() - t0
    t0 = time.time()
    run_inference_port(model, tok, prompt, prompt_len + new_tokens, timeout=1e9, greedy=greedy)
    t_total = time.time() - t0
    dec = max(t_total - t_prefill, 1e-9)
    return {
        "decode_tok_s": (new_tokens - 1) / dec,
        "prefill_tok_s": prompt_len / max(t_prefill, 1e-9),
        "t_total_s": t_total,
        "t_prefill_s": t_prefill,
        "threads": torch.get_num_threads(),
        "sample": f"1 prompt x {prompt_len} tokens -> {new_tokens} new tokens, "
                  f"{'greedy' if greedy else 'reference sampling'}, {str(dtype).replace('torch.', '')} on CPU",
    }
