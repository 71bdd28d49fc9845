"""Parity at FULL DEPTH against the reference's own forward on the same B200.

The reference loads `AutoModelForCausalLM` (reference worker/app.py:121-124, `USE_GPU=1` :23-26) and calls
`model.generate(...)` (:297-305); with `do_sample=False` that is the parity mode north-star names ("greedy
token IDs must match exactly, logits within a stated bf16 tolerance").  The other engine tests pin the CUDA path on
2-4-layer models; here the WHOLE models of BASELINE.json run side by side, from one seeded state dict:

  * Llama-3-8B   (32 layers), 2 048-token prompt -> 64 greedy tokens  (configs[2], the headline)
  * TinyLlama-1.1B (22 layers), 512-token prompt -> 64 tokens         (configs[1])
  * Mistral-7B   (32 layers), 4 200-token prompt -> 24 tokens, so the 4 096-token window bites (configs[3])
  * GPT-2 small  (12 layers), 32-token prompt -> 64 tokens: configs[0]'s model and shape on the CUDA path

three ways: this engine (bf16, CUDA-graph decode = the path bench.py times), transformers bf16 sdpa (what the
reference would run on this GPU) and transformers fp32 ("truth").  Stated tolerance (the one tests/test_engine_gpu.py
uses on the small models):
    max|ours - fp32| <= 2 x max|HF-bf16 - fp32| + 4e-3     (teacher-forced along HF-bf16's tokens, every step)
    token == HF-fp32's token at every step whose fp32 top-1/top-2 margin exceeds twice the bf16 error on the
      leading candidates (margin-safe steps), and over all steps as often as HF-bf16's own token does (- 10 %)
    free-running tokens == HF-bf16's up to the first step that is not margin-safe
Random-init weights give nearly flat logits: at 32 layers and a 2k context the bf16 noise on a logit
(~0.1, up to 0.5 somewhere in a 128k-entry row) is of the order of the median top-1/top-2 margin (~0.2), so
transformers' own bf16 run picks fp32's token on only ~83 % of the steps (measured: 53 of 64, this engine 53 of 64)
and "exact" can only be asked where the margin clears the noise.  Margins, noise and the counts are printed (-s),
written to gpurun_out/full_depth_parity.jsonl and asserted to be non-trivial.
"""
import gc
import json
import os

import pytest
import torch

pytestmark = pytest.mark.gpu
BF = torch.bfloat16
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _hf_config(transformers, name):
    from dli_b200.engine import NAMED_SPECS

    s = NAMED_SPECS[name]
    if s.arch == "gpt2":
        return s, transformers.GPT2Config(vocab_size=s.vocab, n_embd=s.d_model, n_layer=s.n_layers, n_head=s.n_heads,
                                          n_positions=s.max_pos, n_inner=s.d_ff, layer_norm_epsilon=s.eps)
    kw = dict(vocab_size=s.vocab, hidden_size=s.d_model, intermediate_size=s.d_ff, num_hidden_layers=s.n_layers,
              num_attention_heads=s.n_heads, num_key_value_heads=s.n_kv_heads, head_dim=s.head_dim,
              max_position_embeddings=s.max_pos, rms_norm_eps=s.eps, tie_word_embeddings=False,
              rope_parameters={"rope_type": "default", "rope_theta": s.rope_theta})
    if s.sliding_window:
        return s, transformers.MistralConfig(sliding_window=s.sliding_window, **kw)
    return s, transformers.LlamaConfig(**kw)


def _seeded_state_dict(spec, device, seed):
    """HF-named bf16 tensors, N(0, 0.02) like transformers' default init, norms = 1, generated on the GPU."""
    g = torch.Generator(device=device).manual_seed(seed)

    def n(*shape):
        return (0.02 * torch.randn(*shape, device=device, generator=g)).to(BF)

    d, hd = spec.d_model, spec.head_dim
    sd = {"model.embed_tokens.weight": n(spec.vocab, d), "model.norm.weight": torch.ones(d, device=device, dtype=BF),
          "lm_head.weight": n(spec.vocab, d)}
    for i in range(spec.n_layers):
        p = f"model.layers.{i}."
        sd[p + "input_layernorm.weight"] = torch.ones(d, device=device, dtype=BF)
        sd[p + "post_attention_layernorm.weight"] = torch.ones(d, device=device, dtype=BF)
        sd[p + "self_attn.q_proj.weight"] = n(spec.n_heads * hd, d)
        sd[p + "self_attn.k_proj.weight"] = n(spec.n_kv_heads * hd, d)
        sd[p + "self_attn.v_proj.weight"] = n(spec.n_kv_heads * hd, d)
        sd[p + "self_attn.o_proj.weight"] = n(d, spec.n_heads * hd)
        sd[p + "mlp.gate_proj.weight"] = n(spec.d_ff, d)
        sd[p + "mlp.up_proj.weight"] = n(spec.d_ff, d)
        sd[p + "mlp.down_proj.weight"] = n(d, spec.d_ff)
    return sd


def _gpt2_state_dict(transformers, cfg, device, seed):
    """GPT-2 keeps transformers' own (seeded) initialisation — Conv1D layout, learned positions, tied head — rounded
    to bf16 so that all three runs hold the same weights."""
    torch.manual_seed(seed)
    with torch.device(device):
        m = transformers.AutoModelForCausalLM.from_config(cfg, dtype=torch.float32)
    return {k: v.detach().to(BF) for k, v in m.state_dict().items()}


def _hf_model(transformers, cfg, sd, dtype, device):
    with torch.device(device):  # parameters and the rotary buffers are created on the GPU
        m = transformers.AutoModelForCausalLM.from_config(cfg, dtype=dtype)
    m.load_state_dict(sd, strict=True)  # copies (and widens for fp32)
    return m.eval()


@pytest.mark.parametrize("name,T,n_new", [("llama-3-8b", 2048, 64), ("tinyllama-1.1b", 512, 64),
                                          ("mistral-7b", 4200, 24), ("gpt2", 32, 64)])  # gpt2: configs[0]'s shape
def test_full_depth_model_matches_transformers_on_the_same_gpu(cuda, name, T, n_new):
    transformers = pytest.importorskip("transformers")
    from dli_b200.engine import ShardEngine

    spec, cfg = _hf_config(transformers, name)
    free = torch.cuda.mem_get_info()[0]
    w_bf16 = spec.n_layers * spec.layer_weight_bytes() + 2 * spec.head_weight_bytes()
    need = int(3.3 * w_bf16) + (8 << 30)  # peak: the bf16 state dict + the fp32 transformers model
    if free < need:
        pytest.skip(f"needs ~{need >> 30} GiB of free HBM, {free >> 30} GiB available")
    sd = _gpt2_state_dict(transformers, cfg, cuda, 2024) if spec.arch == "gpt2" else _seeded_state_dict(spec, cuda, seed=2024)
    ids = torch.randint(0, spec.vocab, (1, T), generator=torch.Generator().manual_seed(1))
    max_len = T + n_new

    # ---- the reference's path on this GPU: transformers generate, bf16 sdpa, greedy -----------------------------
    with torch.no_grad():
        hf = _hf_model(transformers, cfg, sd, BF, cuda)
        gen = hf.generate(ids.to(cuda), attention_mask=torch.ones_like(ids).to(cuda), max_length=max_len,
                          do_sample=False, pad_token_id=0, output_logits=True, return_dict_in_generate=True)
        hf_tokens = gen.sequences[0, T:].cpu()                                    # [n_new]
        hf_logits = torch.stack([l[0].float().cpu() for l in gen.logits])         # [n_new, V]
        del hf, gen
        gc.collect()
        torch.cuda.empty_cache()
        # ---- truth: fp32 forward over prompt + HF's tokens, logits of the n_new predicting positions -----------
        hf32 = _hf_model(transformers, cfg, sd, torch.float32, cuda)
        full = torch.cat((ids[0], hf_tokens[:-1]))[None].to(cuda)
        truth = hf32(full, logits_to_keep=n_new).logits[0].float().cpu()          # [n_new, V]
        del hf32
        gc.collect()
        torch.cuda.empty_cache()

    # ---- this engine: teacher-forced along HF's tokens (eager steps), then free-running (CUDA-graph steps) ------
    eng = ShardEngine(spec, device=str(cuda), max_batch=1, max_seq_len=max_len + 8)
    eng.load_state_dict(sd)
    del sd
    forced = hf_tokens[None, :-1]
    ours_tf, logits = eng.generate(ids, max_len, return_logits=True, forced=forced)
    ours_logits = torch.stack([l[0] for l in logits])                             # [n_new, V] fp32
    ours_free = eng.generate(ids, max_len)[0, T:]
    eng.unload()

    # ---- noise and margins ------------------------------------------------------------------------------------------
    hf_err = (hf_logits - truth).abs().amax(-1)          # transformers' own bf16 noise per step (max over the vocabulary)
    our_err = (ours_logits - truth).abs().amax(-1)
    top = truth.topk(17, dim=-1)
    margin = top.values[:, 0] - top.values[:, 1]
    # A step is "margin-safe" when no bf16 implementation can pick another token than fp32 does: the error either
    # implementation makes on the 16 best fp32 candidates is below half the top-1/top-2 margin, and nothing outside
    # those 16 comes within reach (gap to the 17th > the two worst errors anywhere in the row).
    idx = top.indices[:, :16]
    local = torch.maximum((ours_logits.gather(1, idx) - top.values[:, :16]).abs().amax(-1),
                          (hf_logits.gather(1, idx) - top.values[:, :16]).abs().amax(-1))
    safe = (margin > 2.0 * local) & ((top.values[:, 0] - top.values[:, 16]) > 2.0 * torch.maximum(hf_err, our_err))
    scale = truth.abs().max().item()
    truth_tok = truth.argmax(-1)
    ours_vs_truth = ours_logits.argmax(-1) == truth_tok
    hf_vs_truth = hf_logits.argmax(-1) == truth_tok
    agree_hf = ours_tf[0, T:] == hf_tokens
    first_unsafe = int((~safe).float().argmax().item()) if bool((~safe).any()) else n_new
    same_hist = int((ours_free == hf_tokens).long().cumprod(0).sum())
    report = {"model": name, "prompt": T, "new": n_new, "logit_scale": scale, "max_err_ours_vs_fp32": our_err.max().item(),
              "max_err_hf_bf16_vs_fp32": hf_err.max().item(), "median_err_ours": our_err.median().item(),
              "median_err_hf_bf16": hf_err.median().item(), "fp32_margin_min": margin.min().item(),
              "fp32_margin_median": margin.median().item(), "margin_safe_steps": int(safe.sum()),
              "teacher_forced_equal_hf_bf16": int(agree_hf.sum()), "teacher_forced_equal_fp32": int(ours_vs_truth.sum()),
              "hf_bf16_equal_fp32": int(hf_vs_truth.sum()), "free_running_equal_hf_bf16_prefix": same_hist,
              "margin_safe_prefix": first_unsafe}
    print("\n[full-depth parity] " + json.dumps(report))
    try:  # kept with the run's other outputs when the box has the directory (gpurun merges it back)
        os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
        with open(os.path.join(ROOT, "gpurun_out", "full_depth_parity.jsonl"), "a") as f:
            f.write(json.dumps(report) + "\n")
    except OSError:
        pass
    # logits: within twice transformers' own bf16 error against fp32 truth, worst step and typical step
    assert our_err.max() <= 2.0 * hf_err.max() + 4e-3
    assert our_err.median() <= 2.0 * hf_err.median() + 4e-3
    # tokens: exact wherever the margin clears the noise ...
    assert bool(ours_vs_truth[safe].all()) and bool(agree_hf[safe].all())
    # ... and overall this engine picks fp32's token about as often as transformers' own bf16 run does (on
    # random-init weights neither can be exact: the median margin is of the order of the bf16 noise)
    assert int(ours_vs_truth.sum()) >= int(hf_vs_truth.sum()) - max(3, n_new // 10)
    # free-running (the CUDA-graph path): identical to HF-bf16 over the margin-safe prefix
    assert torch.equal(ours_free[:first_unsafe], hf_tokens[:first_unsafe])
    # the graph path and the eager path of this engine agree step by step while they are fed the same history
    k = min(same_hist + 1, n_new)
    assert torch.equal(ours_free[:k], ours_tf[0, T: T + k])
    # the statement must not be vacuous
    assert int(safe.sum()) >= max(3, n_new // 10), "too few margin-safe steps to call this a parity check"
