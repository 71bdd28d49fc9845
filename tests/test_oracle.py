"""Pin the oracle (oracle/decoder_ref.py) to the reference's arithmetic:
 (a) the committed fixtures produced by transformers' forward/generate and by the
     reference's own sharder (tests/golden, oracle/gen_golden.py);
 (b) transformers itself, imported here, on freshly seeded weights.
Tolerances: fp32 runs agree to 2e-4 absolute on logits of magnitude ~4 (different
summation order only); bf16 runs are compared with the error transformers' own
bf16 run has against fp32 (stored per step in the fixture)."""
import json
import os

import pytest
import torch

from golden_util import GOLD, MODEL_CASES, load_case, spec_kwargs
from oracle import decoder_ref as R


def _spec(case):
    from dli_b200.engine import ModelSpec as ES

    e = ES.from_hf_config(spec_kwargs(case["config"]))
    return R.ModelSpec(**e.__dict__)


@pytest.mark.parametrize("name", MODEL_CASES)
def test_oracle_fp32_matches_transformers_fixture(name):
    c = load_case(name)
    spec = _spec(c)
    sd = {k: v.float() for k, v in R.seeded_state_dict(spec, c["seed"], c["std"]).items()}
    full = torch.cat((c["prompts"], c["tokens"]), dim=1)
    T = c["prompts"].shape[1]
    for step in (0, c["tokens"].shape[1] // 2, c["tokens"].shape[1] - 1):
        lg = R.forward(full[:, : T + step], sd, spec)
        ref = c["logits_fp32"][step]
        # fixture stores fp32 logits as fp16 (2^-11 relative): 4e-3 absolute at |logit| < 8
        torch.testing.assert_close(lg, ref, rtol=0, atol=4e-3)


@pytest.mark.parametrize("name", MODEL_CASES)
def test_oracle_bf16_within_transformers_own_bf16_error(name):
    c = load_case(name)
    spec = _spec(c)
    sd = R.seeded_state_dict(spec, c["seed"], c["std"])
    n_new = c["tokens"].shape[1]
    # teacher-forced along transformers' bf16 greedy path, through the oracle's KV cache
    cache, logits = R.KVCache(), []
    cur = c["prompts"]
    for i in range(n_new):
        logits.append(R.forward(cur, sd, spec, cache).float())
        cur = c["tokens"][:, i:i + 1]
    lg = torch.stack(logits)
    err_vs_fp32 = (lg - c["logits_fp32"]).abs().amax(-1)
    # the oracle is a bf16 forward: its distance to fp32 truth stays within 2x the worst
    # distance transformers' own bf16 forward has on this case (+ the fp16 storage error)
    assert err_vs_fp32.max() <= 2.0 * c["hf_err"].max() + 4e-3
    # and wherever the fp32 margin clears that noise the greedy token is the same
    safe = c["margin"] > 2.0 * (c["hf_err"].max() + err_vs_fp32.max())
    same = lg.argmax(-1) == c["logits_fp32"].argmax(-1)
    assert safe.any() and bool(same[safe].all())


@pytest.mark.parametrize("name", ["llama_tiny", "gpt2_tiny"])
def test_oracle_shard_boundary_and_staged_walk(name):
    c = load_case(name)
    spec = _spec(c)
    sd = R.seeded_state_dict(spec, c["seed"], c["std"])
    b1, b2 = [], []
    mono = R.forward(c["prompts"], sd, spec, boundaries=b1)
    staged = R.forward(c["prompts"], sd, spec, shards=R.shard_ranges(spec.n_layers, 2), boundaries=b2)
    assert torch.equal(mono, staged)  # staged layer-range walk is bit-identical to monolithic
    assert c["boundary_layer"] == R.shard_ranges(spec.n_layers, 2)[1][0]
    # hidden state leaving shard 0 vs transformers' hidden_states[boundary]: bf16 residual
    # stream of magnitude ~1-4; 0.1 absolute covers the per-backend attention rounding
    torch.testing.assert_close(b2[0].float(), c["boundary_hidden"], rtol=0, atol=0.1)


def test_shard_formula_matches_reference_sharder():
    for case in json.load(open(os.path.join(GOLD, "shard_ranges.json"))):
        got = R.shard_ranges(case["n_layers"], case["n_shards"])
        want = [(m["start_layer"], m["end_layer"]) for m in case["metadata"]]
        assert got == want, case
        for i, m in enumerate(case["metadata"]):
            assert m["shard_id"] == i and m["num_shards"] == case["n_shards"] and m["total_layers"] == case["n_layers"]


def test_oracle_matches_transformers_live():
    """Same check against transformers imported in this process (fp32, fresh seed)."""
    transformers = pytest.importorskip("transformers")
    cfg = transformers.LlamaConfig(vocab_size=256, hidden_size=128, intermediate_size=256, num_hidden_layers=3,
                                   num_attention_heads=2, num_key_value_heads=1, max_position_embeddings=64,
                                   rms_norm_eps=1e-5, rope_theta=10000.0, tie_word_embeddings=False)
    spec = R.ModelSpec.from_hf(cfg)
    sd = R.seeded_state_dict(spec, seed=99, std=0.08, dtype=torch.float32)
    m = transformers.AutoModelForCausalLM.from_config(cfg, dtype=torch.float32).eval()
    m.load_state_dict(sd, strict=True)
    ids = R.seeded_prompts(spec.vocab, 2, 11, seed=3)
    with torch.no_grad():
        ref = m(ids).logits[:, -1]
        g_ref = m.generate(ids, attention_mask=torch.ones_like(ids), max_length=24, do_sample=False, pad_token_id=0)
    torch.testing.assert_close(R.forward(ids, sd, spec), ref, rtol=0, atol=2e-4)
    assert torch.equal(R.greedy_generate(ids, sd, spec, 24, shards=R.shard_ranges(3, 3)), g_ref)


# ------------------------------------------------------------------------------------- RoPE scaling
_LLAMA31_ROPE = {"rope_type": "llama3", "factor": 8.0, "low_freq_factor": 1.0, "high_freq_factor": 4.0,
                 "original_max_position_embeddings": 64, "rope_theta": 10000.0}


def _scaled_cfg(transformers, rope):
    return transformers.LlamaConfig(vocab_size=256, hidden_size=128, intermediate_size=256, num_hidden_layers=2,
                                    num_attention_heads=2, num_key_value_heads=1, max_position_embeddings=512,
                                    rms_norm_eps=1e-5, tie_word_embeddings=False, rope_parameters=dict(rope))


@pytest.mark.parametrize("rope", [_LLAMA31_ROPE, {"rope_type": "linear", "factor": 4.0, "rope_theta": 10000.0}])
def test_rope_scaling_matches_transformers(rope):
    """`llama3` / `linear` rope scaling (SURVEY.md a9 "only if config says so"): the engine's host-side
    inverse frequencies equal transformers' bit for bit, from the new-style config and from a
    Llama-3.1-style config.json dict; the oracle's forward with the scaled table equals transformers'."""
    transformers = pytest.importorskip("transformers")
    from transformers.modeling_rope_utils import ROPE_INIT_FUNCTIONS

    from dli_b200.engine import ModelSpec as ES, rope_inv_freq

    cfg = _scaled_cfg(transformers, rope)
    want, factor = ROPE_INIT_FUNCTIONS[rope["rope_type"]](cfg, "cpu")
    assert factor == 1.0
    assert torch.equal(rope_inv_freq(ES.from_hf_config(cfg)), want)
    legacy = {"model_type": "llama", "hidden_size": 128, "num_attention_heads": 2, "num_key_value_heads": 1,
              "num_hidden_layers": 2, "intermediate_size": 256, "vocab_size": 256, "rms_norm_eps": 1e-5,
              "max_position_embeddings": 512, "rope_theta": 10000.0,
              "rope_scaling": {k: v for k, v in rope.items() if k != "rope_theta"}}
    assert torch.equal(rope_inv_freq(ES.from_hf_config(legacy)), want)

    spec = R.ModelSpec.from_hf(cfg)
    sd = R.seeded_state_dict(spec, seed=5, std=0.08, dtype=torch.float32)
    m = transformers.AutoModelForCausalLM.from_config(cfg, dtype=torch.float32).eval()
    m.load_state_dict(sd, strict=True)
    ids = R.seeded_prompts(spec.vocab, 2, 150, seed=4)  # long enough that scaled and plain tables differ visibly
    with torch.no_grad():
        ref = m(ids).logits[:, -1]
    torch.testing.assert_close(R.forward(ids, sd, spec), ref, rtol=0, atol=2e-4)
    plain = R.ModelSpec(**{**spec.__dict__, "rope_scaling": None})
    assert (R.forward(ids, sd, plain) - ref).abs().max() > 1e-2  # the scaling is not a no-op on this case


def test_unknown_rope_scaling_is_refused():
    """A checkpoint whose RoPE variant is not implemented must not load as plain RoPE (it would
    generate wrong tokens silently)."""
    from dli_b200.engine import ModelSpec as ES

    base = {"model_type": "llama", "hidden_size": 128, "num_attention_heads": 2, "num_hidden_layers": 2,
            "intermediate_size": 256, "vocab_size": 256}
    for bad in ({"rope_type": "yarn", "factor": 4.0}, {"type": "dynamic", "factor": 2.0},
                {"rope_type": "longrope", "short_factor": [1.0], "long_factor": [2.0]}):
        with pytest.raises(ValueError, match="rope_type"):
            ES.from_hf_config({**base, "rope_scaling": bad})
    with pytest.raises(ValueError, match="llama3 needs"):
        ES.from_hf_config({**base, "rope_scaling": {"rope_type": "llama3", "factor": 8.0}})
