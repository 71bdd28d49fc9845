"""End-to-end parity of the CUDA path (ShardEngine over the C-ABI) against
 (a) the committed transformers fixtures (tests/golden) and
 (b) the oracle run in this process on the same seeded weights.

Tolerance (bf16, stated by north_star as "a stated bf16 tolerance"): our
last-position logits must lie within 2x of the distance that transformers' OWN
bf16 forward has from the fp32 forward on the same case (worst step), + 4e-3 for
the fixture's fp16 storage.  Greedy tokens must equal transformers' wherever the
fp32 top-1/top-2 margin clears that noise (teacher-forced), and the free-running
sequences must be identical on the cases whose whole path clears it.
"""
import pytest
import torch

from golden_util import MODEL_CASES, load_case, spec_kwargs

pytestmark = pytest.mark.gpu


def _engine(case, **kw):
    from dli_b200.engine import ModelSpec, ShardEngine
    from oracle import decoder_ref as R

    spec = ModelSpec.from_hf_config(spec_kwargs(case["config"]))
    sd = R.seeded_state_dict(R.ModelSpec(**spec.__dict__), case["seed"], case["std"])
    eng = ShardEngine(spec, max_batch=4, max_seq_len=128, **kw)
    eng.load_state_dict(sd)
    return eng, spec, sd


@pytest.mark.parametrize("name", MODEL_CASES)
def test_logits_and_tokens_teacher_forced(cuda, name):
    c = load_case(name)
    eng, spec, _ = _engine(c)
    T, N = c["prompts"].shape[1], c["tokens"].shape[1]
    out, logits = eng.generate(c["prompts"], T + N, return_logits=True, forced=c["tokens"])
    lg = torch.stack(logits)  # [N, B, V]
    err32 = (lg - c["logits_fp32"]).abs().amax(-1)
    err16 = (lg - c["logits_bf16"]).abs().amax(-1)
    scale = c["logits_fp32"].abs().max().item()
    print(f"\n{name}: max|ours-fp32|={err32.max():.4f} max|ours-hf_bf16|={err16.max():.4f} "
          f"(hf_bf16 vs fp32 {c['hf_err'].max():.4f}; logit scale {scale:.2f}; "
          f"rel {err32.max().item() / scale:.4f})")
    assert err32.max() <= 2.0 * c["hf_err"].max() + 4e-3
    noise = c["hf_err"].max() + err32.max()
    safe = c["margin"] > 2.0 * noise
    ours = out[:, T:].t()  # [N, B] this engine's own argmax at each step
    assert safe.float().mean() > 0.15
    assert bool((ours == c["tokens"].t())[safe].all())
    # overall agreement with transformers' bf16 greedy tokens, reported and bounded
    agree = (ours == c["tokens"].t()).float().mean().item()
    print(f"{name}: teacher-forced token agreement {agree:.3f} (safe steps {safe.float().mean():.2f})")
    assert agree >= 0.9


@pytest.mark.parametrize("name", MODEL_CASES)
def test_free_running_greedy_graph_equals_eager(cuda, name):
    """CUDA-graph replayed decode == un-graphed decode, token for token."""
    c = load_case(name)
    eng, spec, _ = _engine(c)
    T, N = c["prompts"].shape[1], c["tokens"].shape[1]
    a = eng.generate(c["prompts"], T + N)
    b, _ = eng.generate(c["prompts"], T + N, return_logits=True)
    a2 = eng.generate(c["prompts"], T + N)  # replay of the already captured graph
    assert torch.equal(a, b) and torch.equal(a, a2)
    assert torch.equal(a[:, :T], c["prompts"])


@pytest.mark.parametrize("name", MODEL_CASES)
@pytest.mark.parametrize("chunk", [3, 5])
def test_chunked_prefill_matches_whole_prompt_prefill(cuda, name, chunk):
    """Feeding the prompt a few tokens at a time (attention over the paged cache) gives the same first-token
    logits as one prefill within bf16 noise, writes the same KV cache, and decoding continues identically."""
    c = load_case(name)
    eng, spec, _ = _engine(c)
    prompts = c["prompts"]
    B, T = prompts.shape
    N = 6
    ids = prompts.to(torch.int32).reshape(-1).cuda()
    eng.start_batch(B, T, T + N)
    eng.prefill(B, T, ids=ids)
    lg_a = eng._state(B)["logits"].float().clone()
    toks_a = [eng._state(B)["tokens"].clone()]
    for _ in range(N - 1):
        toks_a.append(eng.decode_step(B).clone())
    eng.start_batch(B, T, T + N)
    eng.prefill_chunked(B, T, chunk, ids=ids)
    lg_b = eng._state(B)["logits"].float().clone()
    toks_b = [eng._state(B)["tokens"].clone()]
    for _ in range(N - 1):
        toks_b.append(eng.decode_step(B).clone())
    scale = lg_a.abs().max().item()
    err = (lg_a - lg_b).abs().max().item()
    print(f"\n{name} chunk={chunk}: max|logits whole - chunked| = {err:.4f} (scale {scale:.2f})")
    # two bf16 attention kernels with different tile orders: same bound as the parity bar uses for HF's own noise
    assert err <= 2.0 * c["hf_err"].max().item() + 4e-3
    agree = (torch.stack(toks_a) == torch.stack(toks_b)).float().mean().item()
    assert agree >= 0.8, agree


def test_generate_with_prefill_chunk_setting(cuda):
    """engine.prefill_chunk: generate() feeds long prompts in chunks; tokens stay those of the whole-prompt
    prefill on the head_dim-128 case (same tcgen05 attention arithmetic either way)."""
    c = load_case("llama_hd128")
    eng, spec, _ = _engine(c)
    T = c["prompts"].shape[1]
    whole = eng.generate(c["prompts"], T + 8)
    eng.prefill_chunk = 4
    assert torch.equal(eng.generate(c["prompts"], T + 8), whole)


def test_free_running_exact_match_where_margin_allows(cuda):
    """Greedy token IDs identical to transformers' on the sequences whose every step
    has an fp32 margin above the bf16 noise floor of the case."""
    hits = 0
    for name in MODEL_CASES:
        c = load_case(name)
        eng, spec, _ = _engine(c)
        T, N = c["prompts"].shape[1], c["tokens"].shape[1]
        out = eng.generate(c["prompts"], T + N)[:, T:]
        noise = 2.0 * c["hf_err"].max()
        for b in range(out.shape[0]):
            # longest prefix of steps that are all safe
            safe = (c["margin"][:, b] > 2.0 * noise).int()
            n_safe = int(safe.cumprod(0).sum())
            if n_safe:
                hits += n_safe
                assert torch.equal(out[b, :n_safe], c["tokens"][b, :n_safe]), (name, b, n_safe)
    assert hits >= 10


@pytest.mark.parametrize("name", ["llama_tiny", "gpt2_tiny"])
def test_engine_matches_oracle_live(cuda, name):
    """Same seeded weights through the oracle in this process: prefill logits and
    2-shard boundary hidden states."""
    from dli_b200.engine import ShardEngine
    from oracle import decoder_ref as R

    c = load_case(name)
    eng, spec, sd = _engine(c)
    rspec = R.ModelSpec(**spec.__dict__)
    ids = c["prompts"]
    B, T = ids.shape
    bounds = []
    ref = R.forward(ids, sd, rspec, shards=R.shard_ranges(spec.n_layers, 2), boundaries=bounds).float()
    _, logits = eng.generate(ids, T + 1, return_logits=True)
    assert (logits[0] - ref).abs().max() <= 2.0 * c["hf_err"].max() + 4e-3
    # two ShardEngines wired by hand: hidden states leaving shard 0 and final tokens
    (a0, a1), (b0, b1) = R.shard_ranges(spec.n_layers, 2)
    s0 = ShardEngine(spec, a0, a1, max_batch=4, max_seq_len=128)
    s1 = ShardEngine(spec, b0, b1, max_batch=4, max_seq_len=128)
    s0.load_state_dict(sd), s1.load_state_dict(sd)
    for s in (s0, s1):
        s.start_batch(B, T, T + 4)
    dev_ids = ids.to("cuda", torch.int32).reshape(-1)
    h = s0.prefill(B, T, ids=dev_ids)
    torch.testing.assert_close(h.float().cpu().reshape(B, T, -1), bounds[0].float(), rtol=0, atol=0.1)
    assert torch.equal(h.float().cpu().reshape(B, T, -1), c["boundary_hidden"]) or \
        (h.float().cpu().reshape(B, T, -1) - c["boundary_hidden"]).abs().max() <= 0.1
    tok = s1.prefill(B, T, hidden=h).clone()
    full = eng.generate(ids, T + 3)
    assert torch.equal(tok.cpu().long(), full[:, T])
    # two decode steps through the 2-stage chain == single engine
    for i in range(2):
        s0._state(B)["tokens"].copy_(tok)
        h = s0.decode_step(B)
        tok = s1.decode_step(B, hidden_in=h).clone()
        assert torch.equal(tok.cpu().long(), full[:, T + 1 + i])


def test_batch_rows_are_independent(cuda):
    """Each sequence of a batch gets the tokens it gets alone (data-parallel rows)."""
    c = load_case("llama_tiny")
    eng, spec, _ = _engine(c)
    T = c["prompts"].shape[1]
    both = eng.generate(c["prompts"], T + 8)
    for b in range(c["prompts"].shape[0]):
        solo = eng.generate(c["prompts"][b:b + 1], T + 8)
        assert torch.equal(solo[0], both[b])


def test_unload_frees_memory(cuda):
    c = load_case("llama_tiny")
    eng, _, _ = _engine(c)
    eng.generate(c["prompts"], c["prompts"].shape[1] + 2)
    before = torch.cuda.memory_allocated()
    eng.unload()
    assert torch.cuda.memory_allocated() < before
    assert len(eng.kv.free) == eng.kv.num_pages


def test_gpt2_small_shape_against_oracle(cuda):
    """BASELINE configs[0] shape (GPT-2 small: L12 d768 H12 V50257, tied head, learned positions,
    LayerNorm, Conv1D biases, gelu_new) through the CUDA path vs the oracle on the same seeded
    weights: last-position logits within 2x the oracle's own bf16-vs-fp32 error."""
    from dli_b200.engine import NAMED_SPECS, ShardEngine
    from oracle import decoder_ref as R

    spec = NAMED_SPECS["gpt2"]
    rspec = R.ModelSpec(**spec.__dict__)
    sd = R.seeded_state_dict(rspec, seed=5, std=0.04)
    ids = R.seeded_prompts(spec.vocab, 1, 32, seed=6)
    ref16 = R.forward(ids, sd, rspec).float()
    ref32 = R.forward(ids, {k: v.float() for k, v in sd.items()}, rspec)
    eng = ShardEngine(spec, max_batch=1, max_seq_len=128)
    eng.load_state_dict(sd)
    out, logits = eng.generate(ids, 32 + 8, return_logits=True)
    err_ours = (logits[0] - ref32).abs().max().item()
    err_oracle = (ref16 - ref32).abs().max().item()
    print(f"gpt2-small: |ours-fp32| {err_ours:.4f}  |oracle_bf16-fp32| {err_oracle:.4f}  scale {ref32.abs().max():.2f}")
    assert err_ours <= 2.0 * err_oracle + 0.02
    assert out.shape == (1, 40) and torch.equal(eng.generate(ids, 40), out)  # graph == eager


@pytest.mark.parametrize("name", ["gpt2_tiny", "mistral_tiny"])
def test_batched_decode_other_families(cuda, name):
    """Batch 8 (tcgen05 split-K projections; GPT-2: bias + gelu_new epilogues, Mistral: sliding
    window) gives every row the tokens transformers gives it, over the margin-safe prefix."""
    c = load_case(name)
    eng, spec, _ = _engine(c)
    eng.unload()
    from dli_b200.engine import ShardEngine
    from oracle import decoder_ref as R

    eng = ShardEngine(spec, max_batch=8, max_seq_len=128)
    eng.load_state_dict(R.seeded_state_dict(R.ModelSpec(**spec.__dict__), c["seed"], c["std"]))
    B0, T = c["prompts"].shape
    N = c["tokens"].shape[1]
    reps = 8 // B0
    out = eng.generate(c["prompts"].repeat(reps, 1), T + N)
    for r in range(1, reps):
        assert torch.equal(out[:B0], out[r * B0:(r + 1) * B0])
    noise = 2.0 * c["hf_err"].max()
    checked = 0
    for b in range(B0):
        n_safe = int((c["margin"][:, b] > 2.0 * noise).int().cumprod(0).sum())
        checked += n_safe
        assert torch.equal(out[b, T:T + n_safe], c["tokens"][b, :n_safe]), (name, b, n_safe)
    assert checked > 0
