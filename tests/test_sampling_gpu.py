"""The reference's default sampling epilogue (temperature 0.8 -> top-k 50 -> top-p 0.95 ->
multinomial) on the GPU: the kept set and its probabilities equal what transformers' own
logits warpers produce; draws follow those probabilities; seeds reproduce."""
import pytest
import torch

pytestmark = pytest.mark.gpu
BF = torch.bfloat16


def _hf_kept(logits_bf16, temperature, top_k, top_p):
    """transformers' warpers on the fp32 copy, as _sample does (TF:generation/utils.py:2762-2791)."""
    from transformers.generation.logits_process import (TemperatureLogitsWarper, TopKLogitsWarper,
                                                          TopPLogitsWarper)

    scores = logits_bf16.float().cpu()
    ids = torch.zeros(scores.shape[0], 1, dtype=torch.long)
    for w in (TemperatureLogitsWarper(temperature), TopKLogitsWarper(top_k), TopPLogitsWarper(top_p)):
        scores = w(ids, scores)
    return torch.softmax(scores, dim=-1)  # zeros outside the kept set


@pytest.mark.parametrize("B,V,scale", [(4, 128256, 2.0), (3, 32000, 6.0), (2, 1000, 0.3), (1, 50257, 12.0)])
def test_kept_set_and_probabilities_match_transformers(cuda, B, V, scale):
    from dli_b200 import ops

    g = torch.Generator(device="cuda").manual_seed(V)
    logits = (torch.randn(B, V, device=cuda, generator=g) * scale).to(BF)
    tok, probs, ids = ops.sample(logits, debug_cap=64)
    ref = _hf_kept(logits, 0.8, 50, 0.95)
    for b in range(B):
        kept_ref = set(ref[b].nonzero().flatten().tolist())
        mine = [i for i in ids[b].tolist() if i >= 0]
        # same set, except that among EQUAL logits at the nucleus boundary torch.sort's tie order
        # decides which tokens stay: the symmetric difference may only hold tokens tied with the
        # smallest kept logit, and the sizes and the kept logit values must agree
        row = logits[b].float().cpu()
        assert len(mine) == len(kept_ref)
        diff = set(mine) ^ kept_ref
        assert all(row[i] == row[mine[-1]] for i in diff), (b, sorted(diff))
        assert sorted(row[mine].tolist()) == sorted(row[list(kept_ref)].tolist())
        p_ref = torch.sort(ref[b][ref[b] > 0], descending=True).values
        # fp32 softmax in a different summation order: 1e-5 absolute on probabilities
        torch.testing.assert_close(probs[b, : len(mine)].cpu(), p_ref, rtol=1e-4, atol=1e-5)
        assert row[int(tok[b])] >= row[mine[-1]]
        assert (probs[b, : len(mine) - 1] >= probs[b, 1: len(mine)]).all()  # descending


def test_top_k_keeps_ties_at_the_threshold_like_transformers(cuda):
    from dli_b200 import ops

    V = 4096
    logits = torch.full((1, V), -3.0, device=cuda).to(BF)
    logits[0, 100:140] = 2.0  # 40 tokens above
    logits[0, 500:530] = 1.0  # 30 tied tokens straddle the k = 50 threshold: all are kept
    _, probs, ids = ops.sample(logits, temperature=1.0, top_k=50, top_p=1.0, debug_cap=128)
    kept = sorted(i for i in ids[0].tolist() if i >= 0)
    assert kept == list(range(100, 140)) + list(range(500, 530))
    ref = _hf_kept(logits, 1.0, 50, 1.0)
    assert set(ref[0].nonzero().flatten().tolist()) == set(kept)


def test_draws_follow_the_distribution_and_seeds_reproduce(cuda):
    from dli_b200 import ops

    g = torch.Generator(device="cuda").manual_seed(5)
    V, n = 2000, 20000
    row = (torch.randn(1, V, device=cuda, generator=g) * 1.5).to(BF)
    logits = row.expand(n, V).contiguous()  # n independent rows with the same logits
    seed = torch.tensor([1234], device=cuda, dtype=torch.int64)
    step = torch.tensor([7], device=cuda, dtype=torch.int32)
    a = ops.sample(logits, seed=seed, step=step)
    b = ops.sample(logits, seed=seed, step=step)
    assert torch.equal(a, b)  # same (seed, step, row) -> same draw
    c = ops.sample(logits, seed=seed + 1, step=step)
    assert not torch.equal(a, c)
    _, probs, ids = ops.sample(row, debug_cap=64)
    kept = [(i, p) for i, p in zip(ids[0].tolist(), probs[0].tolist()) if i >= 0]
    counts = torch.bincount(a.long().cpu(), minlength=V)
    assert counts.sum() == n and set(counts.nonzero().flatten().tolist()) <= {i for i, _ in kept}
    chi2 = sum((counts[i].item() - n * p) ** 2 / (n * p) for i, p in kept if n * p >= 5)
    dof = sum(1 for _, p in kept if n * p >= 5) - 1
    assert chi2 < dof + 6 * (2 * dof) ** 0.5, (chi2, dof)  # far inside the chi-square tail


def test_engine_sampling_is_reproducible_and_differs_from_greedy(cuda):
    from golden_util import load_case, spec_kwargs
    from dli_b200.engine import ModelSpec, ShardEngine
    from oracle import decoder_ref as R

    c = load_case("llama_tiny")
    spec = ModelSpec.from_hf_config(spec_kwargs(c["config"]))
    sd = R.seeded_state_dict(R.ModelSpec(**spec.__dict__), c["seed"], c["std"])
    eng = ShardEngine(spec, max_batch=2, max_seq_len=128)
    eng.load_state_dict(sd)
    T = c["prompts"].shape[1]
    greedy = eng.generate(c["prompts"], T + 24)
    s1 = eng.generate(c["prompts"], T + 24, do_sample=True, seed=11)
    s2 = eng.generate(c["prompts"], T + 24, do_sample=True, seed=11)
    s3 = eng.generate(c["prompts"], T + 24, do_sample=True, seed=12)
    assert torch.equal(s1, s2) and not torch.equal(s1, s3) and not torch.equal(s1, greedy)
    assert torch.equal(eng.generate(c["prompts"], T + 24), greedy)  # switching back re-captures the greedy graph
    assert int(s1.max()) < spec.vocab and torch.equal(s1[:, :T], c["prompts"])
