"""Continuous batching: concurrent requests of different prompt and output lengths share decode
steps and each gets exactly the tokens it gets when it is the only request in the batcher."""
import threading

import pytest
import torch

from golden_util import load_case, spec_kwargs

pytestmark = pytest.mark.gpu


def _engine(max_batch):
    from dli_b200.engine import ModelSpec, ShardEngine
    from oracle import decoder_ref as R

    c = load_case("llama_tiny")
    spec = ModelSpec.from_hf_config(spec_kwargs(c["config"]))
    sd = R.seeded_state_dict(R.ModelSpec(**spec.__dict__), c["seed"], c["std"])
    eng = ShardEngine(spec, max_batch=max_batch, max_seq_len=128)
    eng.load_state_dict(sd)
    return eng, spec, c


def test_requests_join_and_leave_without_changing_each_other(cuda):
    from dli_b200.batching import ContinuousBatcher

    eng, spec, c = _engine(4)
    g = torch.Generator().manual_seed(3)
    reqs = [(torch.randint(0, spec.vocab, (T,), generator=g), T + n) for T, n in
            [(5, 9), (17, 20), (9, 1), (30, 12), (3, 25), (12, 7), (21, 16)]]
    # alone: one request at a time through the same batcher (same kernels, same batch width)
    b = ContinuousBatcher(eng, max_batch=4)
    alone = []
    for ids, ml in reqs:
        f = b.submit(ids, ml)
        while not f.done():
            b.step()
        alone.append(f.result())
    # together: all submitted at once; 7 requests over 4 slots, so slots are reused mid-flight
    futs = [b.submit(ids, ml) for ids, ml in reqs]
    steps0 = b.steps
    while not all(f.done() for f in futs):
        b.step()
    for (ids, ml), a, f in zip(reqs, alone, futs):
        out = f.result()
        assert out.shape == (ml,) and torch.equal(out[: len(ids)], ids)
        assert torch.equal(out, a)
    # sharing steps: far fewer decode steps than the sum of the requests' lengths
    assert b.steps - steps0 < sum(ml - len(ids) for ids, ml in reqs) * 0.6
    # and the first fixture prompt matches the plain engine over its margin-safe prefix
    T = c["prompts"].shape[1]
    f = b.submit(c["prompts"][0], T + 12)
    while not f.done():
        b.step()
    noise = 2.0 * c["hf_err"].max()
    n_safe = min(12, int((c["margin"][:, 0] > 2.0 * noise).int().cumprod(0).sum()))
    assert torch.equal(f.result()[T:T + n_safe], c["tokens"][0, :n_safe])
    assert len(eng.kv.free) == eng.kv.num_pages  # every sequence's pages were returned


def test_background_thread_serves_concurrent_submitters(cuda):
    from dli_b200.batching import ContinuousBatcher

    eng, spec, _ = _engine(4)
    b = ContinuousBatcher(eng).start()
    g = torch.Generator().manual_seed(4)
    prompts = [torch.randint(0, spec.vocab, (6 + i,), generator=g) for i in range(6)]
    results = [None] * 6

    def client(i):
        results[i] = b.submit(prompts[i], len(prompts[i]) + 10 + i).result(timeout=60)

    threads = [threading.Thread(target=client, args=(i,)) for i in range(6)]
    [t.start() for t in threads]
    [t.join(timeout=90) for t in threads]
    b.stop()
    for i, r in enumerate(results):
        assert r is not None and r.shape == (len(prompts[i]) + 10 + i,) and torch.equal(r[: len(prompts[i])], prompts[i])
    with pytest.raises(ValueError, match="Input length of input_ids is"):  # transformers' own refusal
        b.submit(prompts[0], 3).result()
    with pytest.raises(ValueError):
        b.submit(prompts[0], 1000).result()


def test_request_ends_at_its_eos(cuda):
    """A request leaves the batch right after it produced an EOS id (the EOS is part of the result, as in
    transformers), the others keep going."""
    from dli_b200.batching import ContinuousBatcher

    eng, spec, c = _engine(4)
    g = torch.Generator().manual_seed(8)
    reqs = [(torch.randint(0, spec.vocab, (T,), generator=g), T + 30) for T in (6, 11, 19)]
    b = ContinuousBatcher(eng, max_batch=4)
    futs = [b.submit(ids, ml) for ids, ml in reqs]
    while not all(f.done() for f in futs):
        b.step()
    free = [f.result() for f in futs]
    eos = int(free[0][6 + 7])  # the 8th new token of request 0
    eng.eos_token_ids = (eos,)
    futs = [b.submit(ids, ml) for ids, ml in reqs]
    while not all(f.done() for f in futs):
        b.step()
    for (ids, ml), fr, f in zip(reqs, free, futs):
        new = fr[len(ids):]
        hits = (new == eos).nonzero()
        cut = len(ids) + (int(hits[0]) + 1 if len(hits) else len(new))
        assert torch.equal(f.result(), fr[:cut])
    assert len(futs[0].result()) <= 6 + 8
    assert len(eng.kv.free) == eng.kv.num_pages


def test_sampled_requests_are_reproducible_and_independent_of_the_batch(cuda):
    """do_sample requests (the reference's default call) draw with their own seed, keyed by token position:
    the same request gives the same tokens alone or next to other requests, in whichever row; greedy requests
    in the same batch are unaffected; another seed gives another continuation."""
    from dli_b200.batching import ContinuousBatcher

    eng, spec, c = _engine(4)
    g = torch.Generator().manual_seed(21)
    prompts = [torch.randint(0, spec.vocab, (T,), generator=g) for T in (7, 13, 5, 10)]
    b = ContinuousBatcher(eng, max_batch=4)

    def run(jobs):
        futs = [b.submit(p, len(p) + 24, do_sample=ds, seed=sd) for p, ds, sd in jobs]
        while not all(f.done() for f in futs):
            b.step()
        return [f.result() for f in futs]

    alone_s = run([(prompts[0], True, 123)])[0]
    alone_g = run([(prompts[1], False, 0)])[0]
    together = run([(prompts[2], True, 9), (prompts[1], False, 0), (prompts[3], True, 77), (prompts[0], True, 123)])
    assert torch.equal(together[3], alone_s)  # same seed, other row, other neighbours
    assert torch.equal(together[1], alone_g)  # greedy row untouched by the sampling kernel
    assert not torch.equal(run([(prompts[0], True, 124)])[0], alone_s)
    assert not torch.equal(run([(prompts[0], False, 0)])[0], alone_s)  # and it is not just the argmax


def test_kv_pages_are_taken_on_demand_from_a_shared_pool(cuda):
    """A page pool smaller than max_batch x max_seq_len: sequences reserve pages as they grow, a request that finds
    the pool empty waits for a running one to leave, and every request still gets the tokens it gets alone."""
    from dli_b200.batching import ContinuousBatcher
    from dli_b200.engine import ModelSpec, ShardEngine
    from oracle import decoder_ref as R

    c = load_case("llama_tiny")
    spec = ModelSpec.from_hf_config(spec_kwargs(c["config"]))
    sd = R.seeded_state_dict(R.ModelSpec(**spec.__dict__), c["seed"], c["std"])
    g = torch.Generator().manual_seed(9)
    reqs = [(torch.randint(0, spec.vocab, (T,), generator=g), T + n) for T, n in [(20, 30), (18, 40), (25, 22), (9, 50), (30, 31)]]
    full = ShardEngine(spec, max_batch=4, max_seq_len=128)
    full.load_state_dict(sd)
    b = ContinuousBatcher(full, max_batch=4)
    alone = []
    for ids, ml in reqs:
        f = b.submit(ids, ml)
        while not f.done():
            b.step()
        alone.append(f.result())
    small = ShardEngine(spec, max_batch=4, max_seq_len=128, kv_pool_tokens=160)  # 10 pages of 16 for up to 4 x 8
    small.load_state_dict(sd)
    assert small.kv.num_pages == 10
    b2 = ContinuousBatcher(small, max_batch=4)
    b2.grow = 16
    futs = [b2.submit(ids, ml) for ids, ml in reqs]
    low = small.kv.num_pages
    for _ in range(2000):
        if all(f.done() for f in futs):
            break
        b2.step()
        low = min(low, len(small.kv.free))
    assert all(f.done() for f in futs)
    for f, want in zip(futs, alone):
        assert torch.equal(f.result(), want)
    assert low <= 2 and len(small.kv.free) == 10  # the pool really ran (nearly) dry, and everything came back
    # a long request takes most of the pool once it has drained
    big = b2.submit(torch.randint(0, spec.vocab, (100,), generator=g), 127)
    for _ in range(400):
        if big.done():
            break
        b2.step()
    assert big.done() and big.exception() is None  # 127 tokens = 8 pages <= 10: fits once the pool is empty again
