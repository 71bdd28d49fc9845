"""The decode pipeline over per-tick mailboxes (pipeline.MailboxRunner + ShardEngine.attach_mailbox), on ONE GPU:
N layer-range engines live in this process and the per-tick all-gather is replaced by local copies
(pipeline.LocalExchange), so the whole multi-stage data path — kernels reading a neighbour's hidden state straight out
of the incoming mailbox, writing their output straight into the outgoing mail, the vocabulary-parallel head slices, the
strided candidate merge, the per-slot CUDA graphs — is checked token for token against the single full-range engine
on the box the driver has.  (The same runner over NCCL: tests/test_pipeline_gpu.py, needs >= 2 GPUs.)"""
import pytest
import torch

from golden_util import load_case, spec_kwargs

pytestmark = pytest.mark.gpu


def _engines(case, n_stages, vp, max_batch, max_len, device):
    from dli_b200.engine import ModelSpec, ShardEngine
    from dli_b200.sharding import shard_ranges
    from oracle import decoder_ref as R

    spec = ModelSpec.from_hf_config(spec_kwargs(case["config"]))
    sd = R.seeded_state_dict(R.ModelSpec(**spec.__dict__), case["seed"], case["std"])
    full = ShardEngine(spec, device=str(device), max_batch=max_batch, max_seq_len=max_len)
    full.load_state_dict(sd)
    stages = []
    for r, (a, b) in enumerate(shard_ranges(spec.n_layers, n_stages)):
        e = ShardEngine(spec, a, b, device=str(device), max_batch=max_batch, max_seq_len=max_len,
                        head_shard=(r, n_stages) if vp else None)
        e.load_state_dict(sd)
        stages.append(e)
    return spec, full, stages


def _prefill_through_stages(stages, b, T, ids, mb, vp):
    """Prompt pass of slot mb through the stage chain; leaves the slot's first token in stage 0's state."""
    x = stages[0].prefill(b, T, ids=ids, mb=mb)
    for e in stages[1:]:
        x = e.prefill(b, T, hidden=x, mb=mb)
    if not vp:
        first = x.clone()  # the last stage's tokens
    else:
        for r, e in enumerate(stages):  # every stage: its head slice on the slot's final hidden state
            e._state(b, mb)["hf"].copy_(x)
            cand = e.head_slice_step(b, mb)
            stages[0]._state(b, mb)["cands"][r].copy_(cand)
        first = stages[0].combine_candidates(b, mb).clone()
    stages[0]._state(b, mb)["tokens"].copy_(first)
    for e in stages:
        e._state(b, mb)["step"].zero_()
    return first


@pytest.mark.parametrize("name,n_stages,vp,hop,b,n_slots", [
    ("llama_tiny", 2, True, 2, 1, 6), ("llama_tiny", 2, True, 1, 2, 3), ("llama_tiny", 2, False, 2, 1, 4),
    ("llama_hd128", 2, True, 2, 1, 2), ("mistral_tiny", 2, False, 1, 2, 2), ("llama_tiny", 1, True, 2, 1, 4),
])
def test_mailbox_pipeline_on_one_gpu_equals_the_single_engine(cuda, name, n_stages, vp, hop, b, n_slots):
    from dli_b200.pipeline import LocalExchange, MailboxRunner, MailStage, mailbox_period

    case = load_case(name)
    T, rounds = 9, 7
    max_len = T + rounds + 4
    spec, full, stages = _engines(case, n_stages, vp, n_slots * b, max_len, cuda)
    if len(stages) != n_stages:
        pytest.skip("model has fewer layers than stages")
    g = torch.Generator().manual_seed(5)
    prompts = [torch.randint(0, spec.vocab, (b, T), generator=g) for _ in range(n_slots)]
    want = [full.generate(p, T + rounds + 1)[:, T:] for p in prompts]  # [b, rounds + 1] per slot

    firsts = []
    for mb in range(n_slots):
        for e in stages:
            e.start_batch(b, T, max_len, mb=mb)
        ids = prompts[mb].to(device=cuda, dtype=torch.int32).reshape(-1).contiguous()
        firsts.append(_prefill_through_stages(stages, b, T, ids, mb, vp).cpu())

    P = mailbox_period(n_slots, n_stages, hop, vp)
    ex = LocalExchange()
    mstages = [MailStage(e, b, n_slots, P, hop, r, n_stages) for r, e in enumerate(stages)]
    ex.stages = mstages
    runners = [MailboxRunner(ms, r, n_stages, n_slots, hop=hop, vp=vp, exchange=ex.bind(r)) for r, ms in enumerate(mstages)]
    for rn in runners:
        for j in range(n_slots):
            rn.open_slot(j, rounds)
    while any(rn.busy() for rn in runners):
        for rn in runners:  # one tick on every stage, in rank order, like N processes in lockstep
            rn.tick()
    torch.cuda.synchronize()

    for mb in range(n_slots):
        if vp:  # stage 0 logs the token it consumes; the last one taken in is still in its input buffer
            st = stages[0]._state(b, mb)
            got = torch.cat((st["token_log"][:rounds], st["tokens"][None]), dim=0).t().cpu().long()
        else:   # the last stage logs the tokens it produces
            st = stages[-1]._state(b, mb)
            got = torch.cat((firsts[mb][None].to(cuda), st["token_log"][:rounds]), dim=0).t().cpu().long()
        assert torch.equal(got, want[mb]), (name, mb, got.tolist(), want[mb].tolist())
    # a second block continues the same generations (the serving loop runs block after block)
    more = 2
    if T + rounds + 1 + more <= max_len:
        want2 = [full.generate(p, T + rounds + 1 + more)[:, T + rounds + 1:] for p in prompts]
        for rn in runners:
            for j in range(n_slots):
                rn.open_slot(j, more)
        while any(rn.busy() for rn in runners):
            for rn in runners:
                rn.tick()
        torch.cuda.synchronize()
        for mb in range(n_slots):
            if vp:
                st = stages[0]._state(b, mb)
                got = torch.cat((st["token_log"][rounds + 1:rounds + more], st["tokens"][None]), dim=0).t().cpu().long()
            else:
                got = stages[-1]._state(b, mb)["token_log"][rounds:rounds + more].t().cpu().long()
            assert torch.equal(got, want2[mb]), (name, mb, "second block", got.tolist(), want2[mb].tolist())
