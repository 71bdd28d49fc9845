"""The stopping rule of generate(): the reference calls transformers' generate (worker/app.py:297-305),
which ends a sequence at an EOS id, feeds pad_token_id afterwards and stops when every sequence is done.
CPU: apply_eos == transformers on a tiny random model.  GPU: the engine stops early and returns the same."""
import pytest
import torch

from golden_util import MODEL_CASES, load_case, spec_kwargs


def _tiny_hf():
    from transformers import LlamaConfig, LlamaForCausalLM

    torch.manual_seed(3)
    cfg = LlamaConfig(vocab_size=97, hidden_size=32, intermediate_size=64, num_hidden_layers=2, num_attention_heads=4,
                      num_key_value_heads=2, max_position_embeddings=64)
    return LlamaForCausalLM(cfg).eval()


@pytest.mark.parametrize("pad", [None, 0])
def test_apply_eos_is_transformers_rule(pad):
    from dli_b200.engine import apply_eos

    m = _tiny_hf()
    g = torch.Generator().manual_seed(5)
    prompts = torch.randint(3, 97, (3, 6), generator=g)
    n = 20
    with torch.no_grad():
        free = m.generate(prompts, attention_mask=torch.ones_like(prompts), max_length=6 + n, do_sample=False,
                          eos_token_id=None, pad_token_id=0)[:, 6:]
    # choose EOS ids that actually occur: one early in sequence 0, one that another sequence reaches later
    candidates = [int(free[0, 3]), int(free[1, 9])]
    for eos in (candidates[:1], candidates):
        with torch.no_grad():
            want = m.generate(prompts, attention_mask=torch.ones_like(prompts), max_length=6 + n, do_sample=False,
                              eos_token_id=eos, pad_token_id=eos[0] if pad is None else pad)[:, 6:]
        got = apply_eos(free.clone(), tuple(eos), pad)
        assert got.shape == want.shape and torch.equal(got, want), (eos, got.tolist(), want.tolist())


def test_apply_eos_edge_cases():
    from dli_b200.engine import all_finished, apply_eos

    new = torch.tensor([[5, 2, 7, 8], [1, 1, 1, 1]])
    assert torch.equal(apply_eos(new, (), None), new)
    assert torch.equal(apply_eos(new, (9,), None), new)            # never produced
    assert apply_eos(new, (2,), 0).tolist() == [[5, 2, 0, 0], [1, 1, 1, 1]]  # one unfinished: full length
    assert apply_eos(new, (2, 1), 0).tolist() == [[5, 2], [1, 0]]  # cut after the last sequence's EOS
    assert apply_eos(new[:, :0], (2,), 0).shape == (2, 0)
    assert all_finished(new.t(), (2, 1)) and not all_finished(new.t(), (2,))


@pytest.mark.gpu
@pytest.mark.parametrize("name", MODEL_CASES[:2])
def test_engine_stops_at_eos(cuda, name):
    from dli_b200.engine import ModelSpec, ShardEngine, apply_eos
    from oracle import decoder_ref as R

    c = load_case(name)
    spec = ModelSpec.from_hf_config(spec_kwargs(c["config"]))
    eng = ShardEngine(spec, max_batch=4, max_seq_len=128)
    eng.load_state_dict(R.seeded_state_dict(R.ModelSpec(**spec.__dict__), c["seed"], c["std"]))
    prompts = c["prompts"]
    T = prompts.shape[1]
    n = 40
    free = eng.generate(prompts, T + n)
    steps_free = eng.kv  # noqa: F841 (keeps the engine alive)
    new = free[:, T:]
    # every sequence reaches one of these ids: the token each produces at step 10 + its index
    eos = tuple(sorted({int(new[b, 10 + b]) for b in range(new.shape[0])}))
    eng.eos_token_ids, eng.pad_token_id = eos, 0
    from dli_b200 import ops

    before = ops.launch_count
    got = eng.generate(prompts, T + n)
    used = ops.launch_count - before
    want = torch.cat((prompts.cpu().long(), apply_eos(new.clone(), eos, 0)), dim=1)
    assert torch.equal(got, want)
    assert got.shape[1] < T + n  # cut after the last sequence's EOS ...
    eng.eos_token_ids = ()
    before = ops.launch_count
    eng.generate(prompts, T + n)
    assert used < ops.launch_count - before  # ... and the engine really stopped early
