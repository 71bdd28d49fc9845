"""One batch-32 decode step of Llama-3-8B (8 layers are enough for a per-kernel launch list) for ncu."""
import os, sys, torch
sys.path.insert(0, '.')
from dli_b200.engine import ShardEngine, NAMED_SPECS
import dataclasses
B, ctx = int(os.environ.get("B", 32)), int(os.environ.get("CTX", 2048))
spec = dataclasses.replace(NAMED_SPECS['llama-3-8b'], n_layers=8)
eng = ShardEngine(spec, max_batch=B, max_seq_len=ctx + 16, page_size=16)
eng.load_random(seed=0)
g = torch.Generator().manual_seed(2)
ids = torch.randint(0, spec.vocab, (B * 64,), generator=g, dtype=torch.int32).cuda()
eng.start_batch(B, ctx, ctx + 16)
# fake a long context without paying a 2k prefill: prefill 64 tokens, then set the lengths
eng.prefill(B, 64, ids=ids)
st = eng._state(B)
st["positions"].fill_(ctx); st["seq_lens"].fill_(ctx + 1)
for _ in range(4):
    eng.decode_step(B)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(8): eng.decode_step(B)
e1.record(); torch.cuda.synchronize()
print("B", B, "ms/step (8 layers)", e0.elapsed_time(e1) / 8)
