"""One 2048-token prefill of the first layers of the 8B shape (for ncu launch lists): python scratch/prefill_only.py [layers]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from dli_b200.engine import ModelSpec, ShardEngine

L = int(sys.argv[1]) if len(sys.argv) > 1 else 4
T = int(sys.argv[2]) if len(sys.argv) > 2 else 2048
spec = ModelSpec.from_hf_config(dict(model_type="llama", hidden_size=4096, intermediate_size=14336, num_hidden_layers=32,
                                     num_attention_heads=32, num_key_value_heads=8, vocab_size=128256, rope_theta=500000.0,
                                     rms_norm_eps=1e-5, max_position_embeddings=8192))
eng = ShardEngine(spec, 0, L - 1, max_batch=1, max_seq_len=T + 64)
eng.load_random(seed=0)
ids = torch.randint(0, spec.vocab, (T,), dtype=torch.int32, device="cuda")
for _ in range(3):
    eng.start_batch(1, T, T + 64)
    eng.prefill(1, T, ids=ids)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
best = 1e9
for _ in range(5):
    eng.start_batch(1, T, T + 64)
    e0.record()
    eng.prefill(1, T, ids=ids)
    e1.record()
    torch.cuda.synchronize()
    best = min(best, e0.elapsed_time(e1))
print(f"prefill of {L} layers, T={T}: {best * 1e3:.1f} us = {best * 1e3 / L:.1f} us per layer")
