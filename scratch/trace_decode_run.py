"""Prints the timeline of one decode layer (batch 1, 2 048-token context, 8B shape) from scratch/_abl/libdli_tl.so."""
import ctypes
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from dli_b200 import _lib
_lib.LIB_PATH = os.path.join(ROOT, "scratch", "_abl", "libdli_tl.so")
from dli_b200.engine import ModelSpec, ShardEngine

L, T = 6, 2048
B = int(sys.argv[1]) if len(sys.argv) > 1 else 1
spec = ModelSpec.from_hf_config(dict(model_type="llama", hidden_size=4096, intermediate_size=14336, num_hidden_layers=32,
                                     num_attention_heads=32, num_key_value_heads=8, vocab_size=128256, rope_theta=500000.0,
                                     rms_norm_eps=1e-5, max_position_embeddings=8192))
eng = ShardEngine(spec, 0, L - 1, max_batch=B, max_seq_len=T + 64)
eng.load_random(seed=0)
ids = torch.randint(0, spec.vocab, (B * T,), dtype=torch.int32, device="cuda")
eng.start_batch(B, T, T + 64)
eng.prefill(B, T, ids=ids)
for _ in range(4):
    eng.decode_step(B)
torch.cuda.synchronize()
lib = ctypes.CDLL(_lib.LIB_PATH)


def drain(fn):
    buf = (ctypes.c_ulonglong * (1 << 16))()
    n = ctypes.c_uint(0)
    assert fn(buf, ctypes.byref(n)) == 0
    return [(buf[4 * i] >> 32, buf[4 * i] & 0xffffffff, buf[4 * i + 1], buf[4 * i + 2], buf[4 * i + 3]) for i in range(min(n.value, 1 << 14))]


drain(lib.dli_dbg_tl_gemv), drain(lib.dli_dbg_tl_attn), drain(lib.dli_dbg_tl_gemm)  # discard warm-up
eng.decode_step(B)
torch.cuda.synchronize()
ev = drain(lib.dli_dbg_tl_gemv) + drain(lib.dli_dbg_tl_attn) + drain(lib.dli_dbg_tl_gemm)
ev = [e for e in ev if e[2] or e[3]]
ev.sort(key=lambda e: e[2] if e[0] != 2 else e[2])
t0 = min(e[2] for e in ev)
names = {106144: "qkv gemm", 104096: "o/down gemm", 128672: "gate/up gemm", 6144: "qkv gemv", 4096: "o/down gemv", 14336: "gate/up gemv", 1: "attention", 2: "attention (K/V issued, first tile landed, exit)", 3: "merge", 4: "attention (Q rotated, first P.V done, partials in smem)"}
print("kernel, CTA: entry / after dependency wait / exit   (ns since the step's first stamp)")
for k, blk, a, b, c in ev:
    print(f"{names.get(k, str(k)):48s} cta {blk:4d}: {a - t0:8d} {b - t0:8d} {c - t0:8d}   (wait {b - a:6d}, body {c - b:6d})")
