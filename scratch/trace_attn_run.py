import ctypes
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from dli_b200 import _lib
_lib.LIB_PATH = os.path.join(ROOT, "scratch", "_abl", "libdli_trace.so")
from dli_b200 import ops

H, KV, hd, T = 32, 8, 128, 2048
qkv = torch.randn(T, (H + 2 * KV) * hd, device="cuda").to(torch.bfloat16)
out = torch.empty(T, H * hd, device="cuda", dtype=torch.bfloat16)
for _ in range(3):
    ops.attn_prefill(qkv, out, 1, T, H, KV, hd, hd ** -0.5)
torch.cuda.synchronize()
buf = (ctypes.c_longlong * (3 * 64 * 8))()
lib = ctypes.CDLL(_lib.LIB_PATH)
assert lib.dli_dbg_trace(buf) == 0
tr = [[[buf[(r * 64 + t) * 8 + p] for p in range(8)] for t in range(64)] for r in range(3)]
t0 = tr[1][0][0]
print("tile | softmax: s_full seen, S loaded, max exchanged, exps done, P st issued, P st done, arrived | mma: loop top, next S issued, p_full seen, v_full seen, PV issued | producer: K issue, V issue   (cycles since the MMA thread's first loop top)")
for t in range(16):
    sm = [tr[0][t][p] - t0 for p in range(7)]
    mm = [tr[1][t][p] - t0 for p in range(5)]
    pr = [tr[2][t][p] - t0 for p in range(2)]
    print(f"{t:3d} | " + " ".join(f"{v:6d}" for v in sm) + " | " + " ".join(f"{v:6d}" for v in mm) + " | " + " ".join(f"{v:6d}" for v in pr))
