"""Times scratch/_abl/libdli_<variant>.so's prefill attention (one subprocess per variant: the library is loaded once)."""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if len(sys.argv) > 1:
    sys.path.insert(0, ROOT)
    import torch
    from dli_b200 import _lib
    _lib.LIB_PATH = os.path.join(ROOT, "scratch", "_abl", f"libdli_{sys.argv[1]}.so")
    from dli_b200 import ops
    H, KV, hd = 32, 8, 128
    res = []
    for T in (2048, 4096):
        qkv = [torch.randn(T, (H + 2 * KV) * hd, device="cuda").to(torch.bfloat16) for _ in range(8)]
        out = torch.empty(T, H * hd, device="cuda", dtype=torch.bfloat16)
        for x in qkv:
            ops.attn_prefill(x, out, 1, T, H, KV, hd, hd ** -0.5)
        torch.cuda.synchronize()
        best = 1e9
        for _ in range(5):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for x in qkv:
                ops.attn_prefill(x, out, 1, T, H, KV, hd, hd ** -0.5)
            e1.record()
            torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1) / len(qkv))
        res.append(f"T={T}: {best * 1e3:.1f} us")
    print(sys.argv[1], " ".join(res), flush=True)
else:
    for f in sorted(os.listdir(os.path.join(ROOT, "scratch", "_abl"))):
        if f.startswith("libdli_") and f.endswith(".so"):
            subprocess.run([sys.executable, __file__, f[len("libdli_"):-3]], timeout=120)
