"""Time of the tcgen05 prefill attention alone (Llama-3-8B heads, T = 2048): CUDA events over back-to-back launches."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dli_b200 import ops  # noqa: E402

H, KV, hd = 32, 8, 128
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
    flops = 4 * (T * T / 2) * H * hd
    print(f"T={T}: {best * 1e3:.1f} us per launch, {flops / best / 1e9:.0f} TFLOP/s (causal)")
