"""Chunk attention (paged K/V) throughput: tcgen05 vs mma.sync kernel (DLI_PREFILL_ATTN=mma)."""
import os, sys, torch
sys.path.insert(0, '.')
from dli_b200 import ops
H, KV, hd, page = 32, 8, 128, 16
for B, T, prefix in [(1, 512, 1536), (1, 2048, 0), (4, 1024, 7168), (1, 1024, 3072)]:
    S = prefix + T
    mp = (S + page - 1) // page
    pt = torch.randperm(B * mp, device='cuda').int().reshape(B, mp).contiguous()
    kc = torch.randn(B * mp, KV, page, hd, device='cuda').to(torch.bfloat16); vc = torch.randn_like(kc)
    qkv = torch.randn(B * T, (H + 2 * KV) * hd, device='cuda').to(torch.bfloat16)
    out = torch.empty(B * T, H * hd, device='cuda', dtype=torch.bfloat16)
    f = lambda: ops.attn_prefill_paged(qkv, out, kc, vc, pt, B, T, prefix, H, KV, hd, hd ** -0.5)
    for _ in range(3): f()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10): f()
    e1.record(); torch.cuda.synchronize()
    us = e0.elapsed_time(e1) / 10 * 1000
    fl = 4 * B * H * hd * (T * prefix + T * T / 2)
    print(f"{os.environ.get('DLI_PREFILL_ATTN','tc'):3s} B={B} T={T} prefix={prefix}: {us:8.1f} us  {fl / us / 1e6:6.1f} TFLOP/s")
