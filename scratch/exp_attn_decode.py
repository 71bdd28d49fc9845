"""Decode attention alone at batch 32 (268 MB of K/V per launch): time vs number of KV splits."""
import sys, torch
sys.path.insert(0, '.')
from dli_b200 import ops
torch.cuda.set_device(0)
H, KV, hd, page = 32, 8, 128, 16
for B, S in [(8, 2048), (16, 2048), (32, 2048), (64, 2048), (8, 8192)]:
    max_pages = (S + page - 1) // page
    n_pages = B * max_pages
    pt = torch.randperm(n_pages, device='cuda').int().reshape(B, max_pages).contiguous()
    caches = [(torch.randn(n_pages, KV, page, hd, device='cuda').to(torch.bfloat16),
               torch.randn(n_pages, KV, page, hd, device='cuda').to(torch.bfloat16)) for _ in range(2)]
    qkv = torch.randn(B, (H + 2 * KV) * hd, device='cuda').to(torch.bfloat16)
    out = torch.empty(B, H * hd, device='cuda', dtype=torch.bfloat16)
    seq = torch.full((B,), S, device='cuda', dtype=torch.int32)
    mb = B * S * KV * hd * 2 * 2 / 1e6
    for splits in (1, 2, 4):
        nb = ops.attn_decode_scratch_bytes(B, H, hd, splits)
        scratch = torch.empty(max(nb // 4, 1), device='cuda', dtype=torch.float32)
        def run():
            for kc, vc in caches:
                ops.paged_attn_decode(qkv, qkv.shape[1], kc, vc, pt, seq, out, H, KV, hd, hd ** -0.5,
                                      num_splits=splits, scratch=scratch)
        for pdl in (1,):
            ops.set_pdl(bool(pdl))
            run(); torch.cuda.synchronize()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                for _ in range(4): run()
            g.replay(); torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); g.replay(); g.replay(); e1.record(); torch.cuda.synchronize()
            us = e0.elapsed_time(e1) / 16 * 1000
            print(f'B={B} S={S} splits={splits:2d} pdl={pdl}: {us:7.1f} us  {mb / us:5.2f} TB/s   ({mb:.0f} MB)')
