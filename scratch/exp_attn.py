import sys, torch
sys.path.insert(0, '.')
from dli_b200 import ops
torch.cuda.set_device(0)
H, KV, hd = 32, 8, 128
for (B, T) in [(1, 2048), (4, 2048), (1, 8192)]:
    bufs = [torch.randn(B * T, (H + 2 * KV) * hd, device='cuda').to(torch.bfloat16) for _ in range(8)]
    out = torch.empty(B * T, H * hd, device='cuda', dtype=torch.bfloat16)
    for tc in (True, False):
        for q in bufs: ops.attn_prefill(q, out, B, T, H, KV, hd, hd ** -0.5, tensor_core=tc)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(4):
            for q in bufs: ops.attn_prefill(q, out, B, T, H, KV, hd, hd ** -0.5, tensor_core=tc)
        e1.record(); torch.cuda.synchronize()
        us = e0.elapsed_time(e1) / 32 * 1000
        fl = 4 * B * (T * T / 2) * H * hd
        print(f'B={B} T={T} tc={tc}: {us:8.1f} us  {fl / us / 1e6:7.1f} TFLOP/s (causal-counted)')
