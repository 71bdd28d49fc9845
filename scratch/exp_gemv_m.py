"""GEMV bandwidth vs number of activation rows M (1..4 CUDA-core path; 5..16 split-K tcgen05 path)."""
import sys, torch
sys.path.insert(0, '.')
from dli_b200 import ops
torch.cuda.set_device(0)
ops.set_pdl(True)
def rnd(*s): return (torch.randn(*s, device='cuda') * 0.02).to(torch.bfloat16)
shapes = {'gate/up 8B': (28672, 4096, True), 'down 8B': (4096, 14336, False), 'qkv 8B': (6144, 4096, False),
          'gate/up 70B': (57344, 8192, True), 'down 70B': (8192, 28672, False), 'o 70B': (8192, 8192, False)}
sk = torch.zeros(ops.gemm_splitk_scratch_bytes(128, 57344), dtype=torch.uint8, device='cuda')
for name, (N, K, gated) in shapes.items():
    ws = [rnd(N, K) for _ in range(3)]
    for M in (1, 2, 3, 4, 8, 16):
        x = rnd(M, K)
        out = torch.empty(M, N // 2 if gated else N, device='cuda', dtype=torch.bfloat16)
        def run():
            for w in ws:
                if M <= 4:
                    (ops.linear_gated_silu(x, w, out=out) if gated else ops.linear(x, w, out=out))
                else:
                    (ops.linear_splitk_gated_silu(x, w, sk, out=out) if gated else ops.linear_splitk(x, w, sk, out=out))
        run(); torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            for _ in range(4): run()
        g.replay(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); g.replay(); e1.record(); torch.cuda.synchronize()
        us = e0.elapsed_time(e1) / 12 * 1000
        line = f'{name:12s} M={M:2d}: {us:7.1f} us  {N * K * 2 / us / 1e6:5.2f} TB/s'
        if M >= 2:
            def run2():
                for w in ws:
                    (ops.gemv_mma_gated_silu(x, w, out=out) if gated else ops.gemv_mma(x, w, out=out))
            run2(); torch.cuda.synchronize()
            g2 = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g2):
                for _ in range(4): run2()
            g2.replay(); torch.cuda.synchronize()
            e0.record(); g2.replay(); e1.record(); torch.cuda.synchronize()
            us2 = e0.elapsed_time(e1) / 12 * 1000
            line += f'   | mma stream {us2:7.1f} us  {N * K * 2 / us2 / 1e6:5.2f} TB/s'
        print(line)
