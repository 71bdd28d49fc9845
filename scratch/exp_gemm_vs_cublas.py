"""Prefill GEMM shapes of the 8B model: this library's persistent tcgen05 kernel against torch.matmul (cuBLAS), us per call
over rotating weight copies (inputs larger than L2 in total)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from dli_b200 import ops

M = int(sys.argv[1]) if len(sys.argv) > 1 else 2048


def timeit(fn, n):
    for i in range(n):
        fn(i)
    torch.cuda.synchronize()
    best = 1e9
    for _ in range(5):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(n):
            fn(i)
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1) / n)
    return best * 1e3


for name, N, K in (("qkv", 6144, 4096), ("o", 4096, 4096), ("down", 4096, 14336), ("gate_up(ungated)", 28672, 4096)):
    n = 4
    ws = [(torch.randn(N, K, device="cuda") * 0.02).to(torch.bfloat16) for _ in range(n)]
    x = torch.randn(M, K, device="cuda").to(torch.bfloat16)
    y = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
    t_ours = timeit(lambda i: ops.linear(x, ws[i % n], out=y), n)
    t_cublas = timeit(lambda i: torch.matmul(x, ws[i % n].t(), out=y), n)
    ref = torch.matmul(x, ws[0].t())
    ops.linear(x, ws[0], out=y)
    err = (y.float() - ref.float()).abs().max().item()
    fl = 2.0 * M * N * K
    print(f"[tile={os.environ.get('DLI_GEMM_TILE', 'auto')}] max|ours - cuBLAS| = {err:.4f}  ", end="")
    print(f"{name:18s} M={M} N={N} K={K}: ours {t_ours:7.1f} us ({fl / t_ours / 1e6:6.0f} TF)   cuBLAS {t_cublas:7.1f} us ({fl / t_cublas / 1e6:6.0f} TF)")

F, K = 14336, 4096
n = 4
wg = [ops.pack_gate_up((torch.randn(F, K, device="cuda") * 0.02).to(torch.bfloat16), (torch.randn(F, K, device="cuda") * 0.02).to(torch.bfloat16)) for _ in range(n)]
x = torch.randn(M, K, device="cuda").to(torch.bfloat16)
y = torch.empty(M, F, device="cuda", dtype=torch.bfloat16)
t = timeit(lambda i: ops.linear_gated_silu(x, wg[i % n], out=y), n)
print(f"[tile={os.environ.get('DLI_GEMM_TILE', 'auto')}] gated gate/up M={M} F={F} K={K}: ours {t:7.1f} us ({4.0 * M * F * K / t / 1e6:6.0f} TF)")
