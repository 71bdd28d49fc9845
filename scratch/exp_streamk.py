"""Stream-K vs uniform split-K weight stream (DLI_STREAMK=1/0, one process each): error pattern per 128-row tile
against an fp32 reference, and time per launch on the Llama-3-8B projection shapes at M rows."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dli_b200 import ops  # noqa: E402

dev = torch.device("cuda")
BF = torch.bfloat16
mode = os.environ.get("DLI_STREAMK", "1")


def check(M, N, K):
    g = torch.Generator(device="cuda").manual_seed(N + K + M)
    x = torch.randn(M, K, device=dev, generator=g).to(BF)
    w = (torch.randn(N, K, device=dev, generator=g) * 0.02).to(BF)
    sk = torch.zeros(ops.gemm_splitk_scratch_bytes(M, N), dtype=torch.uint8, device=dev)
    y = ops.linear_splitk(x, w, sk)
    y2 = ops.linear_splitk(x, w, sk)
    ref = (x.float() @ w.float().t())
    bad = ((y.float() - ref).abs() > 0.05 + 0.02 * ref.abs()).any(dim=0)  # per output column n
    tiles = bad.view(-1, 128).any(dim=1) if N % 128 == 0 else None
    total_kb = K // 64
    n_tiles = (N + 127) // 128
    total = n_tiles * total_kb
    G = min(148, total)
    info = []
    if tiles is not None:
        for t in tiles.nonzero().flatten().tolist()[:12]:
            lo, hi = t * total_kb, (t + 1) * total_kb
            owners = sorted({next(c for c in range(G) if total * c // G <= i < total * (c + 1) // G) for i in range(lo, hi)})
            cuts = [max(lo, total * c // G) - lo for c in owners]
            info.append((t, owners, cuts))
    print(f"streamk={mode} M={M} N={N} K={K}: bad columns {int(bad.sum())}, bad tiles {int(tiles.sum()) if tiles is not None else '?'}"
          f" of {n_tiles}, deterministic {bool(torch.equal(y, y2))}; first bad tiles (tile, owners, first kb of each): {info}")


def timeit(M, N, K, gated=False, reps=16):
    g = torch.Generator(device="cuda").manual_seed(1)
    x = torch.randn(M, K, device=dev, generator=g).to(BF)
    ws = [(torch.randn(N, K, device=dev, generator=g) * 0.02).to(BF) for _ in range(reps)]
    sk = torch.zeros(ops.gemm_splitk_scratch_bytes(M, N), dtype=torch.uint8, device=dev)
    out = torch.empty(M, N // 2 if gated else N, device=dev, dtype=BF)
    fn = (lambda w: ops.linear_splitk_gated_silu(x, w, sk, out=out)) if gated else (lambda w: ops.linear_splitk(x, w, sk, out=out))
    for w in ws:
        fn(w)
    torch.cuda.synchronize()
    best = 1e9
    for _ in range(3):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for w in ws:
            fn(w)
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1) / reps)
    print(f"streamk={mode} M={M} {'gate/up' if gated else 'linear'} N={N} K={K}: {best * 1e3:.1f} us, {N * K * 2 / best / 1e9:.2f} TB/s")


ops.set_pdl(True)
for shape in [(3, 1024, 256), (3, 128256, 256), (32, 4096, 4096), (5, 4096, 14336), (32, 6144, 4096)]:
    check(*shape)
for M in (32, 16):
    timeit(M, 6144, 4096)
    timeit(M, 4096, 4096)
    timeit(M, 4096, 14336)
    timeit(M, 28672, 4096, gated=True)
