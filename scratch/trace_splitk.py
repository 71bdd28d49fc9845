"""Phase trace of the swap-AB split-K kernel (temporary instrumentation build only)."""
import ctypes, sys, torch, numpy as np
sys.path.insert(0, '.')
from dli_b200 import ops, _lib
lib = _lib.load()
raw = ctypes.CDLL(lib._name)
M = 32
def rnd(*s): return (torch.randn(*s, device='cuda') * 0.02).to(torch.bfloat16)
cases = {'o': (4096, 4096), 'qkv': (6144, 4096), 'down': (4096, 14336)}
NCTA = {'o': 128, 'qkv': 144, 'down': 128}
x4 = rnd(M, 4096); x14 = rnd(M, 14336)
ws = {k: [rnd(n, kk) for _ in range(4)] for k, (n, kk) in cases.items()}
sk = torch.zeros(ops.gemm_splitk_scratch_bytes(128, 28672), dtype=torch.uint8, device='cuda')
names = ['entry', 'setup', 'ring+wait', 'first_full', 'last_mma_issued', 'acc_ready', 'partials_out', 'atomic_back', 'end']
for k, (n, kk) in cases.items():
    x = x4 if kk == 4096 else x14
    res = rnd(M, n)
    for w in ws[k]:
        ops.linear_splitk(x, w, sk, residual=res, out=res)
    torch.cuda.synchronize()
    buf = (ctypes.c_longlong * (4096 * 10))()
    raw.dli_debug_trace(buf, 4096 * 10)
    t = np.frombuffer(buf, dtype=np.int64).reshape(4096, 10)
    ncta = NCTA[k]
    t = t[:ncta]
    g0 = t[:, 9].min()
    print(f'--- {k}: N={n} K={kk}, {ncta} CTAs; CTA start spread (globaltimer) {(t[:,9]-g0).max()} ns')
    rel = (t[:, :9] - t[:, :1]) / 1.9e3   # us at ~1.9 GHz
    last = rel[:, 8] > 0
    for i, nm in enumerate(names):
        col = rel[:, i]
        ok = t[:, i] != 0
        print(f'  {nm:16s} median {np.median(col[ok]):7.2f} us   max {col[ok].max():7.2f}   (n={int(ok.sum())})')
    fin = (t[:, 8] != 0)
    red = rel[:, 8] - rel[:, 7]
    print(f'  reducers: {int((red > 1.0).sum())} CTAs, reduce+epilogue median {np.median(red[red > 1.0]) if (red>1.0).any() else 0:.2f} us')
