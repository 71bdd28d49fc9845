"""Phase trace of the tcgen05 prefill attention (temporary instrumentation build only)."""
import ctypes, sys, torch, numpy as np
sys.path.insert(0, '.')
from dli_b200 import ops, _lib
raw = ctypes.CDLL(_lib.load()._name)
H, KV, hd, B, T = 32, 8, 128, 1, 2048
qkv = torch.randn(B * T, (H + 2 * KV) * hd, device='cuda').to(torch.bfloat16)
out = torch.empty(B * T, H * hd, device='cuda', dtype=torch.bfloat16)
for _ in range(3): ops.attn_prefill(qkv, out, B, T, H, KV, hd, hd ** -0.5, tensor_core=True)
torch.cuda.synchronize()
buf = (ctypes.c_longlong * (64 * 16))()
raw.dli_debug_atrace(buf, 64 * 16)
t = np.frombuffer(buf, dtype=np.int64).reshape(64, 16)
n = int((t[:, 0] != 0).sum())
names = ['wait S', 'ld S', '(arrive)+mask+max', 'bar.sync', 'exp+sum', 'wait PV(it-1)', 'P store', 'O rescale', 'fence+arrive']
print('iterations traced:', n, ' (CTA 0 = query tile 15, head 0; clocks)')
d = np.diff(t[:n, :10], axis=1)
for i, nm in enumerate(names):
    print(f'  {nm:20s} median {np.median(d[1:, i]):8.0f}   first it {d[0, i]:8.0f}   max {d[:, i].max():8.0f}')
per_it = np.diff(t[:n, 0])
print('  per iteration (start to start): median', np.median(per_it), 'clk')
