import os, sys, torch
sys.path.insert(0, '.')
from dli_b200 import ops
from dli_b200.engine import NAMED_SPECS
torch.cuda.set_device(0)
spec = NAMED_SPECS['llama-3-8b']
d, ff, H, KV, hd = spec.d_model, spec.d_ff, spec.n_heads, spec.n_kv_heads, spec.head_dim
L = 16
M = int(sys.argv[1]) if len(sys.argv) > 1 else 32
def rnd(*s): return (torch.randn(*s, device='cuda') * 0.02).to(torch.bfloat16)
layers = [dict(ln1=rnd(d)+1, wqkv=rnd((H+2*KV)*hd, d), wo=rnd(d, H*hd), ln2=rnd(d)+1, wgu=rnd(2*ff, d), wdown=rnd(d, ff)) for _ in range(L)]
h = rnd(M, d); x = rnd(M, d); qkv = torch.empty(M, (H+2*KV)*hd, device='cuda', dtype=torch.bfloat16); attn = rnd(M, H*hd); act = rnd(M, ff)
sk = torch.zeros(ops.gemm_splitk_scratch_bytes(128, 2*ff), dtype=torch.uint8, device='cuda')
def chain(which):
    for Lw in layers:
        if 'q' in which: ops.linear_splitk(x, Lw['wqkv'], sk, out=qkv)
        if 'o' in which: ops.linear_splitk(attn, Lw['wo'], sk, residual=h, out=h)
        if 'g' in which: ops.linear_splitk_gated_silu(x, Lw['wgu'], sk, out=act)
        if 'd' in which: ops.linear_splitk(act, Lw['wdown'], sk, residual=h, out=h)
def timeit(which):
    chain(which); torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g): chain(which)
    for _ in range(3): g.replay()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    for _ in range(10): g.replay()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / 10 * 1000 / L
MB = dict(q=(H+2*KV)*hd*d*2/1e6, o=d*H*hd*2/1e6, g=2*ff*d*2/1e6, d=d*ff*2/1e6)
ops.set_pdl(True)
for which in ['q', 'o', 'g', 'd', 'qogd']:
    mb = sum(MB[c] for c in which)
    us = timeit(which)
    print(f'M={M} {which:5s} {us:7.2f} us/layer  {mb:7.1f} MB  {mb/us/1e3:.2f} TB/s')
