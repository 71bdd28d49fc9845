import os, sys, torch, json
sys.path.insert(0, '.')
from dli_b200 import ops
from dli_b200.engine import NAMED_SPECS
torch.cuda.set_device(0)
spec = NAMED_SPECS['llama-3-8b']
d, ff, H, KV, hd = spec.d_model, spec.d_ff, spec.n_heads, spec.n_kv_heads, spec.head_dim
L = 32
def rnd(*s): return (torch.randn(*s, device='cuda') * 0.02).to(torch.bfloat16)
layers = [dict(ln1=rnd(d)+1, wqkv=rnd((H+2*KV)*hd, d), wo=rnd(d, H*hd), ln2=rnd(d)+1, wgu=rnd(2*ff, d), wdown=rnd(d, ff)) for _ in range(L)]
h = rnd(1, d); qkv = torch.empty(1, (H+2*KV)*hd, device='cuda', dtype=torch.bfloat16); attn = rnd(1, H*hd); act = torch.empty(1, ff, device='cuda', dtype=torch.bfloat16)
def chain(which):
    for Lw in layers:
        if 'q' in which: ops.rmsnorm_linear(h, Lw['ln1'], 1e-5, Lw['wqkv'], out=qkv)
        if 'o' in which: ops.linear(attn, Lw['wo'], residual=h, out=h)
        if 'g' in which: ops.rmsnorm_linear_gated_silu(h, Lw['ln2'], 1e-5, Lw['wgu'], out=act)
        if 'd' in which: ops.linear(act, Lw['wdown'], residual=h, out=h)
def timeit(which, pdl):
    ops.set_pdl(pdl)
    chain(which); torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g): chain(which)
    for _ in range(3): g.replay()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    for _ in range(20): g.replay()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / 20 * 1000 / L
MB = dict(q=(H+2*KV)*hd*d*2/1e6, o=d*H*hd*2/1e6, g=2*ff*d*2/1e6, d=d*ff*2/1e6)
for which in ['q', 'o', 'g', 'd', 'qogd']:
    mb = sum(MB[c] for c in which)
    for pdl in (False, True):
        us = timeit(which, pdl)
        print(f'{which:5s} pdl={int(pdl)} {us:7.2f} us/layer  {mb:7.1f} MB  {mb/us/1e3*1e3:7.1f} GB/s -> {mb/us*1e-3:.2f} TB/s')
