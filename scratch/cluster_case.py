import sys, torch
sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
from dli_b200 import ops
import test_kernels_gpu as T
cuda = torch.device('cuda:0')
B,H,KV,hd,page,lens,splits,window = 1,32,8,128,16,[2049],8,0
gen = torch.Generator(device="cuda").manual_seed(31 + H)
page_table, kc, vc = T._paged_cache(B, max(lens), KV, hd, page, cuda, gen)
qkv = torch.randn(B, (H + 2 * KV) * hd, device=cuda, generator=gen).to(torch.bfloat16)
cos, sin = T._hf_rope_tables(hd, 10000.0, max(lens) + 1, cuda)
seq_lens = torch.tensor(lens, device=cuda, dtype=torch.int32)
out = torch.empty(B, H * hd, device=cuda, dtype=torch.bfloat16)
ops.paged_attn_decode_fused(qkv.clone(), cos, sin, kc, vc, page_table, seq_lens, out, H, KV, hd, hd ** -0.5, num_splits=splits, scratch=None, window=window)
torch.cuda.synchronize()
print('ok', out.float().abs().mean().item())
scratch = torch.empty(max(ops.attn_decode_scratch_bytes(B, H, hd, splits), 4) // 4, device=cuda, dtype=torch.float32)
for rep in range(3):
    for sc in (scratch, None):
        kc2, vc2 = kc.clone(), vc.clone()
        out = torch.empty(B, H * hd, device=cuda, dtype=torch.bfloat16)
        ops.paged_attn_decode_fused(qkv.clone(), cos, sin, kc2, vc2, page_table, seq_lens, out, H, KV, hd, hd ** -0.5, num_splits=splits, scratch=sc, window=window)
        torch.cuda.synchronize()
        print('rep', rep, 'scratch' if sc is not None else 'cluster', out.float().abs().mean().item(), flush=True)
