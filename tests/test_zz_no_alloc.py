"""Decode attention without a caller-provided scratch buffer.

`dli_paged_attn_decode_static_bf16` exists for users of the library who cannot
allocate device memory at the point of the call (all HBM already handed to their
own pools, a binding with no allocator, or a loop that cannot afford cudaMalloc's
device synchronisation): it keeps the split-KV partials in a fixed array inside the
library.  Here it is called through the C-ABI twice in a row — the second call with
other inputs and another split count, so whatever the first call left in the shared
array must not matter — and each result is compared with the ordinary entry point
given a scratch tensor AND with an fp32 PyTorch evaluation of the same attention, at
the tolerance the other attention tests use (the two entry points run the same kernel,
so agreeing with each other alone would only show the wiring).  The second call is also
bracketed by cudaMemGetInfo readings: a call that allocates nothing leaves the device's
free memory where it was.

This file sorts last on purpose: a run that stops at the first failure has then
already executed every other test.  It also passes alone in a fresh process (the
first call performs the per-device set-up).
"""
import pytest
import torch

pytestmark = pytest.mark.gpu

BF = torch.bfloat16


def _case(seed, B, H, KV, hd, page, lens, device):
    g = torch.Generator(device="cuda").manual_seed(seed)
    max_pages = (max(lens) + page - 1) // page
    n_pages = B * max_pages + 2
    page_table = torch.randperm(n_pages, device=device, generator=g)[: B * max_pages].int().reshape(B, max_pages).contiguous()
    kc = torch.randn(n_pages, KV, page, hd, device=device, generator=g).to(BF)
    vc = torch.randn(n_pages, KV, page, hd, device=device, generator=g).to(BF)
    q = torch.randn(B, H * hd, device=device, generator=g).to(BF)
    seq_lens = torch.tensor(lens, device=device, dtype=torch.int32)
    return q, kc, vc, page_table, seq_lens


def _fp32_reference(case, B, H, KV, hd, page):
    """softmax(q k^T / sqrt(hd)) v per sequence in fp32 on the gathered pages (what test_kernels_gpu.py
    checks dli_paged_attn_decode_bf16 against)."""
    q, kc, vc, page_table, seq_lens = case
    G, scale, rows = H // KV, hd ** -0.5, []
    for b in range(B):
        S = int(seq_lens[b])
        pages = page_table[b, : (S + page - 1) // page].long()
        k = kc[pages].permute(1, 0, 2, 3).reshape(KV, -1, hd)[:, :S].float()
        v = vc[pages].permute(1, 0, 2, 3).reshape(KV, -1, hd)[:, :S].float()
        p = torch.softmax(torch.einsum("kgd,ksd->kgs", q[b].float().reshape(KV, G, hd), k) * scale, dim=-1)
        rows.append(torch.einsum("kgs,ksd->kgd", p, v).reshape(-1))
    return torch.stack(rows)


def _both(lib, _lib, case, B, H, KV, hd, page, splits, expect_no_alloc=False):
    from dli_b200 import ops

    q, kc, vc, page_table, seq_lens = case
    stream = torch.cuda.current_stream().cuda_stream
    scale = hd ** -0.5
    out_static = torch.full((B, H * hd), float("nan"), device=q.device, dtype=BF)
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info()[0]
    rc = lib.dli_paged_attn_decode_static_bf16(
        q.data_ptr(), H * hd, kc.data_ptr(), vc.data_ptr(), page_table.data_ptr(), page_table.shape[1],
        seq_lens.data_ptr(), out_static.data_ptr(), B, H, KV, hd, page, splits, scale, 0, stream)
    _lib.check("dli_paged_attn_decode_static_bf16", rc)
    torch.cuda.synchronize()
    if expect_no_alloc:  # not on a device's first call: that one loads the kernel image lazily
        assert torch.cuda.mem_get_info()[0] == free_before, "the no-allocation entry point changed free device memory"
    out_ref = torch.empty_like(out_static)
    nbytes = ops.attn_decode_scratch_bytes(B, H, hd, splits)
    scratch = torch.empty(max(nbytes, 4) // 4, device=q.device, dtype=torch.float32)
    ops.paged_attn_decode(q, H * hd, kc, vc, page_table, seq_lens, out_ref, H, KV, hd, scale,
                          num_splits=splits, scratch=scratch)
    torch.cuda.synchronize()
    return out_static, out_ref


def test_static_scratch_entry_point_matches_on_two_calls_in_a_row(cuda):
    from dli_b200 import _lib

    lib = _lib.load()
    assert lib.dli_paged_attn_decode_static_max_bytes() == 8 << 20
    # first call: Llama-3-8B head geometry, long context, many splits
    case_a = _case(1, 2, 32, 8, 128, 16, [2048, 700], cuda)
    a, ref_a = _both(lib, _lib, case_a, 2, 32, 8, 128, 16, 16)
    torch.testing.assert_close(a.float(), ref_a.float(), rtol=2e-2, atol=2e-2)
    torch.testing.assert_close(a.float(), _fp32_reference(case_a, 2, 32, 8, 128, 16), rtol=2e-2, atol=2e-2)
    # second call straight after, different shapes and split count: stale partials of the
    # first call are still in the array and must not leak into this result
    case_b = _case(2, 3, 32, 4, 64, 32, [130, 5, 300], cuda)
    _both(lib, _lib, case_b, 3, 32, 4, 64, 32, 3)  # brings in the head_dim-64 kernel image (a lazy module load)
    b, ref_b = _both(lib, _lib, case_b, 3, 32, 4, 64, 32, 3, expect_no_alloc=True)
    torch.testing.assert_close(b.float(), ref_b.float(), rtol=2e-2, atol=2e-2)
    torch.testing.assert_close(b.float(), _fp32_reference(case_b, 3, 32, 4, 64, 32), rtol=2e-2, atol=2e-2)
    assert not torch.isnan(a.float()).any() and not torch.isnan(b.float()).any()


def test_static_scratch_rejects_oversized_requests(cuda):
    from dli_b200 import _lib

    lib = _lib.load()
    one = torch.zeros(8, device=cuda, dtype=torch.int32)
    # B=64, H=64, hd=128, 32 splits -> 68 MB of partials, far beyond the 8 MiB array
    rc = lib.dli_paged_attn_decode_static_bf16(
        one.data_ptr(), 64 * 128, one.data_ptr(), one.data_ptr(), one.data_ptr(), 1, one.data_ptr(), one.data_ptr(),
        64, 64, 8, 128, 16, 32, 1.0, 0, None)
    assert rc == -1
    msg = _lib.last_error()
    assert "built-in array holds 8388608" in msg and "needs" in msg
