"""One process, two GPUs: the library keeps its one-time kernel set-up (shared-memory attributes, occupancy, the
no-allocation entry point's scratch address) per DEVICE, so a host that drives several GPUs from one process — any
binding of include/dli.h that is not one-rank-per-GPU — can call every kernel family on the second device after the
first.  Skipped on a one-GPU box."""
import pytest
import torch

pytestmark = pytest.mark.gpu
BF = torch.bfloat16


def _run_all(dev):
    from dli_b200 import _lib, ops

    torch.cuda.set_device(dev)
    g = torch.Generator(device=dev).manual_seed(3)
    out = {}
    x = torch.randn(2, 4096, device=dev, generator=g).to(BF)
    w = (torch.randn(1024, 4096, device=dev, generator=g) * 0.02).to(BF)
    ln = torch.ones(4096, device=dev, dtype=BF)
    out["gemv"] = ops.rmsnorm_linear(x, ln, 1e-5, w)                                  # > 48 KB dynamic shared memory
    xm = torch.randn(300, 4096, device=dev, generator=g).to(BF)
    out["gemm"] = ops.gemm(xm, w)                                                     # tcgen05 prefill GEMM
    x32 = torch.randn(32, 4096, device=dev, generator=g).to(BF)
    sk = torch.zeros(ops.gemm_splitk_scratch_bytes(32, 1024), dtype=torch.uint8, device=dev)
    out["splitk"] = ops.linear_splitk(x32, w, sk)                                     # swap-AB split-K
    H, KV, hd, page, S = 32, 8, 128, 16, 600
    qkv = torch.randn(1 * 64, (H + 2 * KV) * hd, device=dev, generator=g).to(BF)
    o = torch.empty(64, H * hd, device=dev, dtype=BF)
    out["prefill_attn"] = ops.attn_prefill(qkv, o, 1, 64, H, KV, hd, hd ** -0.5)      # tcgen05 flash attention
    n_pages = (S + page - 1) // page
    pt = torch.arange(n_pages, device=dev, dtype=torch.int32)[None].contiguous()
    kc = torch.randn(n_pages, KV, page, hd, device=dev, generator=g).to(BF)
    vc = torch.randn(n_pages, KV, page, hd, device=dev, generator=g).to(BF)
    q = torch.randn(1, H * hd, device=dev, generator=g).to(BF)
    lens = torch.tensor([S], device=dev, dtype=torch.int32)
    od = torch.empty(1, H * hd, device=dev, dtype=BF)
    rc = _lib.load().dli_paged_attn_decode_static_bf16(q.data_ptr(), H * hd, kc.data_ptr(), vc.data_ptr(), pt.data_ptr(),
                                                       n_pages, lens.data_ptr(), od.data_ptr(), 1, H, KV, hd, page, 3,
                                                       hd ** -0.5, 0, torch.cuda.current_stream(dev).cuda_stream)
    _lib.check("dli_paged_attn_decode_static_bf16", rc)                               # per-device symbol address
    out["decode_attn_static"] = od
    torch.cuda.synchronize(dev)
    return {k: v.float().cpu() for k, v in out.items()}


def test_every_kernel_family_on_a_second_device_in_one_process(cuda):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs in one process")
    a = _run_all(torch.device("cuda:0"))
    b = _run_all(torch.device("cuda:1"))  # same seeds, other device: same bits
    for k in a:
        assert torch.isfinite(b[k]).all(), k
        assert torch.equal(a[k], b[k]), k
    torch.cuda.set_device(0)
