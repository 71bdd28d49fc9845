"""Kernel-level parity: every C-ABI entry point against a plain PyTorch fp32
restatement of the same op (floating-point kernels; tolerance stated per test).
Integer/index work (embedding gather, KV-cache placement, argmax) is bit-exact.
"""
import math

import pytest
import torch

pytestmark = pytest.mark.gpu

BF = torch.bfloat16


def _ulp_close(a, b, max_ulp=1, min_exact=0.97):
    """bf16 tensors equal up to `max_ulp` bf16 ulps, and mostly bit-identical."""
    a32, b32 = a.float(), b.float()
    ulp = torch.maximum(a32.abs(), b32.abs()).clamp_min(1e-30) * 2.0 ** -7
    bad = (a32 - b32).abs() > max_ulp * ulp + 1e-30
    assert not bad.any(), f"{int(bad.sum())} elements differ by more than {max_ulp} bf16 ulp"
    exact = (a32 == b32).float().mean().item()
    assert exact >= min_exact, f"only {exact:.4f} bit-identical"


# ----------------------------------------------------------------------------- norms
@pytest.mark.parametrize("rows,d", [(1, 4096), (7, 2048), (64, 8192), (3, 768), (5, 64)])
@pytest.mark.parametrize("with_res", [False, True])
def test_add_rmsnorm(cuda, rows, d, with_res):
    from dli_b200 import ops

    g = torch.Generator(device="cuda").manual_seed(rows * 131 + d)
    x = torch.randn(rows, d, device=cuda, generator=g).to(BF)
    res = torch.randn(rows, d, device=cuda, generator=g).to(BF) if with_res else None
    w = (1 + 0.1 * torch.randn(d, device=cuda, generator=g)).to(BF)
    eps = 1e-5
    h_out = torch.empty_like(x)
    y = ops.add_rmsnorm(x, w, eps, residual=res, resid_out=h_out)
    # HF LlamaRMSNorm (modeling_llama.py:62-67) on h = residual + x
    h = x if res is None else (res + x)
    hf = h.float()
    ref = w * (hf * torch.rsqrt(hf.pow(2).mean(-1, keepdim=True) + eps)).to(BF)
    assert torch.equal(h_out, h)
    _ulp_close(y, ref)


@pytest.mark.parametrize("d", [2048, 4096, 8192])
@pytest.mark.parametrize("with_res", [False, True])
def test_rmsnorm_of_many_rows_has_the_bits_of_the_row_by_row_kernel(cuda, d, with_res):
    """>= 64 rows take the warp-per-row kernel (prompt passes), fewer the CTA-per-row kernel (decode steps): a token must
    be normalised to the same bits either way, so the two kernels add the squares in the same order."""
    from dli_b200 import ops

    g = torch.Generator(device="cuda").manual_seed(d + with_res)
    rows = 203
    x = (torch.randn(rows, d, device=cuda, generator=g) * 3).to(BF)
    res = torch.randn(rows, d, device=cuda, generator=g).to(BF) if with_res else None
    w = (1 + 0.1 * torch.randn(d, device=cuda, generator=g)).to(BF)
    h_many = torch.empty_like(x)
    y_many = ops.add_rmsnorm(x, w, 1e-5, residual=res, resid_out=h_many)
    for i in range(0, rows, 29):
        sl = slice(i, min(i + 29, rows))
        h_few = torch.empty_like(x[sl])
        y_few = ops.add_rmsnorm(x[sl].contiguous(), w, 1e-5, residual=None if res is None else res[sl].contiguous(), resid_out=h_few)
        assert torch.equal(y_many[sl], y_few) and torch.equal(h_many[sl], h_few)
    hf = (x if res is None else res + x).float()
    ref = w * (hf * torch.rsqrt(hf.pow(2).mean(-1, keepdim=True) + 1e-5)).to(BF)
    _ulp_close(y_many, ref)


@pytest.mark.parametrize("rows,d", [(1, 768), (9, 1024), (4, 4096)])
def test_add_layernorm(cuda, rows, d):
    from dli_b200 import ops

    g = torch.Generator(device="cuda").manual_seed(d)
    x = torch.randn(rows, d, device=cuda, generator=g).to(BF)
    res = torch.randn(rows, d, device=cuda, generator=g).to(BF)
    w = (1 + 0.1 * torch.randn(d, device=cuda, generator=g)).to(BF)
    b = (0.1 * torch.randn(d, device=cuda, generator=g)).to(BF)
    h_out = torch.empty_like(x)
    y = ops.add_layernorm(x, w, b, 1e-5, residual=res, resid_out=h_out)
    h = res + x
    ref = torch.nn.functional.layer_norm(h.float(), (d,), w.float(), b.float(), 1e-5).to(BF)
    assert torch.equal(h_out, h)
    _ulp_close(y, ref)


# ----------------------------------------------------------------------------- embedding / argmax
def test_embedding_exact(cuda):
    from dli_b200 import ops

    V, d = 1000, 256
    table = torch.randn(V, d, device=cuda).to(BF)
    pos_t = torch.randn(64, d, device=cuda).to(BF)
    ids = torch.randint(0, V, (37,), device=cuda, dtype=torch.int32)
    pos = torch.randint(0, 64, (37,), device=cuda, dtype=torch.int32)
    assert torch.equal(ops.embedding(ids, table), table[ids.long()])
    assert torch.equal(ops.embedding(ids, table, positions=pos, pos_table=pos_t), table[ids.long()] + pos_t[pos.long()])
    assert ops.embedding(ids[:0], table).shape == (0, d)


@pytest.mark.parametrize("B,V", [(1, 128256), (32, 32000), (3, 50257), (2, 17)])
def test_argmax_exact_with_ties(cuda, B, V):
    from dli_b200 import ops

    logits = torch.randn(B, V, device=cuda).to(BF)
    # force exact ties of the maximum at several positions: lowest index must win
    top = logits.float().max().item() + 1.0
    for b in range(B):
        idx = torch.randint(0, V, (3,), device=cuda)
        logits[b, idx] = top
    out = ops.argmax(logits)
    ref = torch.stack([(logits[b].float() == logits[b].float().max()).nonzero()[0, 0] for b in range(B)])
    assert torch.equal(out.long(), ref)
    # strided rows (logits living inside a wider buffer)
    if V % 8 == 0:
        wide = torch.full((B, V + 64), -1e9, device=cuda).to(BF)
        wide[:, :V] = logits
        assert torch.equal(ops.argmax(wide[:, :V]).long(), ref)


# ----------------------------------------------------------------------------- rope + kv append
def _hf_rope_tables(hd, theta, max_pos, device):
    # LlamaRotaryEmbedding (modeling_llama.py:73-135): fp32 angles, cos/sin cast to bf16
    inv_freq = 1.0 / (theta ** (torch.arange(0, hd, 2, dtype=torch.int64).float() / hd))
    t = torch.arange(max_pos, dtype=torch.float32)
    freqs = torch.outer(t, inv_freq)
    return freqs.cos().to(BF).to(device), freqs.sin().to(BF).to(device)


def _rotate_half(x):
    x1, x2 = x[..., : x.shape[-1] // 2], x[..., x.shape[-1] // 2:]
    return torch.cat((-x2, x1), dim=-1)


@pytest.mark.parametrize("H,KV,hd,page", [(32, 8, 128, 16), (32, 4, 64, 32), (4, 4, 64, 8)])
def test_rope_kv_append_exact(cuda, H, KV, hd, page):
    from dli_b200 import ops

    torch.manual_seed(0)
    B, T, max_pos = 3, 21, 64
    rows = B * T
    qkv = torch.randn(rows, (H + 2 * KV) * hd, device=cuda).to(BF)
    cos, sin = _hf_rope_tables(hd, 10000.0, max_pos, cuda)
    start = torch.tensor([0, 5, 40], device=cuda)
    positions = (start[:, None] + torch.arange(T, device=cuda)[None]).reshape(-1).int()
    seq_ids = torch.arange(B, device=cuda).repeat_interleave(T).int()
    max_pages = max_pos // page
    perm = torch.randperm(B * max_pages, device=cuda).int()
    page_table = perm.reshape(B, max_pages).contiguous()
    kc = torch.zeros(B * max_pages, KV, page, hd, device=cuda, dtype=BF)
    vc = torch.zeros_like(kc)
    ref_qkv = qkv.clone()
    ops.rope_kv_append(qkv, cos, sin, positions, seq_ids, page_table, kc, vc, H, KV, hd)

    # HF apply_rotary_pos_emb (modeling_llama.py:138-168), bf16 op by op
    q = ref_qkv[:, : H * hd].reshape(rows, H, hd)
    k = ref_qkv[:, H * hd: (H + KV) * hd].reshape(rows, KV, hd)
    v = ref_qkv[:, (H + KV) * hd:].reshape(rows, KV, hd)
    c = torch.cat((cos, cos), -1)[positions.long()][:, None, :]
    s = torch.cat((sin, sin), -1)[positions.long()][:, None, :]
    q_ref = q * c + _rotate_half(q) * s
    k_ref = k * c + _rotate_half(k) * s
    assert torch.equal(qkv[:, : H * hd].reshape(rows, H, hd), q_ref)
    assert torch.equal(qkv[:, H * hd: (H + KV) * hd].reshape(rows, KV, hd), k_ref)
    assert torch.equal(qkv[:, (H + KV) * hd:], ref_qkv[:, (H + KV) * hd:])
    for r in range(rows):
        p, b = int(positions[r]), int(seq_ids[r])
        pg, off = int(page_table[b, p // page]), p % page
        assert torch.equal(kc[pg, :, off, :], k_ref[r])
        assert torch.equal(vc[pg, :, off, :], v[r])


@pytest.mark.parametrize("B,T,H,KV,K", [(1, 2048, 32, 8, 4096), (3, 101, 4, 2, 256), (2, 130, 64, 8, 512)])
def test_qkv_projection_with_rope_and_append_in_its_epilogue_equals_the_three_steps(cuda, B, T, H, KV, K):
    """dli_gemm_qkv_rope_append_bf16 (prompt passes) == dli_gemm_bf16 + dli_rope_kv_append_bf16, bit for bit: the packed
    qkv matrix (q, k rotated; ragged last row tile) and every cache row written."""
    from dli_b200 import ops

    hd, page = 128, 16
    g = torch.Generator(device="cuda").manual_seed(B * 1000 + T)
    rows, max_pos = B * T, T + 40
    x = torch.randn(rows, K, device=cuda, generator=g).to(BF)
    w = (torch.randn((H + 2 * KV) * hd, K, device=cuda, generator=g) * 0.05).to(BF)
    cos, sin = _hf_rope_tables(hd, 500000.0, max_pos, cuda)
    start = torch.randint(0, 40, (B,), device=cuda, generator=g)
    positions = (start[:, None] + torch.arange(T, device=cuda)[None]).reshape(-1).int()
    seq_ids = torch.arange(B, device=cuda).repeat_interleave(T).int()
    max_pages = (max_pos + page - 1) // page
    page_table = torch.randperm(B * max_pages, device=cuda, generator=g).int().reshape(B, max_pages).contiguous()
    assert ops.qkv_rope_append_supported(rows, K, H, KV, hd, None)

    k1 = torch.zeros(B * max_pages, KV, page, hd, device=cuda, dtype=BF)
    v1 = torch.zeros_like(k1)
    qkv1 = ops.linear(x, w)
    ops.rope_kv_append(qkv1, cos, sin, positions, seq_ids, page_table, k1, v1, H, KV, hd)

    k2, v2 = torch.zeros_like(k1), torch.zeros_like(k1)
    qkv2 = torch.empty_like(qkv1)
    ops.linear_qkv_rope_append(x, w, qkv2, cos, sin, positions, seq_ids, page_table, k2, v2, H, KV, hd)
    assert torch.equal(qkv1, qkv2)
    assert torch.equal(k1, k2) and torch.equal(v1, v2)
    assert k1.abs().sum() > 0 and not ops.qkv_rope_append_supported(64, K, H, KV, hd, None)


def test_kv_append_without_rope(cuda):
    from dli_b200 import ops

    H = KV = 4
    hd, page = 64, 16
    qkv = torch.randn(5, 3 * H * hd, device=cuda).to(BF)
    before = qkv.clone()
    positions = torch.arange(5, device=cuda).int()
    seq_ids = torch.zeros(5, device=cuda).int()
    page_table = torch.tensor([[1, 0]], device=cuda).int()
    kc = torch.zeros(2, KV, page, hd, device=cuda, dtype=BF)
    vc = torch.zeros_like(kc)
    ops.rope_kv_append(qkv, None, None, positions, seq_ids, page_table, kc, vc, H, KV, hd)
    assert torch.equal(qkv, before)
    assert torch.equal(kc[1, :, :5].transpose(0, 1).reshape(5, -1), before[:, H * hd: 2 * H * hd])
    assert torch.equal(vc[1, :, :5].transpose(0, 1).reshape(5, -1), before[:, 2 * H * hd:])


# ----------------------------------------------------------------------------- attention
def _paged_cache(B, S_max, KV, hd, page, device, gen):
    max_pages = (S_max + page - 1) // page
    n_pages = B * max_pages + 3
    perm = torch.randperm(n_pages, device=device, generator=gen)[: B * max_pages].int()
    page_table = perm.reshape(B, max_pages).contiguous()
    # poison unused slots with NaN-free but huge values so leaks show up
    kc = torch.randn(n_pages, KV, page, hd, device=device, generator=gen).to(BF)
    vc = torch.randn(n_pages, KV, page, hd, device=device, generator=gen).to(BF)
    return page_table, kc, vc


def _gather_kv(cache, page_table, b, S):
    page = cache.shape[2]
    pages = page_table[b, : (S + page - 1) // page].long()
    x = cache[pages]  # [np, KV, page, hd]
    return x.permute(1, 0, 2, 3).reshape(cache.shape[1], -1, cache.shape[3])[:, :S]  # [KV, S, hd]


@pytest.mark.parametrize(
    "B,H,KV,hd,page,lens,splits,window",
    [
        (1, 32, 8, 128, 16, [2048], 16, 0),
        (1, 32, 8, 128, 16, [2048], 1, 0),
        (4, 32, 8, 128, 32, [1, 63, 64, 777], 4, 0),
        (3, 32, 4, 64, 16, [130, 5, 300], 3, 0),
        (2, 64, 8, 128, 16, [1000, 513], 8, 0),
        (2, 32, 8, 128, 16, [900, 200], 4, 256),
        (2, 12, 12, 64, 16, [40, 97], 2, 0),
        (1, 64, 8, 128, 16, [8195], 33, 0),        # 70B heads at an 8k context (BASELINE configs[4])
        (4, 64, 8, 128, 16, [8192, 4097, 1, 6000], 9, 0),
        (2, 32, 8, 128, 16, [5000, 4097], 8, 4096),  # Mistral window once the context exceeds it
        (40, 32, 8, 128, 16, [2048 - 7 * i for i in range(40)], 1, 0),  # many sequences: the 3-stage path
    ],
)
def test_paged_attn_decode(cuda, B, H, KV, hd, page, lens, splits, window):
    from dli_b200 import ops

    gen = torch.Generator(device="cuda").manual_seed(1234 + B + H)
    S_max = max(lens)
    page_table, kc, vc = _paged_cache(B, S_max, KV, hd, page, cuda, gen)
    q = torch.randn(B, H * hd, device=cuda, generator=gen).to(BF)
    seq_lens = torch.tensor(lens, device=cuda, dtype=torch.int32)
    out = torch.empty(B, H * hd, device=cuda, dtype=BF)
    nbytes = ops.attn_decode_scratch_bytes(B, H, hd, splits)
    scratch = torch.empty(max(nbytes, 4) // 4, device=cuda, dtype=torch.float32)
    scale = hd ** -0.5
    ops.paged_attn_decode(q, H * hd, kc, vc, page_table, seq_lens, out, H, KV, hd, scale,
                          num_splits=splits, scratch=scratch, window=window)
    G = H // KV
    for b in range(B):
        S = lens[b]
        k = _gather_kv(kc, page_table, b, S).float()  # [KV, S, hd]
        v = _gather_kv(vc, page_table, b, S).float()
        if window and S > window:
            k, v = k[:, S - window:], v[:, S - window:]
        qb = q[b].float().reshape(KV, G, hd)
        p = torch.softmax(torch.einsum("kgd,ksd->kgs", qb, k) * scale, dim=-1)
        ref = torch.einsum("kgs,ksd->kgd", p, v).reshape(-1)
        # fp32 reference; kernel rounds P to bf16 before P.V and the output to bf16:
        # tolerance 2e-2 absolute on O(1) values (|out| <= max|v| ~ 4)
        torch.testing.assert_close(out[b].float(), ref, rtol=2e-2, atol=2e-2)


@pytest.mark.parametrize(
    "B,T,H,KV,hd,window",
    [(1, 2048, 32, 8, 128, 0), (2, 100, 32, 4, 64, 0), (1, 1, 4, 4, 64, 0), (1, 300, 8, 2, 128, 128),
     (3, 65, 12, 12, 64, 0), (2, 129, 8, 8, 128, 0), (1, 1, 4, 2, 128, 0), (2, 700, 16, 4, 128, 200)],
)
@pytest.mark.parametrize("tensor_core", [False, True])
def test_attn_prefill(cuda, B, T, H, KV, hd, window, tensor_core):
    from dli_b200 import ops

    if tensor_core and hd != 128:
        pytest.skip("the tcgen05 kernel covers head_dim 128")
    gen = torch.Generator(device="cuda").manual_seed(7 + T)
    qkv = torch.randn(B * T, (H + 2 * KV) * hd, device=cuda, generator=gen).to(BF)
    out = torch.empty(B * T, H * hd, device=cuda, dtype=BF)
    scale = hd ** -0.5
    ops.attn_prefill(qkv, out, B, T, H, KV, hd, scale, window=window, tensor_core=tensor_core)
    G = H // KV
    x = qkv.float().reshape(B, T, H + 2 * KV, hd)
    q = x[:, :, :H].permute(0, 2, 1, 3)  # [B,H,T,hd]
    k = x[:, :, H: H + KV].permute(0, 2, 1, 3).repeat_interleave(G, dim=1)
    v = x[:, :, H + KV:].permute(0, 2, 1, 3).repeat_interleave(G, dim=1)
    s = torch.einsum("bhtd,bhsd->bhts", q, k) * scale
    i = torch.arange(T, device=cuda)
    mask = i[None, :] > i[:, None]
    if window:
        mask = mask | (i[None, :] <= i[:, None] - window)
    s = s.masked_fill(mask, float("-inf"))
    ref = torch.einsum("bhts,bhsd->bhtd", torch.softmax(s, -1), v).permute(0, 2, 1, 3).reshape(B * T, H * hd)
    torch.testing.assert_close(out.float(), ref, rtol=2e-2, atol=2e-2)


@pytest.mark.parametrize(
    "B,T,prefix,H,KV,hd,page,window",
    [(1, 512, 1536, 32, 8, 128, 16, 0), (2, 100, 37, 32, 4, 64, 16, 0), (3, 5, 0, 12, 12, 64, 8, 0),
     (2, 64, 64, 8, 2, 128, 32, 0), (1, 300, 500, 8, 8, 128, 16, 256), (2, 1, 130, 4, 4, 64, 16, 0),
     (2, 200, 77, 16, 4, 128, 8, 0), (1, 130, 1000, 8, 8, 128, 128, 0), (3, 129, 0, 8, 2, 128, 16, 0)],
)
def test_attn_prefill_paged_chunk(cuda, B, T, prefix, H, KV, hd, page, window):
    """A chunk of T query rows at positions prefix..prefix+T-1 against the paged cache == the rows
    prefix.. of a full causal attention over prefix+T tokens."""
    from dli_b200 import ops

    gen = torch.Generator(device="cuda").manual_seed(11 + T + prefix)
    S = prefix + T
    page_table, kc, vc = _paged_cache(B, S, KV, hd, page, cuda, gen)
    qkv = torch.randn(B * T, (H + 2 * KV) * hd, device=cuda, generator=gen).to(BF)
    out = torch.empty(B * T, H * hd, device=cuda, dtype=BF)
    scale = hd ** -0.5
    ops.attn_prefill_paged(qkv, out, kc, vc, page_table, B, T, prefix, H, KV, hd, scale, window=window)
    G = H // KV
    q = qkv.float().reshape(B, T, H + 2 * KV, hd)[:, :, :H].permute(0, 2, 1, 3)  # [B,H,T,hd]
    for b in range(B):
        k = _gather_kv(kc, page_table, b, S).float().repeat_interleave(G, dim=0)  # [H,S,hd]
        v = _gather_kv(vc, page_table, b, S).float().repeat_interleave(G, dim=0)
        s = torch.einsum("htd,hsd->hts", q[b], k) * scale
        qpos = prefix + torch.arange(T, device=cuda)
        kpos = torch.arange(S, device=cuda)
        mask = kpos[None, :] > qpos[:, None]
        if window:
            mask = mask | (kpos[None, :] <= qpos[:, None] - window)
        s = s.masked_fill(mask, float("-inf"))
        ref = torch.einsum("hts,hsd->htd", torch.softmax(s, -1), v).permute(1, 0, 2).reshape(T, H * hd)
        torch.testing.assert_close(out[b * T:(b + 1) * T].float(), ref, rtol=2e-2, atol=2e-2)


# ----------------------------------------------------------------------------- projections
def _lin_ref(x, w, bias=None, residual=None, gelu=False):
    acc = x.float() @ w.float().t()
    if bias is not None:
        acc = acc + bias.float()
    y = acc.to(BF)
    if gelu:
        y = torch.nn.functional.gelu(y.float(), approximate="tanh").to(BF)
    if residual is not None:
        y = (residual.float() + y.float()).to(BF)
    return y


def _check_linear(y, ref, K, acc_scale=1.0, operand_ulp=0.0):
    # fp32-accumulated bf16 GEMM vs fp32 reference.  Both round the accumulator to bf16
    # once, so a different summation order can move a value sitting on a rounding
    # boundary by one bf16 ulp (2^-7 relative to the ACCUMULATOR, which after a residual
    # add or a gate*up product can be larger than the final value): rtol = 2 ulp, atol =
    # 1 ulp of the typical accumulator magnitude `acc_scale`*sqrt(K/4096).
    # `operand_ulp`: with a bias / residual of magnitude up to ~8 the rounded intermediate
    # bf16(acc + bias) is that large too, and a 1-ulp flip of it (2^-5 in [4, 8)) survives a
    # cancelling residual add in full, whatever K is.
    torch.testing.assert_close(y.float(), ref.float(), rtol=1.6e-2,
                               atol=max(2e-2, 2.0 ** -7 * 4 * acc_scale) * math.sqrt(K / 4096) + operand_ulp)


@pytest.mark.parametrize("M", [1, 2, 3, 4, 7])
@pytest.mark.parametrize("N,K", [(6144, 4096), (4096, 14336), (2560, 2048), (1000, 264)])
def test_gemv(cuda, M, N, K):
    from dli_b200 import ops

    gen = torch.Generator(device="cuda").manual_seed(M * 7 + N)
    x = torch.randn(M, K, device=cuda, generator=gen).to(BF)
    w = (torch.randn(N, K, device=cuda, generator=gen) * 0.02).to(BF)
    bias = torch.randn(N, device=cuda, generator=gen).to(BF)
    res = torch.randn(M, N, device=cuda, generator=gen).to(BF)
    _check_linear(ops.gemv(x, w), _lin_ref(x, w), K)
    _check_linear(ops.gemv(x, w, bias=bias, residual=res), _lin_ref(x, w, bias, res), K)
    _check_linear(ops.gemv(x, w, bias=bias, epilogue=ops.EPI_GELU_NEW), _lin_ref(x, w, bias, gelu=True), K)


@pytest.mark.parametrize("N", [1, 2, 3, 5, 443, 444, 445, 1777, 4099])
def test_gemv_every_output_written_once_for_ragged_row_counts(cuda, N):
    """The GEMV gives every CTA a contiguous share of the outputs, walked in groups of <= 4 rows whose last
    groups can be short: row counts around the grid size (444 CTAs on a B200) and not divisible by 4."""
    from dli_b200 import ops

    K = 512
    gen = torch.Generator(device="cuda").manual_seed(N)
    x = torch.randn(2, K, device=cuda, generator=gen).to(BF)
    w = (torch.randn(N, K, device=cuda, generator=gen) * 0.05).to(BF)
    y = torch.full((2, N + 8), 77.0, device=cuda, dtype=BF)  # guard columns behind the last row
    ops.gemv(x, w, out=y.view(-1)[: 2 * N].view(2, N))
    _check_linear(y.view(-1)[: 2 * N].view(2, N), _lin_ref(x, w), K)
    assert bool((y.view(-1)[2 * N:] == 77.0).all())


@pytest.mark.parametrize("M", [1, 2, 4])
@pytest.mark.parametrize("N,K,norm", [(128256, 4096, True), (32000, 2048, True), (50257, 768, False), (17, 64, False),
                                      (16032, 4096, True)])
def test_lm_head_fused_argmax_equals_argmax_of_its_own_logits(cuda, M, N, K, norm):
    """dli_gemv_argmax_bf16: logits bit-identical to the unfused GEMV, token == torch.argmax of those bf16
    logits (lowest index on ties), candidate form == (max logit, offset + index)."""
    from dli_b200 import ops

    gen = torch.Generator(device="cuda").manual_seed(M + N)
    h = torch.randn(M, K, device=cuda, generator=gen).to(BF)
    w = (torch.randn(N, K, device=cuda, generator=gen) * 0.02).to(BF)
    ln = (1.0 + 0.1 * torch.randn(K, device=cuda, generator=gen)).to(BF) if norm else None
    ref_logits = ops.rmsnorm_linear(h, ln, 1e-5, w) if norm else ops.gemv(h, w)
    logits = torch.empty(M, N, device=cuda, dtype=BF)
    scratch = torch.empty(ops.gemv_argmax_scratch_bytes(M) // 4, device=cuda, dtype=torch.float32)
    tokens = torch.full((M,), -1, device=cuda, dtype=torch.int32)
    ops.linear_argmax(h, w, logits, scratch, tokens=tokens, ln_w=ln, eps=1e-5)
    assert torch.equal(logits, ref_logits)
    assert torch.equal(tokens.long(), torch.argmax(logits.float(), dim=-1))
    cand = torch.zeros(M, 2, device=cuda, dtype=torch.float32)
    ops.linear_argmax(h, w, logits, scratch, cand=cand, ln_w=ln, eps=1e-5, idx_offset=1000)
    assert torch.equal(cand[:, 0], logits.float().max(dim=-1).values)
    assert torch.equal(cand[:, 1].contiguous().view(torch.int32).long(), torch.argmax(logits.float(), dim=-1) + 1000)


def test_lm_head_fused_argmax_ties_take_the_lowest_index(cuda):
    """Identical weight rows give identical logits: the fused choice must be the first of them, wherever the
    duplicates fall among the CTAs' shares."""
    from dli_b200 import ops

    K, N = 256, 5000
    gen = torch.Generator(device="cuda").manual_seed(3)
    h = torch.randn(1, K, device=cuda, generator=gen).to(BF)
    w = (torch.randn(N, K, device=cuda, generator=gen) * 0.02).to(BF)
    best = (h.float() @ w.float().t()).argmax().item()
    for dup in (0, 7, 2499, 2500, 4999):
        w2 = w.clone()
        w2[dup] = w[best]
        logits = torch.empty(1, N, device=cuda, dtype=BF)
        scratch = torch.empty(ops.gemv_argmax_scratch_bytes(1) // 4, device=cuda, dtype=torch.float32)
        tokens = torch.zeros(1, device=cuda, dtype=torch.int32)
        ops.linear_argmax(h, w2, logits, scratch, tokens=tokens)
        assert tokens.item() == min(dup, best) == torch.argmax(logits.float(), dim=-1).item()


def _gated_ref(x, wg, wu):
    g = (x.float() @ wg.float().t()).to(BF)
    u = (x.float() @ wu.float().t()).to(BF)
    return torch.nn.functional.silu(g) * u  # bf16 ops, HF rounding points


@pytest.mark.parametrize("M,F,K", [(1, 14336, 4096), (3, 5632, 2048), (4, 256, 128)])
def test_gemv_gated_silu(cuda, M, F, K):
    from dli_b200 import ops

    gen = torch.Generator(device="cuda").manual_seed(F)
    x = torch.randn(M, K, device=cuda, generator=gen).to(BF)
    wg = (torch.randn(F, K, device=cuda, generator=gen) * 0.03).to(BF)
    wu = (torch.randn(F, K, device=cuda, generator=gen) * 0.03).to(BF)
    y = ops.gemv_gated_silu(x, ops.pack_gate_up(wg, wu))
    torch.testing.assert_close(y.float(), _gated_ref(x, wg, wu).float(), rtol=3.2e-2, atol=2e-2 * math.sqrt(K / 4096))


@pytest.mark.parametrize(
    "M,N,K",
    [(2048, 6144, 4096), (2048, 4096, 4096), (2048, 4096, 14336), (300, 2560, 2048),  # 2k rows: 240-column tiles (129, 512, 128), (1000, 48, 64),
     (5, 4096, 4096), (16, 1024, 512), (32, 6144, 4096), (64, 2048, 2048), (100, 384, 256), (128, 128256, 256),
     (4741, 1280, 256), (8192, 768, 128),  # > 16 row tiles: grouped tile walk, partial last group
     (2000, 4096, 2048), (1900, 2048, 4096), (300, 4096, 8192)],  # 256-row tiles (one round), ragged last row tile
)
def test_gemm_tcgen05(cuda, M, N, K):
    from dli_b200 import ops

    gen = torch.Generator(device="cuda").manual_seed(M + N + K)
    x = torch.randn(M, K, device=cuda, generator=gen).to(BF)
    w = (torch.randn(N, K, device=cuda, generator=gen) * 0.02).to(BF)
    bias = torch.randn(N, device=cuda, generator=gen).to(BF)
    res = torch.randn(M, N, device=cuda, generator=gen).to(BF)
    _check_linear(ops.gemm(x, w), _lin_ref(x, w), K)
    _check_linear(ops.gemm(x, w, bias=bias, residual=res), _lin_ref(x, w, bias, res), K, operand_ulp=2.0 ** -5)
    _check_linear(ops.gemm(x, w, bias=bias, epilogue=ops.EPI_GELU_NEW), _lin_ref(x, w, bias, gelu=True), K,
                  operand_ulp=2.0 ** -5)


@pytest.mark.parametrize("M,F,K", [(2048, 14336, 4096), (200, 5632, 2048), (32, 14336, 4096), (8, 256, 128), (130, 128, 64),
                                   (4741, 640, 256)])
def test_gemm_gated_silu_tcgen05(cuda, M, F, K):
    from dli_b200 import ops

    gen = torch.Generator(device="cuda").manual_seed(M + F)
    x = torch.randn(M, K, device=cuda, generator=gen).to(BF)
    wg = (torch.randn(F, K, device=cuda, generator=gen) * 0.03).to(BF)
    wu = (torch.randn(F, K, device=cuda, generator=gen) * 0.03).to(BF)
    y = ops.gemm_gated_silu(x, ops.pack_gate_up(wg, wu))
    # silu(g)*u: a 1-ulp move of g or u scales with the other factor (|g|,|u| up to ~8 here)
    torch.testing.assert_close(y.float(), _gated_ref(x, wg, wu).float(), rtol=3.2e-2, atol=2e-2 * math.sqrt(K / 4096))


@pytest.mark.parametrize("M", [1, 3, 4, 8, 9, 13, 16])
@pytest.mark.parametrize("N,K", [(6144, 4096), (4096, 14336), (2560, 2048), (1000, 264), (50257, 768)])
def test_gemv_mma(cuda, M, N, K):
    """mma.sync weight stream (3..16 rows): same contract as the GEMV, incl. ragged N and K."""
    from dli_b200 import ops

    gen = torch.Generator(device="cuda").manual_seed(M * 7 + N + K)
    x = torch.randn(M, K, device=cuda, generator=gen).to(BF)
    w = (torch.randn(N, K, device=cuda, generator=gen) * 0.02).to(BF)
    bias = torch.randn(N, device=cuda, generator=gen).to(BF)
    res = torch.randn(M, N, device=cuda, generator=gen).to(BF)
    _check_linear(ops.gemv_mma(x, w), _lin_ref(x, w), K)
    _check_linear(ops.gemv_mma(x, w, bias=bias, residual=res), _lin_ref(x, w, bias, res), K, operand_ulp=2.0 ** -5)
    _check_linear(ops.gemv_mma(x, w, bias=bias, epilogue=ops.EPI_GELU_NEW), _lin_ref(x, w, bias, gelu=True), K,
                  operand_ulp=2.0 ** -5)
    # in-place residual (y aliases residual), as the engine uses it
    h = res.clone()
    ops.gemv_mma(x, w, residual=h, out=h)
    _check_linear(h, _lin_ref(x, w, None, res), K, operand_ulp=2.0 ** -5)
    # deterministic
    assert torch.equal(ops.gemv_mma(x, w), ops.gemv_mma(x, w))


@pytest.mark.parametrize("M,F,K", [(3, 14336, 4096), (8, 5632, 2048), (16, 256, 128), (11, 28672, 8192)])
def test_gemv_mma_gated_silu(cuda, M, F, K):
    from dli_b200 import ops

    gen = torch.Generator(device="cuda").manual_seed(F + M)
    x = torch.randn(M, K, device=cuda, generator=gen).to(BF)
    wg = (torch.randn(F, K, device=cuda, generator=gen) * 0.03).to(BF)
    wu = (torch.randn(F, K, device=cuda, generator=gen) * 0.03).to(BF)
    y = ops.gemv_mma_gated_silu(x, ops.pack_gate_up(wg, wu))
    torch.testing.assert_close(y.float(), _gated_ref(x, wg, wu).float(), rtol=3.2e-2, atol=2e-2 * math.sqrt(K / 4096))


def test_gemv_matches_gemm_closely(cuda):
    """The two projection paths agree to bf16 rounding on the same inputs."""
    from dli_b200 import ops

    x = torch.randn(4, 4096, device=cuda).to(BF)
    w = (torch.randn(4096, 4096, device=cuda) * 0.02).to(BF)
    a, b = ops.gemv(x, w).float(), ops.gemm(x, w).float()
    assert (a - b).abs().max() <= 2.0 ** -7 * a.abs().max()


def test_bad_arguments_raise(cuda):
    from dli_b200 import _lib, ops

    x = torch.randn(2, 100, device=cuda).to(BF)  # K % 8 != 0
    w = torch.randn(16, 100, device=cuda).to(BF)
    with pytest.raises(_lib.DliError) as e:
        ops.gemv(x, w)
    assert "K % 8" in str(e.value)
    with pytest.raises(ValueError):
        ops.add_rmsnorm(x.cpu(), w[0].cpu(), 1e-5)


# ----------------------------------------------------------------------------- fused decode kernels
@pytest.mark.parametrize("B,H,KV,hd,page,lens,splits,window,rope",
                         [(1, 32, 8, 128, 16, [2049], 16, 0, True), (3, 32, 4, 64, 32, [1, 64, 300], 4, 0, True),
                          (2, 64, 8, 128, 16, [513, 1000], 8, 256, True), (2, 12, 12, 64, 16, [40, 97], 2, 0, False)])
def test_fused_rope_append_attention_is_bit_identical(cuda, B, H, KV, hd, page, lens, splits, window, rope):
    """dli_paged_attn_decode_fused == dli_rope_kv_append + dli_paged_attn_decode, bit for bit
    (outputs AND the cache pages it wrote)."""
    from dli_b200 import ops

    gen = torch.Generator(device="cuda").manual_seed(99 + B)
    page_table, kc, vc = _paged_cache(B, max(lens), KV, hd, page, cuda, gen)
    qkv = torch.randn(B, (H + 2 * KV) * hd, device=cuda, generator=gen).to(BF)
    cos, sin = _hf_rope_tables(hd, 500000.0, max(lens) + 1, cuda) if rope else (None, None)
    seq_lens = torch.tensor(lens, device=cuda, dtype=torch.int32)
    positions = (seq_lens - 1).contiguous()
    seq_ids = torch.arange(B, device=cuda, dtype=torch.int32)
    nbytes = ops.attn_decode_scratch_bytes(B, H, hd, splits)
    scratch = torch.empty(max(nbytes, 4) // 4, device=cuda, dtype=torch.float32)
    scale = hd ** -0.5
    # reference: two kernels
    qkv_a, kc_a, vc_a = qkv.clone(), kc.clone(), vc.clone()
    ops.rope_kv_append(qkv_a, cos, sin, positions, seq_ids, page_table, kc_a, vc_a, H, KV, hd)
    out_a = torch.empty(B, H * hd, device=cuda, dtype=BF)
    ops.paged_attn_decode(qkv_a, qkv_a.shape[1], kc_a, vc_a, page_table, seq_lens, out_a, H, KV, hd, scale,
                          num_splits=splits, scratch=scratch, window=window)
    # fused
    qkv_b, kc_b, vc_b = qkv.clone(), kc.clone(), vc.clone()
    out_b = torch.empty_like(out_a)
    ops.paged_attn_decode_fused(qkv_b, cos, sin, kc_b, vc_b, page_table, seq_lens, out_b, H, KV, hd, scale,
                                num_splits=splits, scratch=scratch, window=window)
    assert torch.equal(kc_a, kc_b) and torch.equal(vc_a, vc_b)
    assert torch.equal(out_a, out_b)
    assert torch.equal(qkv_b, qkv)  # the fused kernel leaves the qkv row untouched


@pytest.mark.parametrize("M", [1, 2, 4])
@pytest.mark.parametrize("N,K", [(6144, 4096), (2560, 2048), (128256, 4096), (10240, 8192)])
def test_rmsnorm_gemv_is_bit_identical_to_two_kernels(cuda, M, N, K):
    from dli_b200 import ops

    gen = torch.Generator(device="cuda").manual_seed(M + N)
    h = torch.randn(M, K, device=cuda, generator=gen).to(BF)
    lnw = (1 + 0.1 * torch.randn(K, device=cuda, generator=gen)).to(BF)
    w = (torch.randn(N, K, device=cuda, generator=gen) * 0.02).to(BF)
    res = torch.randn(M, N, device=cuda, generator=gen).to(BF)
    two = ops.gemv(ops.add_rmsnorm(h, lnw, 1e-5), w, residual=res)
    one = ops.rmsnorm_linear(h, lnw, 1e-5, w, residual=res)
    assert torch.equal(one, two)


@pytest.mark.parametrize("M,F,K", [(1, 14336, 4096), (3, 5632, 2048)])
def test_rmsnorm_gemv_gated_is_bit_identical_to_two_kernels(cuda, M, F, K):
    from dli_b200 import ops

    gen = torch.Generator(device="cuda").manual_seed(F)
    h = torch.randn(M, K, device=cuda, generator=gen).to(BF)
    lnw = (1 + 0.1 * torch.randn(K, device=cuda, generator=gen)).to(BF)
    wgu = ops.pack_gate_up((torch.randn(F, K, device=cuda, generator=gen) * 0.03).to(BF),
                           (torch.randn(F, K, device=cuda, generator=gen) * 0.03).to(BF))
    two = ops.gemv_gated_silu(ops.add_rmsnorm(h, lnw, 1e-5), wgu)
    one = ops.rmsnorm_linear_gated_silu(h, lnw, 1e-5, wgu)
    assert torch.equal(one, two)


def test_pdl_does_not_change_results(cuda):
    """Programmatic dependent launch only overlaps kernel boundaries: same tokens."""
    from dli_b200 import ops

    from golden_util import load_case, spec_kwargs
    from dli_b200.engine import ModelSpec, ShardEngine
    from oracle import decoder_ref as R

    c = load_case("llama_tiny")
    spec = ModelSpec.from_hf_config(spec_kwargs(c["config"]))
    sd = R.seeded_state_dict(R.ModelSpec(**spec.__dict__), c["seed"], c["std"])
    outs = []
    for on in (False, True, False):
        eng = ShardEngine(spec, max_batch=2, max_seq_len=128)
        ops.set_pdl(on)
        eng.load_state_dict(sd)
        outs.append(eng.generate(c["prompts"], c["prompts"].shape[1] + 20))
    ops.set_pdl(True)
    assert torch.equal(outs[0], outs[1]) and torch.equal(outs[0], outs[2])


# ----------------------------------------------------------------------------- split-K batched decode
@pytest.mark.parametrize("M", [3, 5, 16, 32, 64, 100, 128])
def test_gemm_splitk_matches_reference_and_is_deterministic(cuda, M):
    from dli_b200 import ops

    gen = torch.Generator(device="cuda").manual_seed(M)
    scratch = torch.zeros(ops.gemm_splitk_scratch_bytes(128, 128256), dtype=torch.uint8, device=cuda)
    for (N, K) in [(6144, 4096), (4096, 14336), (2560, 2048), (128256, 256), (48, 64)]:
        x = torch.randn(M, K, device=cuda, generator=gen).to(BF)
        w = (torch.randn(N, K, device=cuda, generator=gen) * 0.02).to(BF)
        bias = torch.randn(N, device=cuda, generator=gen).to(BF)
        res = torch.randn(M, N, device=cuda, generator=gen).to(BF)
        y1 = ops.linear_splitk(x, w, scratch, bias=bias, residual=res)
        y2 = ops.linear_splitk(x, w, scratch, bias=bias, residual=res)  # same scratch, counters self-reset
        assert torch.equal(y1, y2)
        _check_linear(y1, _lin_ref(x, w, bias, res), K)
        _check_linear(ops.linear_splitk(x, w, scratch), _lin_ref(x, w), K)
        # GPT-2's fc projection (bias + gelu_new) and the engine's in-place residual (y aliases residual)
        _check_linear(ops.linear_splitk(x, w, scratch, bias=bias, epilogue=ops.EPI_GELU_NEW),
                      _lin_ref(x, w, bias, gelu=True), K, operand_ulp=2.0 ** -5)
        h = res.clone()
        ops.linear_splitk(x, w, scratch, residual=h, out=h)
        _check_linear(h, _lin_ref(x, w, None, res), K, operand_ulp=2.0 ** -5)
    ops.linear_splitk(x[:, :64].contiguous(), w[:, :64].contiguous(), scratch)
    n_tiles = (w.shape[0] + 127) // 128
    assert int(scratch[: 4 * n_tiles].view(torch.int32).abs().sum()) == 0  # counters are left at zero


@pytest.mark.parametrize("M,F,K", [(32, 14336, 4096), (8, 5632, 2048), (128, 256, 128)])
def test_gemm_splitk_gated(cuda, M, F, K):
    from dli_b200 import ops

    gen = torch.Generator(device="cuda").manual_seed(M + F)
    scratch = torch.zeros(ops.gemm_splitk_scratch_bytes(128, 2 * F), dtype=torch.uint8, device=cuda)
    x = torch.randn(M, K, device=cuda, generator=gen).to(BF)
    wg = (torch.randn(F, K, device=cuda, generator=gen) * 0.03).to(BF)
    wu = (torch.randn(F, K, device=cuda, generator=gen) * 0.03).to(BF)
    y = ops.linear_splitk_gated_silu(x, ops.pack_gate_up(wg, wu), scratch)
    torch.testing.assert_close(y.float(), _gated_ref(x, wg, wu).float(), rtol=3.2e-2, atol=2e-2 * math.sqrt(K / 4096))
    assert torch.equal(y, ops.linear_splitk_gated_silu(x, ops.pack_gate_up(wg, wu), scratch))


def test_batched_decode_uses_splitk_and_matches_small_batch(cuda):
    """Engine at batch 8 (tcgen05 split-K path) produces the tokens each row produces at
    batch 1 (GEMV path) over the margin-safe prefix."""
    from golden_util import load_case, spec_kwargs
    from dli_b200.engine import ModelSpec, ShardEngine
    from oracle import decoder_ref as R

    c = load_case("llama_tiny")
    spec = ModelSpec.from_hf_config(spec_kwargs(c["config"]))
    sd = R.seeded_state_dict(R.ModelSpec(**spec.__dict__), c["seed"], c["std"])
    eng = ShardEngine(spec, max_batch=8, max_seq_len=128)
    eng.load_state_dict(sd)
    T, N = c["prompts"].shape[1], c["tokens"].shape[1]
    prompts = c["prompts"].repeat(4, 1)  # 8 rows: the two fixture prompts, four times
    out = eng.generate(prompts, T + N)
    assert torch.equal(out[:2], out[2:4]) and torch.equal(out[:2], out[6:8])  # rows independent of position
    noise = 2.0 * c["hf_err"].max()
    for b in range(2):
        n_safe = int((c["margin"][:, b] > 2.0 * noise).int().cumprod(0).sum())
        assert torch.equal(out[b, T:T + n_safe], c["tokens"][b, :n_safe])


@pytest.mark.parametrize("B,H,KV,hd,page,lens,splits,window",
                         [(1, 32, 8, 128, 16, [2049], 8, 0), (2, 32, 8, 128, 16, [700, 64], 4, 0),
                          (1, 32, 4, 64, 32, [300], 3, 0), (1, 64, 8, 128, 16, [1500], 8, 512)])
def test_cluster_dsmem_merge_matches_scratch_merge(cuda, B, H, KV, hd, page, lens, splits, window):
    """scratch == NULL: the splits run as one thread-block cluster and merge through distributed
    shared memory; same result as partials-in-HBM + merge kernel within the attention tolerance."""
    from dli_b200 import ops

    gen = torch.Generator(device="cuda").manual_seed(31 + H)
    page_table, kc, vc = _paged_cache(B, max(lens), KV, hd, page, cuda, gen)
    qkv = torch.randn(B, (H + 2 * KV) * hd, device=cuda, generator=gen).to(BF)
    cos, sin = _hf_rope_tables(hd, 10000.0, max(lens) + 1, cuda)
    seq_lens = torch.tensor(lens, device=cuda, dtype=torch.int32)
    scratch = torch.empty(max(ops.attn_decode_scratch_bytes(B, H, hd, splits), 4) // 4, device=cuda, dtype=torch.float32)
    outs = []
    for sc in (scratch, None):
        kc2, vc2 = kc.clone(), vc.clone()
        out = torch.empty(B, H * hd, device=cuda, dtype=BF)
        ops.paged_attn_decode_fused(qkv.clone(), cos, sin, kc2, vc2, page_table, seq_lens, out, H, KV, hd, hd ** -0.5,
                                    num_splits=splits, scratch=sc, window=window)
        outs.append(out)
    # The scratch path now runs the many-warp kernel (16 tokens per warp instead of 64 per tile), so the two paths
    # round their bf16 probabilities against different running maxima: equal within the attention tolerance (2e-2 on
    # O(1) values, what test_paged_attn_decode allows against fp32), not to the ulp.
    torch.testing.assert_close(outs[1].float(), outs[0].float(), rtol=2e-2, atol=2e-2)
    # too many splits for a cluster and no scratch: refused with a clear message
    from dli_b200 import _lib
    with pytest.raises(_lib.DliError, match="cluster merge"):
        ops.paged_attn_decode_fused(qkv.clone(), cos, sin, kc.clone(), vc.clone(), page_table, seq_lens,
                                    torch.empty(B, H * hd, device=cuda, dtype=BF), H, KV, hd, hd ** -0.5,
                                    num_splits=9, scratch=None)
