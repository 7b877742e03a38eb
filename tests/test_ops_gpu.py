"""LayerNorm / attention / staging kernels through the C ABI vs fp32 torch references on the same (bf16-rounded) inputs.

Tolerances (stated per test): these kernels take bf16 operands and accumulate in fp32; outputs stored as bf16 carry one
rounding (2^-9 relative), so L2 rel-err bounds of 1e-2 (bf16 outputs) / 1e-3 (fp32 outputs) are used.
"""
import math

import pytest
import torch

pytestmark = pytest.mark.gpu


def rel_err(a, b):
    a, b = a.float(), b.float()
    return ((a - b).norm() / b.norm().clamp_min(1e-20)).item()


@pytest.mark.parametrize("M,H", [(60, 128), (1000, 256), (4097, 768), (33, 1024)])
def test_layernorm_fwd_bwd(cuda_device, M, H):
    from mtrl_b200 import _lib

    torch.manual_seed(0)
    x = torch.randn(M, H, device=cuda_device) * 2 + 0.5
    gamma = torch.randn(H, device=cuda_device)
    beta = torch.randn(H, device=cuda_device)
    y = torch.empty(M, H, device=cuda_device, dtype=torch.bfloat16)
    mean = torch.empty(M, device=cuda_device)
    rstd = torch.empty(M, device=cuda_device)
    _lib.layernorm_fwd(x, gamma, beta, 1e-5, y, mean, rstd)
    xr = x.clone().requires_grad_(True)
    gr = gamma.clone().requires_grad_(True)
    br = beta.clone().requires_grad_(True)
    ref = torch.nn.functional.layer_norm(xr, (H,), gr, br, 1e-5)
    assert rel_err(y, ref) < 5e-3
    assert rel_err(mean, x.mean(1)) < 1e-5
    dy = torch.randn(M, H, device=cuda_device).to(torch.bfloat16)
    dres = torch.randn(M, H, device=cuda_device)
    ref.backward(dy.float())
    dx = torch.empty(M, H, device=cuda_device)
    dxb = torch.empty(M, H, device=cuda_device, dtype=torch.bfloat16)
    dg = torch.empty(H, device=cuda_device)
    db = torch.empty(H, device=cuda_device)
    _lib.layernorm_bwd(dy, x, mean, rstd, gamma, dres, dx, dxb, dg, db)
    torch.cuda.synchronize()
    assert rel_err(dx, xr.grad + dres) < 1e-4
    assert rel_err(dxb, xr.grad + dres) < 5e-3
    assert rel_err(dg, gr.grad) < 1e-4
    assert rel_err(db, br.grad) < 1e-4
    # same pass also reducing the column sums of dx (the bias gradient of the c_proj whose output gradient dx is); bit-reproducible
    outs = []
    for _ in range(2):
        dx2, dxb2, dg2, db2, dc = torch.empty_like(dx), torch.empty_like(dxb), torch.empty_like(dg), torch.empty_like(db), torch.empty(H, device=cuda_device)
        _lib.layernorm_bwd(dy, x, mean, rstd, gamma, dres, dx2, dxb2, dg2, db2, dcolsum=dc)
        torch.cuda.synchronize()
        assert torch.equal(dx2, dx) and torch.equal(dxb2, dxb)
        assert rel_err(dg2, dg) < 1e-5 and rel_err(db2, db) < 1e-5  # (the grid, hence the partial-sum grouping, may differ between the two forms)
        assert rel_err(dc, (xr.grad + dres).sum(0)) < 1e-4
        outs.append(dc)
    assert torch.equal(outs[0], outs[1])


def _attn_ref(qkv, mask, nh):
    B, T, H3 = qkv.shape
    H = H3 // 3
    d = H // nh
    q, k, v = qkv.float().split(H, dim=2)
    q = q.view(B, T, nh, d).transpose(1, 2)
    k = k.view(B, T, nh, d).transpose(1, 2)
    v = v.view(B, T, nh, d).transpose(1, 2)
    s = q @ k.transpose(-1, -2) / math.sqrt(d)
    allowed = torch.ones(T, T, dtype=torch.bool, device=qkv.device).tril()[None, None]
    if mask is not None:
        m3 = mask.repeat_interleave(3, dim=1)  # token j uses mask[b, j // 3]
        allowed = allowed & (m3[:, None, None, :] > 0)
    s = s.masked_fill(~allowed, float("-inf"))
    p = torch.softmax(s, dim=-1)
    p = torch.nan_to_num(p, nan=0.0)  # fully masked rows -> 0 output (documented don't-care rows)
    return (p @ v).transpose(1, 2).reshape(B, T, H)


@pytest.mark.parametrize("B,K,nh,d,pad", [(4, 20, 1, 128, True), (3, 30, 1, 128, False), (2, 50, 8, 32, True), (2, 200, 12, 64, True),
                                          (5, 7, 2, 64, True), (3, 64, 4, 64, True), (2, 43, 2, 64, False), (1, 1, 1, 64, False),
                                          (2, 86, 3, 64, True), (2, 100, 2, 128, True), (1, 43, 1, 128, False), (2, 150, 3, 32, True),
                                          (3, 1, 2, 32, False), (2, 60, 1, 32, True), (1, 300, 2, 128, True)])
def test_attention_fwd_bwd(cuda_device, B, K, nh, d, pad):
    from mtrl_b200 import _lib

    T, H = 3 * K, nh * d
    g = torch.Generator().manual_seed(B * 1000 + K)
    qkv = torch.randn(B, T, 3 * H, generator=g).to(cuda_device).to(torch.bfloat16)
    mask = torch.ones(B, K, device=cuda_device)
    if pad:
        for b in range(B):
            mask[b, : (b * 3) % (K // 2 + 1)] = 0
    out = torch.full((B, T, H), float("nan"), device=cuda_device, dtype=torch.bfloat16)
    lse = torch.empty(B, nh, T, device=cuda_device)
    _lib.attention_fwd(qkv, mask, nh, out, lse)
    torch.cuda.synchronize()
    qr = qkv.float().requires_grad_(True)
    ref = _attn_ref(qr, mask, nh)
    valid = mask.repeat_interleave(3, dim=1) > 0  # [B, T] rows that matter
    assert torch.isfinite(out.float()).all()
    assert rel_err(out[valid], ref[valid]) < 1e-2
    # rows at padded positions are fully masked -> exact zeros
    assert (out[~valid].float() == 0).all()

    dout = torch.randn(B, T, H, generator=g).to(cuda_device).to(torch.bfloat16)
    dout[~valid] = 0  # no gradient ever reaches padded rows (SURVEY.md §8a semantics)
    ref.backward(dout.float())
    dqkv = torch.full((B, T, 3 * H), float("nan"), device=cuda_device, dtype=torch.bfloat16)
    delta = torch.empty(B, nh, T, device=cuda_device)
    _lib.attention_bwd(qkv, mask, out, dout, lse, nh, dqkv, delta)
    torch.cuda.synchronize()
    assert torch.isfinite(dqkv.float()).all()
    for i, name in enumerate("qkv"):
        e = rel_err(dqkv[..., i * H:(i + 1) * H], qr.grad[..., i * H:(i + 1) * H])
        assert e < 2e-2, (name, e)


def test_attention_no_mask(cuda_device):
    from mtrl_b200 import _lib

    B, T, nh, d = 2, 100, 2, 64
    qkv = torch.randn(B, T, 3 * nh * d, device=cuda_device).to(torch.bfloat16)
    out = torch.empty(B, T, nh * d, device=cuda_device, dtype=torch.bfloat16)
    lse = torch.empty(B, nh, T, device=cuda_device)
    _lib.attention_fwd(qkv, None, nh, out, lse)
    assert rel_err(out, _attn_ref(qkv, None, nh)) < 1e-2


def test_cast_transpose_and_colsum(cuda_device):
    from mtrl_b200 import _lib

    w = torch.randn(300, 520, device=cuda_device)
    d = torch.empty(300, 520, device=cuda_device, dtype=torch.bfloat16)
    dt = torch.empty(520, 300, device=cuda_device, dtype=torch.bfloat16)
    _lib.cast_transpose(w, d, dt)
    assert torch.equal(d, w.to(torch.bfloat16))  # bit-exact cast
    assert torch.equal(dt, w.t().contiguous().to(torch.bfloat16))
    x = torch.randn(5000, 768, device=cuda_device).to(torch.bfloat16)
    out = torch.ones(768, device=cuda_device)
    _lib.colsum_bf16(x, out, accumulate=True)
    assert rel_err(out, x.float().sum(0) + 1) < 1e-5
