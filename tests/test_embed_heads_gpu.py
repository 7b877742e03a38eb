"""Fused embedding / heads / masked-MSE kernels through the C ABI.

Bit-exact parts (north star): token interleaving order R_k,s_k,a_k, the int64 timestep gather, row selection of the heads
(state token -> action, action token -> state/return) and the mask>0 row selection of the loss.  They are checked with
integer-valued weights and inputs, for which every fp32 sum is exact, so `==` is the assertion.  The LayerNorm'd output is fp32
arithmetic in a different summation order than torch: tolerance 1e-5 relative.
"""
import pytest
import torch

from util import rel_err

pytestmark = pytest.mark.gpu


def _int(shape, lo, hi, g):
    return torch.randint(lo, hi + 1, shape, generator=g).float()


@pytest.mark.parametrize("B,K,sd,ad,H,E", [(3, 5, 11, 3, 128, 50), (2, 30, 17, 4, 96, 40), (2, 7, 64, 18, 256, 1000), (1, 200, 11, 3, 768, 300)])
def test_embed_interleave_and_gather_bit_exact(cuda_device, B, K, sd, ad, H, E):
    from mtrl_b200 import _lib

    g = torch.Generator().manual_seed(B * 100 + K)
    dev = cuda_device
    states, actions, rtg = _int((B * K, sd), -3, 3, g), _int((B * K, ad), -3, 3, g), _int((B * K,), -4, 4, g)
    ts = torch.randint(0, E, (B * K,), generator=g)
    table = _int((E, H), -8, 8, g)
    wR, bR, Ws, bs, Wa, ba = _int((H,), -2, 2, g), _int((H,), -2, 2, g), _int((H, sd), -2, 2, g), _int((H,), -2, 2, g), _int((H, ad), -2, 2, g), _int((H,), -2, 2, g)
    gamma, beta = torch.rand(H, generator=g) + 0.5, torch.randn(H, generator=g)
    P8 = (1 + ad + sd + 7) // 8 * 8
    x = torch.full((B * K * 3, H), float("nan"), device=dev)
    pre = torch.full((B * K * 3, H), float("nan"), device=dev)
    mean, rstd = torch.empty(B * K * 3, device=dev), torch.empty(B * K * 3, device=dev)
    U = torch.full((B * K, P8), float("nan"), device=dev, dtype=torch.bfloat16)
    d = lambda t: t.to(dev)
    _lib.embed_fwd(d(states), None, d(actions), d(rtg), d(ts), d(table), d(wR), d(bR), d(Ws), d(bs), d(Wa), d(ba), d(gamma), d(beta), 1e-5,
                   x, pre, mean, rstd, U)
    torch.cuda.synchronize()
    te = table[ts]
    e_r = rtg[:, None] * wR[None] + bR + te
    e_s = states @ Ws.t() + bs + te
    e_a = actions @ Wa.t() + ba + te
    ref_pre = torch.stack((e_r, e_s, e_a), dim=1).reshape(B * K * 3, H)  # HF:623-627 order
    assert torch.equal(pre.cpu(), ref_pre)  # bit-exact: gather index, interleave order, integer arithmetic
    ref_x = torch.nn.functional.layer_norm(ref_pre, (H,), gamma, beta, 1e-5)
    assert rel_err(x, ref_x) < 1e-5
    assert rel_err(mean, ref_pre.mean(1)) < 1e-6
    ref_U = torch.zeros(B * K, P8)
    ref_U[:, 0], ref_U[:, 1:1 + ad], ref_U[:, 1 + ad:1 + ad + sd] = rtg, actions, states
    assert torch.equal(U.float().cpu(), ref_U)


def test_embed_with_projected_state(cuda_device):
    """Wide-state path: the state projection arrives from the tensor-core GEMM; the kernel adds bias + timestep embedding + LN."""
    from mtrl_b200 import _lib

    g = torch.Generator().manual_seed(5)
    dev = cuda_device
    BK, ad, H, E, sd = 37, 4, 128, 64, 1024
    proj, actions, rtg = torch.randn(BK, H, generator=g), torch.randn(BK, ad, generator=g), torch.randn(BK, generator=g)
    ts = torch.randint(0, E, (BK,), generator=g)
    table = torch.randn(E, H, generator=g)
    wR, bR, bs, ba = (torch.randn(H, generator=g) for _ in range(4))
    Wa = torch.randn(H, ad, generator=g)
    Ws = torch.zeros(H, sd)
    gamma, beta = torch.rand(H, generator=g) + 0.5, torch.randn(H, generator=g)
    x = torch.empty(BK * 3, H, device=dev)
    d = lambda t: t.to(dev)
    _lib.embed_fwd(None, d(proj), d(actions), d(rtg), d(ts), d(table), d(wR), d(bR), d(Ws), d(bs), d(Wa), d(ba), d(gamma), d(beta), 1e-5, x,
                   u_states=False)
    te = table[ts]
    pre = torch.stack((rtg[:, None] * wR + bR + te, proj + bs + te, actions @ Wa.t() + ba + te), dim=1).reshape(BK * 3, H)
    assert rel_err(x, torch.nn.functional.layer_norm(pre, (H,), gamma, beta, 1e-5)) < 1e-5


@pytest.mark.parametrize("BK,H,sd,ad,tanh", [(40, 128, 11, 3, True), (33, 256, 24, 18, False), (9, 768, 11, 3, True)])
def test_heads_row_selection_and_values(cuda_device, BK, H, sd, ad, tanh):
    from mtrl_b200 import _lib

    g = torch.Generator().manual_seed(BK)
    dev = cuda_device
    # integer-valued y and weights: dot products are exact -> row selection errors cannot hide in rounding
    y = _int((BK * 3, H), -2, 2, g)
    Wa, Ws, Wr = _int((ad, H), -1, 1, g), _int((sd, H), -1, 1, g), _int((H,), -1, 1, g)
    ba, bs, br = _int((ad,), -2, 2, g), _int((sd,), -2, 2, g), _int((1,), -2, 2, g)
    ap = torch.empty(BK, ad, device=dev)
    sp = torch.empty(BK, sd, device=dev)
    rp = torch.empty(BK, device=dev)
    d = lambda t: t.to(dev)
    _lib.heads_fwd(d(y).to(torch.bfloat16), d(Wa), d(ba), d(Ws), d(bs), d(Wr), d(br), tanh, ap, sp, rp)
    torch.cuda.synchronize()
    y3 = y.reshape(BK, 3, H)
    ref_a = y3[:, 1] @ Wa.t() + ba  # HF:655 state token -> action
    ref_s = y3[:, 2] @ Ws.t() + bs  # HF:654 action token -> state
    ref_r = y3[:, 2] @ Wr + br      # HF:653
    assert torch.equal(sp.cpu(), ref_s) and torch.equal(rp.cpu(), ref_r)
    if tanh:
        assert rel_err(ap, torch.tanh(ref_a)) < 1e-6
    else:
        assert torch.equal(ap.cpu(), ref_a)


@pytest.mark.parametrize("BK,ad", [(1280, 3), (77, 18), (5, 4)])
def test_masked_mse(cuda_device, BK, ad):
    from mtrl_b200 import _lib

    g = torch.Generator().manual_seed(ad)
    dev = cuda_device
    pred, target = torch.randn(BK, ad, generator=g), torch.randn(BK, ad, generator=g)
    mask = (torch.rand(BK, generator=g) > 0.3).float()
    mask[0] = 1
    loss = torch.empty(2, device=dev)
    dp = torch.empty(BK, ad, device=dev)
    _lib.masked_mse(pred.to(dev), target.to(dev), mask.to(dev), 0.5, loss, dp)
    torch.cuda.synchronize()
    p = pred.clone().requires_grad_(True)
    ref = torch.mean((p[mask > 0] - target[mask > 0]) ** 2)
    ref.backward()
    assert loss[1].item() == float((mask > 0).sum())  # integer-exact row count
    assert abs(loss[0].item() - ref.item()) < 1e-5 * abs(ref.item())
    assert rel_err(dp, 0.5 * p.grad) < 1e-5
    assert (dp.cpu()[mask == 0] == 0).all()
