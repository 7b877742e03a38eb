"""Whole-model parity: the sm_100a DecisionTransformer (through the C ABI) vs the fp32 oracle on the same seeded inputs and weights.

Tolerances are the north star's (BASELINE.json): bf16 tensor-core compute vs fp32 oracle, L2 relative error
  action_preds <= 1e-2, loss <= 2e-2, every gradient tensor <= 2e-2      (valid rows only; masked rows are don't-care),
with one documented exception for the two tensors behind a ReLU mask (util.grad_tolerances: bounded by torch's own bf16 autocast error).
Integer rules (token interleave, timestep gather, mask indexing, row selection) are checked bit-exactly in test_embed_heads_gpu.py.
"""
import pytest
import torch

from util import CONFIGS, build_pair, grad_tolerances, make_inputs, rel_err

pytestmark = pytest.mark.gpu

TOL_ACTION, TOL_LOSS, TOL_GRAD = 1e-2, 2e-2, 2e-2


def _oracle_step(oracle, s, a, r, t, m):
    from oracle.trainer import masked_action_mse

    oracle.zero_grad()
    sp, ap, rp = oracle(s, a, None, r, t, m)
    loss = masked_action_mse(ap, a, m)
    loss.backward()
    return sp, ap, rp, loss


@pytest.mark.parametrize("name,B", [("c1", 64), ("c2", 8), ("c3", 4), ("c4", 2), ("c2", 64), ("c3", 64), ("c4", 8)])
def test_forward_loss_backward_fused(cuda_device, name, B):
    oracle, model, K = build_pair(name, cuda_device)
    cfg = CONFIGS[name]
    onehot = not cfg["action_tanh"]
    s, a, r, t, m = make_inputs(B, K, cfg["state_dim"], cfg["act_dim"], 1000, seed=1234 + B, onehot_actions=onehot)
    sp_o, ap_o, rp_o, loss_o = _oracle_step(oracle, s, a, r, t, m)
    from oracle.trainer import masked_action_mse

    tol = grad_tolerances(oracle, lambda o: masked_action_mse(o(s, a, None, r, t, m)[1].float(), a, m))
    dev = cuda_device
    out = model.loss_and_backward(s.to(dev), a.to(dev), r.to(dev), t.to(dev), m.to(dev))
    torch.cuda.synchronize()
    loss, nvalid = out[0].item(), out[1].item()
    assert nvalid == float((m > 0).sum())  # integer-exact row selection
    assert abs(loss - loss_o.item()) <= TOL_LOSS * abs(loss_o.item()), (loss, loss_o.item())
    worst = ("", 0.0)
    for n, p in oracle.named_parameters():
        g = model.get_parameter(n).grad
        ref = p.grad if p.grad is not None else torch.zeros_like(p)
        e = rel_err(g, ref)
        if e > worst[1]:
            worst = (n, e)
        assert e <= tol(n), (n, e, tol(n))
    print(f"{name}: loss {loss:.6f} vs {loss_o.item():.6f}; worst grad rel-err {worst}")


@pytest.mark.parametrize("name,B", [("c1", 16), ("c3", 2), ("c4", 1)])
def test_forward_outputs_and_autograd_bridge(cuda_device, name, B):
    """model.forward() returns (state_preds, action_preds, return_preds) like the reference (A8) and loss.backward() fills param.grad."""
    oracle, model, K = build_pair(name, cuda_device, seed=3)
    cfg = CONFIGS[name]
    s, a, r, t, m = make_inputs(B, K, cfg["state_dim"], cfg["act_dim"], 1000, seed=77, onehot_actions=not cfg["action_tanh"])
    dev = cuda_device
    sp, ap, rp = model(s.to(dev), a.to(dev), None, r.to(dev), t.to(dev), m.to(dev))
    sp_o, ap_o, rp_o = oracle(s, a, None, r, t, m)
    keep = m > 0
    assert sp.shape == sp_o.shape and ap.shape == ap_o.shape and rp.shape == rp_o.shape
    assert rel_err(ap[keep.to(dev)], ap_o[keep]) <= TOL_ACTION
    assert rel_err(sp[keep.to(dev)], sp_o[keep]) <= TOL_ACTION
    assert rel_err(rp[keep.to(dev)], rp_o[keep]) <= TOL_ACTION
    # a loss that touches all three heads, through torch autograd
    w = keep.float()

    def total(sp_, ap_, rp_, w_):
        return ((ap_ ** 2).sum(-1) * w_).sum() + 0.5 * ((sp_ ** 2).sum(-1) * w_).sum() / sp_.shape[-1] + 0.3 * (rp_.squeeze(-1) * w_).sum()

    oracle.zero_grad()
    total(sp_o, ap_o, rp_o, w).backward()

    def ac_step(o):
        sp_, ap_, rp_ = o(s, a, None, r, t, m)
        return total(sp_.float(), ap_.float(), rp_.float(), w)

    tol = grad_tolerances(oracle, ac_step)
    model.zero_grad(set_to_none=True)
    total(sp, ap, rp, w.to(dev)).backward()
    torch.cuda.synchronize()
    for n, p in oracle.named_parameters():
        e = rel_err(model.get_parameter(n).grad, p.grad if p.grad is not None else torch.zeros_like(p))
        assert e <= tol(n), (n, e, tol(n))


def test_no_mask_equals_all_ones(cuda_device):
    oracle, model, K = build_pair("c1", cuda_device, seed=5)
    s, a, r, t, m = make_inputs(4, K, 11, 3, 1000, seed=9, pad=False)
    dev = cuda_device
    _, ap1, _ = model(s.to(dev), a.to(dev), None, r.to(dev), t.to(dev), None)
    _, ap2, _ = model(s.to(dev), a.to(dev), None, r.to(dev), t.to(dev), torch.ones(4, K, device=dev, dtype=torch.long))
    assert torch.equal(ap1, ap2)
    _, ap_o, _ = oracle(s, a, None, r, t, None)
    assert rel_err(ap1, ap_o) <= TOL_ACTION


def test_masked_inputs_do_not_leak(cuda_device):
    """Changing states/actions under the mask must not change unmasked predictions at all (bit-exact; SURVEY.md §8a semantics)."""
    _, model, K = build_pair("c1", cuda_device, seed=6)
    s, a, r, t, m = make_inputs(6, K, 11, 3, 1000, seed=10)
    dev = cuda_device
    _, ap1, _ = model(s.to(dev), a.to(dev), None, r.to(dev), t.to(dev), m.to(dev))
    s2, a2 = s.clone(), a.clone()
    s2[m == 0] = 123.0
    a2[m == 0] = -7.0
    _, ap2, _ = model(s2.to(dev), a2.to(dev), None, r.to(dev), t.to(dev), m.to(dev))
    keep = (m > 0).to(dev)
    assert torch.equal(ap1[keep], ap2[keep])


@pytest.mark.parametrize("t_hist", [1, 7, 20, 33])
def test_get_action(cuda_device, t_hist):
    """A9: last-K window, LEFT zero padding, prediction of the last state token."""
    oracle, model, K = build_pair("c1", cuda_device, seed=7)
    g = torch.Generator().manual_seed(t_hist)
    s = torch.randn(t_hist, 11, generator=g)
    a = torch.rand(t_hist, 3, generator=g) * 2 - 1
    a[-1] = 0  # caller appends a zero placeholder for the current step
    r = torch.rand(t_hist, 1, generator=g) * 3
    ts = torch.arange(t_hist)
    dev = cuda_device
    got = model.get_action(s.to(dev), a.to(dev), None, r.to(dev), ts.to(dev))
    ref = oracle.get_action(s, a, None, r, ts)
    assert got.shape == ref.shape == (3,)
    assert rel_err(got, ref) <= TOL_ACTION


def test_trainer_train_step_matches_oracle(cuda_device):
    """SequenceTrainer.train_step(): same loss trajectory as the oracle trainer over a few AdamW steps (fused path and loss_fn path)."""
    from mtrl_b200.trainer import SequenceTrainer
    from oracle.trainer import SequenceTrainer as OracleTrainer

    dev = cuda_device
    for use_loss_fn in (False, True):
        oracle, model, K = build_pair("c1", dev, seed=11)
        batches = [make_inputs(16, K, 11, 3, 1000, seed=100 + i) for i in range(4)]

        def mk(device):
            it = iter(batches)

            def get_batch(bs):
                s, a, r, t, m = next(it)
                rtg = torch.cat([r, r[:, -1:]], dim=1)  # K+1 entries, as upstream get_batch
                z = torch.zeros(s.shape[0], K, 1)
                return tuple(x.to(device) for x in (s, a, z, z, rtg, t, m))
            return get_batch

        opt_o = torch.optim.AdamW(oracle.parameters(), lr=1e-3, weight_decay=1e-4)
        opt_m = torch.optim.AdamW(model.parameters(), lr=1e-3, weight_decay=1e-4)
        loss_fn = (lambda s_hat, a_hat, r_hat, s, a, r: torch.mean((a_hat - a) ** 2)) if use_loss_fn else None
        tr_o = OracleTrainer(oracle, opt_o, 16, mk("cpu"))
        tr_m = SequenceTrainer(model, opt_m, 16, mk(dev), loss_fn=loss_fn)
        for _ in range(4):
            lo, lm = tr_o.train_step(), tr_m.train_step()
            assert abs(lo - lm) <= 3e-2 * abs(lo), (use_loss_fn, lo, lm)


def test_cpu_tensor_raises(cuda_device):
    from mtrl_b200 import _lib

    _, model, K = build_pair("c1", cuda_device)
    s, a, r, t, m = make_inputs(2, K, 11, 3, 1000, seed=1)
    with pytest.raises(_lib.DtbError):
        model(s, a, None, r, t, m)


def test_against_committed_golden_fixture(cuda_device):
    """The CUDA path vs tests/golden/dt_oracle_small.pt (oracle outputs / loss / gradients committed by tests/golden/make_golden.py)."""
    import os

    from mtrl_b200.model import DecisionTransformer

    fx = torch.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "dt_oracle_small.pt"), weights_only=False)
    dev = cuda_device
    m = DecisionTransformer(11, 3, 64, max_length=10, max_ep_len=100, n_layer=2, n_head=2, activation_function="gelu_new", resid_pdrop=0.0,
                            attn_pdrop=0.0, device=dev)
    m.load_state_dict(fx["state"])
    s, a, r, t, mask = (x.to(dev) for x in fx["inputs"])
    sp, ap, rp = m(s, a, None, r, t, mask)
    keep = mask > 0
    for mine, ref in zip((sp, ap, rp), fx["outputs"]):
        assert rel_err(mine[keep], ref.to(dev)[keep]) <= TOL_ACTION
    out = m.loss_and_backward(s, a, r, t, mask)
    torch.cuda.synchronize()
    assert abs(out[0].item() - fx["loss"].item()) <= TOL_LOSS * abs(fx["loss"].item())
    for n, g in fx["grads"].items():
        assert rel_err(m.get_parameter(n).grad, g) <= TOL_GRAD, n


def test_get_action_graph_reuse_and_weight_update(cuda_device):
    """The cached CUDA graph of get_action must follow (a) a changing history length across calls (episode start -> full window ->
    sliding window) and (b) in-place weight updates (bf16 staging happens outside the graph)."""
    oracle, model, K = build_pair("c1", cuda_device, seed=3)
    dev = cuda_device
    g = torch.Generator().manual_seed(0)
    T = K + 6
    s, a, r, ts = torch.randn(T, 11, generator=g), torch.rand(T, 3, generator=g) * 2 - 1, torch.rand(T, 1, generator=g) * 3, torch.arange(T)
    for t in (1, 2, K - 1, K, K + 1, T):
        aa = a[:t].clone(); aa[-1] = 0
        got = model.get_action(s[:t].to(dev), aa.to(dev), None, r[:t].to(dev), ts[:t].to(dev))
        ref = oracle.get_action(s[:t], aa, None, r[:t], ts[:t])
        assert rel_err(got, ref) <= TOL_ACTION, t
    with torch.no_grad():
        for (n, p), (_, q) in zip(oracle.named_parameters(), model.named_parameters()):
            p.mul_(1.05)
        model.load_state_dict(oracle.state_dict())
    got = model.get_action(s.to(dev), a.to(dev), None, r.to(dev), ts.to(dev))
    ref = oracle.get_action(s, a, None, r, ts)
    assert rel_err(got, ref) <= TOL_ACTION


@pytest.mark.parametrize("t,K", [(0, 5), (1, 5), (3, 5), (5, 5), (17, 30)])
def test_pad_left_bit_exact(cuda_device, t, K):
    """dtb_pad_left: window + LEFT zero padding + mask, integer-exact for every dtype it moves (fp32, int64, uint8 frames)."""
    import ctypes as C

    from mtrl_b200 import _lib

    dev = cuda_device
    srcs = [torch.randn(t, 11, device=dev), torch.arange(t, device=dev, dtype=torch.long) * 7 + 3,
            torch.randint(0, 256, (t, 3, 8, 8), device=dev, dtype=torch.uint8)]
    dsts = [torch.full((K, *x.shape[1:]), 77, device=dev, dtype=x.dtype) for x in srcs]
    mask = torch.full((K,), -1.0, device=dev)
    n = len(srcs)
    sp = (C.c_void_p * n)(*[x.data_ptr() for x in srcs])
    dp = (C.c_void_p * n)(*[x.data_ptr() for x in dsts])
    rb = (C.c_int * n)(*[x[0].numel() * x.element_size() for x in dsts])
    _lib.check(_lib.load().dtb_pad_left(_lib.stream_ptr(), n, sp, dp, rb, t, K, _lib.ptr(mask)))
    torch.cuda.synchronize()
    for x, d in zip(srcs, dsts):
        assert torch.equal(d[K - t:], x) and (d[:K - t] == 0).all()
    assert torch.equal(mask, torch.cat([torch.zeros(K - t), torch.ones(t)]).to(dev))


@pytest.mark.parametrize("name,B", [("c1", 64), ("c3", 4)])
def test_gradients_are_bit_reproducible(cuda_device, name, B):
    """SURVEY.md §8b "Determinism": fixed reduction orders everywhere (split-K slices and column partials summed in index order by the
    last-arriving CTA, embedding-table rows accumulated in timestep order, no atomics on data) -> the whole flat gradient buffer, the
    loss and the predictions are bitwise identical from run to run."""
    _, model, K = build_pair(name, cuda_device, seed=13)
    cfg = CONFIGS[name]
    s, a, r, t, m = make_inputs(B, K, cfg["state_dim"], cfg["act_dim"], 1000, seed=5, onehot_actions=not cfg["action_tanh"])
    t[:, :] = t[:1, :]  # every trajectory gathers the same embed_timestep rows: maximal collisions in the table gradient
    dev = cuda_device
    runs = []
    for _ in range(3):
        out = model.loss_and_backward(s.to(dev), a.to(dev), r.to(dev), t.to(dev), m.to(dev))
        torch.cuda.synchronize()
        runs.append((out.clone(), model.flat_grads.clone()))
    for out, g in runs[1:]:
        assert torch.equal(out, runs[0][0])
        assert torch.equal(g, runs[0][1])


def test_gradient_wrt_vector_states(cuda_device):
    """states.requires_grad with state_dim <= 64: d loss / d states through the autograd bridge vs the oracle."""
    oracle, model, K = build_pair("c1", cuda_device, seed=17)
    s, a, r, t, m = make_inputs(8, K, 11, 3, 1000, seed=31)
    dev = cuda_device
    so = s.clone().requires_grad_(True)
    _, ap_o, _ = oracle(so, a, None, r, t, m)
    w = (m > 0).float()
    ((ap_o ** 2).sum(-1) * w).sum().backward()
    sg = s.to(dev).requires_grad_(True)
    _, ap, _ = model(sg, a.to(dev), None, r.to(dev), t.to(dev), m.to(dev))
    ((ap ** 2).sum(-1) * w.to(dev)).sum().backward()
    torch.cuda.synchronize()
    keep = m > 0  # gradients at padded timesteps are don't-care
    assert sg.grad is not None and sg.grad.dtype == torch.float32
    assert rel_err(sg.grad[keep.to(dev)], so.grad[keep]) <= TOL_GRAD
