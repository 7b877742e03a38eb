"""Fused clip_grad_norm_ + AdamW (dtb_clip_adamw_step) against torch's own clip_grad_norm_ + torch.optim.AdamW on the same flat buffers,
and the trainer's train_step() with the fused optimizer against the oracle trainer with torch's."""
import pytest
import torch

from util import CONFIGS, build_pair, make_inputs

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("max_norm", [0.25, 1e9])
def test_fused_clip_adamw_matches_torch(cuda_device, max_norm):
    from mtrl_b200.optim import FusedClipAdamW

    _, model, K = build_pair("c1", cuda_device)
    _, model2, _ = build_pair("c3", cuda_device, seed=1)  # a second module: one global norm across both
    mods = [model, model2]
    opt = FusedClipAdamW(mods, lr=1e-3, weight_decay=1e-2, max_norm=max_norm)
    ref_p = [torch.nn.Parameter(m.flat_params.detach().clone()) for m in mods]
    ref_opt = torch.optim.AdamW(ref_p, lr=1e-3, weight_decay=1e-2)
    g = torch.Generator(device="cpu").manual_seed(3)
    for m in mods:
        m._stage()
    for it in range(4):
        for m, rp in zip(mods, ref_p):
            grad = (torch.randn(m.flat_params.numel(), generator=g) * (0.01 if it % 2 else 3e-4)).to(cuda_device)
            m.flat_grads.copy_(grad)
            rp.grad = grad.clone()
        ref_norm = torch.nn.utils.clip_grad_norm_(ref_p, max_norm)
        ref_opt.step()
        opt.step()
        torch.cuda.synchronize()
        assert abs(opt.grad_norm().item() - ref_norm.item()) <= 2e-6 * ref_norm.item()
        for i, (m, rp) in enumerate(zip(mods, ref_p)):
            torch.testing.assert_close(m.flat_grads, rp.grad, rtol=3e-6, atol=0)           # clipped in place, like clip_grad_norm_
            torch.testing.assert_close(m.flat_params, rp.data, rtol=2e-6, atol=2e-9)
            st = ref_opt.state[rp]
            torch.testing.assert_close(opt._state["exp_avg"][i], st["exp_avg"], rtol=3e-6, atol=1e-10)   # ~1 ulp of a typical moment
            torch.testing.assert_close(opt._state["exp_avg_sq"][i], st["exp_avg_sq"], rtol=3e-6, atol=1e-17)
            assert torch.equal(m._flat_bf16, m.flat_params.to(torch.bfloat16))  # the bf16 shadow is re-staged in the same pass
    assert opt.step_count == 4


def test_device_side_lr_warmup_matches_lambdalr(cuda_device):
    """warmup_steps: lr * min(t / warmup, 1) from the device step count == upstream's LambdaLR(min((steps + 1) / warmup, 1))."""
    from mtrl_b200.optim import FusedClipAdamW

    _, model, _ = build_pair("c1", cuda_device)
    opt = FusedClipAdamW(model, lr=1e-3, weight_decay=1e-4, max_norm=0.25, warmup_steps=3)
    rp = torch.nn.Parameter(model.flat_params.detach().clone())
    ref_opt = torch.optim.AdamW([rp], lr=1e-3, weight_decay=1e-4)
    sched = torch.optim.lr_scheduler.LambdaLR(ref_opt, lambda steps: min((steps + 1) / 3, 1))
    g = torch.Generator(device="cpu").manual_seed(5)
    for _ in range(5):
        grad = (torch.randn(model.flat_params.numel(), generator=g) * 1e-3).to(cuda_device)
        model.flat_grads.copy_(grad)
        rp.grad = grad.clone()
        torch.nn.utils.clip_grad_norm_([rp], 0.25)
        ref_opt.step()
        sched.step()
        opt.step()
        torch.testing.assert_close(model.flat_params, rp.data, rtol=2e-6, atol=2e-9)


def test_train_step_with_fused_optimizer_tracks_oracle_trainer(cuda_device):
    """3 train_step()s: sm_100a model + FusedClipAdamW vs the fp32 oracle + clip_grad_norm_ + torch.optim.AdamW on the same batches."""
    from mtrl_b200.optim import FusedClipAdamW
    from mtrl_b200.trainer import SequenceTrainer
    from oracle.trainer import SequenceTrainer as OracleTrainer

    oracle, model, K = build_pair("c1", cuda_device)
    cfg = CONFIGS["c1"]
    batches = [make_inputs(32, K, cfg["state_dim"], cfg["act_dim"], 1000, seed=100 + i) for i in range(3)]

    def feeder(dev):
        it = iter(batches)

        def get_batch(bs):
            s, a, r, t, m = [x.to(dev) for x in next(it)]
            return s, a, torch.zeros_like(r), torch.zeros_like(m), r, t, m
        return get_batch

    tr = SequenceTrainer(model, FusedClipAdamW(model, lr=1e-3, weight_decay=1e-4), 32, feeder(cuda_device))
    otr = OracleTrainer(oracle, torch.optim.AdamW(oracle.parameters(), lr=1e-3, weight_decay=1e-4), 32, feeder("cpu"))
    for step in range(3):
        l, lo = tr.train_step(), otr.train_step()
        assert abs(l - lo) <= 3e-2 * abs(lo), (step, l, lo)
    # Adam moves every element by ~lr whatever the gradient's scale, so bf16-vs-fp32 sign flips of near-zero gradients show up here:
    # a sanity bound, not a tolerance claim (the step itself is pinned against torch in the test above)
    worst = 0.0
    for n, p in oracle.named_parameters():
        q = model.get_parameter(n).detach().cpu()
        e = (q - p.detach()).norm().item() / max(p.detach().norm().item(), 1e-12)
        worst = max(worst, e)
        assert e <= 0.1, (n, e)
    print("worst parameter rel-diff after 3 fused steps:", worst)


def test_to_device_then_torch_adamw_restages_the_bf16_shadow(cuda_device):
    """the canonical upstream flow ``model = model.to(device)``; ``torch.optim.AdamW(model.parameters())``; step — the
    kernels must see the updated weights (bf16 shadow re-staged, forward output changes)."""
    from mtrl_b200.model import DecisionTransformer

    torch.manual_seed(0)
    model = DecisionTransformer(11, 3, 128, max_length=20, max_ep_len=1000, n_layer=2, n_head=1, resid_pdrop=0.0, attn_pdrop=0.0, device="cpu")
    model = model.to(cuda_device)
    model.eval()
    s, a, r, t, m = (x.to(cuda_device) for x in make_inputs(4, 20, 11, 3, 1000, seed=2))
    _, ap0, _ = model(s, a, None, r, t, m)
    shadow0 = model._flat_bf16.clone()
    opt = torch.optim.AdamW(model.parameters(), lr=1e-2)
    loss = (model(s, a, None, r, t, m)[1] ** 2).mean()
    loss.backward()
    opt.step()
    _, ap1, _ = model(s, a, None, r, t, m)
    torch.cuda.synchronize()
    assert not torch.equal(model._flat_bf16, shadow0)
    assert not torch.equal(ap0, ap1)
    assert torch.equal(model._flat_bf16, model._flat.to(torch.bfloat16))


def test_loss_fn_path_with_fused_optimizer_tracks_torch_optimizer(cuda_device):
    """SequenceTrainer(loss_fn=...) + FusedClipAdamW — autograd must accumulate into the flat views the optimizer reads,
    cleared every step (same loss trajectory as clip_grad_norm_ + torch.optim.AdamW on an identical model)."""
    from mtrl_b200.optim import FusedClipAdamW
    from mtrl_b200.trainer import SequenceTrainer
    from util import rel_err

    dev = cuda_device
    _, m_fused, K = build_pair("c1", dev, seed=21)
    _, m_torch, _ = build_pair("c1", dev, seed=21)
    batches = [make_inputs(16, K, 11, 3, 1000, seed=300 + i) for i in range(5)]

    def mk():
        it = iter(batches)

        def get_batch(bs):
            s, a, r, t, m = next(it)
            rtg = torch.cat([r, r[:, -1:]], dim=1)
            z = torch.zeros(s.shape[0], K, 1)
            return tuple(x.to(dev) for x in (s, a, z, z, rtg, t, m))
        return get_batch

    loss_fn = lambda s_hat, a_hat, r_hat, s, a, r: torch.mean((a_hat - a) ** 2)  # noqa: E731
    tr_f = SequenceTrainer(m_fused, FusedClipAdamW([m_fused], lr=1e-3, weight_decay=1e-4), 16, mk(), loss_fn=loss_fn)
    tr_t = SequenceTrainer(m_torch, torch.optim.AdamW(m_torch.parameters(), lr=1e-3, weight_decay=1e-4), 16, mk(), loss_fn=loss_fn)
    for i in range(5):
        lf, lt = tr_f.train_step(), tr_t.train_step()
        assert abs(lf - lt) <= 2e-3 * abs(lt), (i, lf, lt)
    assert rel_err(m_fused.flat_params, m_torch.flat_params) < 1e-3  # 5 AdamW steps of bf16-kernel gradients (measured 1.8e-4)
