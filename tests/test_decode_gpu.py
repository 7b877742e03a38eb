"""KV-cached incremental get_action (C ABI dtb_dt_decode): one cluster-kernel launch per rollout step.

Exactness claim under test: while the history fits the window (t <= K) appending to the cache is the same function as the reference's
left-padded recompute (A9) — checked against the fp32 oracle's get_action at every step of a rollout, against the library's own recompute
path (tensor-core graph), and bitwise between "step by step" and "whole history in one call" (the per-token arithmetic does not depend
on how the tokens are chunked into launches)."""
import pytest
import torch

from util import CONFIGS, build_pair, rel_err

pytestmark = pytest.mark.gpu


def _history(cfg, T, seed):
    g = torch.Generator().manual_seed(seed)
    s = torch.randn(T, cfg["state_dim"], generator=g)
    a = torch.rand(T, cfg["act_dim"], generator=g) * 2 - 1
    r = torch.rand(T, 1, generator=g) * 3
    ts = torch.arange(T) + 5
    return s, a, r, ts


@pytest.mark.parametrize("name", ["c1", "c4"])
def test_cached_rollout_matches_oracle_and_recompute(cuda_device, name):
    dev = cuda_device
    oracle, model, K = build_pair(name, dev, seed=23)
    model.eval()
    oracle.eval()
    cfg = CONFIGS[name]
    T = min(K, 12) if name == "c4" else K
    s, a, r, ts = _history(cfg, T, 3)
    sd, ad = s.to(dev), a.to(dev)
    rd, td = r.to(dev), ts.to(dev)
    worst = 0.0
    for t in range(1, T + 1):
        at = ad[:t].clone()
        at[-1] = 0  # upstream: the newest action is a zero placeholder
        got = model.get_action(sd[:t], at, None, rd[:t], td[:t], use_cache=True)
        assert model._dc is not None and model._dc["t"] == t  # the decode path ran and cached t timesteps
        a_ref = a[:t].clone()
        a_ref[-1] = 0
        ref = oracle.get_action(s[:t], a_ref, None, r[:t], ts[:t])
        e = rel_err(got, ref)
        worst = max(worst, e)
        assert e <= 1e-2, (t, e)
        if t in (1, 2, T // 2, T):  # the library's own recompute path (pad_left + tensor-core forward graph)
            rec = model.get_action(sd[:t], at, None, rd[:t], td[:t], use_cache=False)
            assert rel_err(got, rec) <= 1e-2, (t, rel_err(got, rec))
    print(f"{name}: worst rel-err of the cached rollout vs the oracle over {T} steps: {worst:.2e}")


def test_stepwise_equals_one_shot_bitwise_and_resets(cuda_device):
    dev = cuda_device
    _, model, K = build_pair("c1", dev, seed=29)
    model.eval()
    cfg = CONFIGS["c1"]
    s, a, r, ts = (x.to(dev) for x in _history(cfg, K, 7))

    def call(t):
        at = a[:t].clone()
        at[-1] = 0
        return model.get_action(s[:t], at, None, r[:t], ts[:t], use_cache=True)

    step = [call(t) for t in range(1, K + 1)]          # append one timestep per call
    for t in (K, 5, 1, 13):                             # a shorter (or equal) history restarts the cache: whole history in one call
        one = call(t)
        assert torch.equal(one, step[t - 1]), t
    # in-place weight update invalidates the cache (stamp of the bf16 staging)
    before = call(K)
    with torch.no_grad():
        model.get_parameter("transformer.h.0.mlp.c_fc.weight").mul_(1.5)
    call(K - 1)
    after = call(K)  # t = t_cached + 1 would append — but the parameters changed since: recomputed from scratch
    assert not torch.equal(before, after)
    model.reset_kv_cache()
    assert torch.equal(call(K), after)
    # beyond the window the call falls back to the recompute path (and the cache is left alone)
    s2, a2, r2, ts2 = (x.to(dev) for x in _history(cfg, K + 7, 9))
    a2[-1] = 0
    out = model.get_action(s2, a2, None, r2, ts2, use_cache=True)
    assert out.shape == (cfg["act_dim"],) and torch.isfinite(out).all()


def test_multimodal_cached_get_action(cuda_device):
    """c2: towers on the new frame only + one decode kernel, vs the recompute path over the whole history."""
    from mtrl_b200.encoder import MultimodalDecisionTransformer, MultimodalStateEncoder

    dev = cuda_device
    torch.manual_seed(0)
    enc = MultimodalStateEncoder((3, 64, 64), (1, 64, 64), "big", "concat", 1.0 / 255.0, device=dev)
    _, dt, K = build_pair("c2", dev, seed=31)
    model = MultimodalDecisionTransformer(enc, dt)
    model.eval()
    g = torch.Generator().manual_seed(1)
    T = 9
    frames = torch.randint(0, 256, (T, 3, 64, 64), generator=g, dtype=torch.uint8).to(dev)
    audio = torch.randn(T, 1, 64, 64, generator=g).clamp_(-1, 1).to(dev)
    acts = torch.nn.functional.one_hot(torch.randint(0, 4, (T,), generator=g), 4).float().to(dev)
    rtg = (torch.rand(T, 1, generator=g) * 3).to(dev)
    ts = torch.arange(T).to(dev)
    for t in range(1, T + 1):
        at = acts[:t].clone()
        at[-1] = 0
        got = model.get_action(frames[:t], audio[:t], at, None, rtg[:t], ts[:t], use_cache=True)
        assert dt._dc["t"] == t
        rec = model.get_action(frames[:t], audio[:t], at, None, rtg[:t], ts[:t], use_cache=False)
        assert rel_err(got, rec) <= 1e-2, (t, rel_err(got, rec))
