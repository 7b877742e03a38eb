"""Modality attention (REF src/agents.py:58-75 CaslAttention, :78-90 NewAttention, call site :135-136) through
the C ABI (dtb_modal_attn_forward / _backward) vs the oracle's restatement, which tests/golden/modal_attn_ref.pt pins bit-exactly
to the reference's own classes (test_oracle_cpu.py).

The op consumes the towers' bf16 features in place, so the oracle is evaluated on the same bf16-rounded features (as fp32): what is
left is the bf16 rounding of the weights / hidden operand of the tensor-core GEMM and of the bf16 outputs.  Tolerances: states and
attention weights rel-err <= 1e-2, every gradient <= 2e-2 (north star).  The ReLU mask of the towers (feat > 0) that the backward
folds into d_feat is an exact integer rule and is checked with `==` on the zero pattern.
"""
import ctypes as C

import pytest
import torch

from util import rel_err

pytestmark = pytest.mark.gpu


def _oracle(kind, F):
    from oracle.encoders import CaslAttention, NewAttention

    return (CaslAttention if kind == "casl" else NewAttention)(F)


def _flat_params(oracle, kind, cfg, dev):
    from mtrl_b200 import _lib

    lib = _lib.load()
    n = _lib.check(lib.dtb_modal_attn_param_entries(C.byref(cfg)))
    offs, nums = (C.c_longlong * n)(), (C.c_longlong * n)()
    total = int(lib.dtb_modal_attn_param_layout(C.byref(cfg), offs, nums))
    mods = ["audio_fc", "video_fc", "state_fc", "attention"] if kind == "casl" else ["fc", "attention"]
    names = [f"{m}.{k}" for m in mods for k in ("weight", "bias")]
    flat = torch.zeros(total, dtype=torch.float32)
    sd = oracle.state_dict()
    for name, off, num in zip(names, offs, nums):
        assert sd[name].numel() == num, (name, num)
        flat[off:off + num] = sd[name].reshape(-1)
    return flat.to(dev), names, [int(o) for o in offs], [int(x) for x in nums]


def _run(kind, F, fusion, N, with_hidden, dev, seed=0, state=None, inputs=None):
    from mtrl_b200 import _lib

    lib = _lib.load()
    torch.manual_seed(seed)
    oracle = _oracle(kind, F)
    if state is not None:
        oracle.load_state_dict(state)
    cfg = _lib.ModalAttnConfig(0 if kind == "casl" else 1, F, 0 if fusion == "concat" else 1)
    flat, names, offs, nums = _flat_params(oracle, kind, cfg, dev)
    g = torch.Generator().manual_seed(seed + 1)
    if inputs is None:
        v, a = torch.relu(torch.randn(N, F, generator=g)), torch.relu(torch.randn(N, F, generator=g))
        h = torch.randn(N, 128, generator=g) if with_hidden else None
    else:
        v, a, h = inputs
    feat = torch.cat([v, a], dim=1).to(torch.bfloat16)
    S = 2 * F if fusion == "concat" else F
    assert lib.dtb_modal_attn_state_dim(C.byref(cfg)) == S
    dS = torch.randn(N, S, generator=g).to(torch.bfloat16)

    # oracle on the bf16-rounded features
    vo, ao = feat[:, :F].float().requires_grad_(True), feat[:, F:].float().requires_grad_(True)
    v2, a2, w_o = oracle(vo, ao, h)
    s_o = torch.cat([v2, a2], dim=1) if fusion == "concat" else v2 + a2
    oracle.zero_grad()
    (s_o * dS.float()).sum().backward()
    dfeat_o = torch.cat([vo.grad, ao.grad], dim=1) * (feat.float() > 0)

    ws = torch.empty(int(lib.dtb_modal_attn_workspace_bytes(C.byref(cfg), N)), dtype=torch.uint8, device=dev)
    feat_d, dS_d = feat.to(dev), dS.to(dev)
    h_d = None if h is None else h.to(dev).contiguous()
    states = torch.empty(N, S, dtype=torch.bfloat16, device=dev)
    w = torch.empty(N, 2, dtype=torch.float32, device=dev)
    _lib.check(lib.dtb_modal_attn_forward(C.byref(cfg), _lib.stream_ptr(), _lib.ptr(flat), N, _lib.ptr(feat_d), _lib.ptr(h_d), _lib.ptr(ws),
                                          ws.numel(), _lib.ptr(states), _lib.ptr(w)))
    dfeat = torch.empty(N, 2 * F, dtype=torch.bfloat16, device=dev)
    grads = torch.full_like(flat, float("nan"))
    _lib.check(lib.dtb_modal_attn_backward(C.byref(cfg), _lib.stream_ptr(), _lib.ptr(flat), N, _lib.ptr(feat_d), int(h is not None),
                                           _lib.ptr(dS_d), _lib.ptr(ws), ws.numel(), _lib.ptr(dfeat), _lib.ptr(grads)))
    torch.cuda.synchronize()
    got = {name: grads[off:off + num].cpu() for name, off, num in zip(names, offs, nums)}
    return dict(oracle=oracle, states=states.cpu(), w=w.cpu(), s_o=s_o.detach(), w_o=w_o.detach(), dfeat=dfeat.cpu(), dfeat_o=dfeat_o,
                grads=got, feat=feat, has_hidden=h is not None)


def _check(r):
    assert rel_err(r["states"], r["s_o"]) <= 1e-2
    assert rel_err(r["w"], r["w_o"]) <= 1e-2
    assert torch.allclose(r["w"].sum(-1), torch.ones(r["w"].shape[0]), atol=1e-5)
    # integer rule: the towers' ReLU mask is folded into d_feat
    assert bool((r["dfeat"].float()[r["feat"].float() <= 0] == 0).all())
    assert rel_err(r["dfeat"], r["dfeat_o"]) <= 2e-2
    for n, p in r["oracle"].named_parameters():
        g = r["grads"][n]
        assert torch.isfinite(g).all(), n
        if not r["has_hidden"] and n == "state_fc.weight":
            assert bool((g == 0).all())
            continue
        ref = p.grad.reshape(-1)
        if not r["has_hidden"] and n == "fc.weight":  # hidden columns: exactly zero without a hidden state
            F2 = p.shape[1] - 128
            assert bool((g.view(p.shape)[:, F2:] == 0).all())
        e = rel_err(g, ref)
        assert e <= 2e-2, (n, e)


@pytest.mark.parametrize("kind,F", [("casl", 512), ("new", 512), ("casl", 256), ("new", 256)])
@pytest.mark.parametrize("fusion", ["concat", "sum"])
@pytest.mark.parametrize("N,with_hidden", [(1920, True), (1920, False), (37, True), (1, False)])
def test_modal_attn_vs_oracle(cuda_device, kind, F, fusion, N, with_hidden):
    _check(_run(kind, F, fusion, N, with_hidden, cuda_device, seed=N + F))


def test_modal_attn_reference_golden(cuda_device):
    """Inputs, weights and outputs of the REFERENCE's own classes (tests/golden/modal_attn_ref.pt, made by make_golden.py)."""
    import os

    fx = torch.load(os.path.join(os.path.dirname(__file__), "golden", "modal_attn_ref.pt"))
    for kind, F in [("casl", 512), ("new", 256)]:
        d = fx[kind]
        n = d["v"].shape[0]
        h = d["h"][0].expand(n, 128).contiguous() if d["h"].shape[1] == 1 else d["h"][0]
        r = _run(kind, F, "concat", n, True, cuda_device, state=d["state"], inputs=(d["v"], d["a"], h))
        ref = torch.cat([d["v_out"], d["a_out"]], dim=1)
        assert rel_err(r["states"], ref) <= 1e-2
        assert rel_err(r["w"], d["w"]) <= 1e-2


@pytest.mark.parametrize("attn,conv_type,fusion", [("casl", "big", "concat"), ("new", "big", "sum"), ("new", "small", "concat"),
                                                   ("casl", "small", "sum")])
def test_encoder_with_modal_attention(cuda_device, attn, conv_type, fusion):
    """Towers -> modality attention -> fusion through MultimodalStateEncoder(attn_type=...) vs the oracle encoder (autograd)."""
    from mtrl_b200.encoder import MultimodalStateEncoder
    from oracle.encoders import MultimodalStateEncoder as OracleEncoder
    from test_encoder_gpu import _autocast_tol, _frames, _spec

    dev = cuda_device
    fs, as_ = (3, 64, 64), (1, 64, 64)
    torch.manual_seed(0)
    oracle = OracleEncoder(fs, as_, conv_type, fusion, 1.0, attn_type=attn)
    with torch.no_grad():
        for n, p in oracle.named_parameters():
            if n.endswith(".bias") and not n.startswith("attn."):
                p.normal_(0, 0.01)
    enc = MultimodalStateEncoder(fs, as_, conv_type, fusion, 1.0, device=dev, attn_type=attn)
    assert sorted(n for n, _ in enc.named_parameters()) == sorted(n for n, _ in oracle.named_parameters())
    enc.load_state_dict(oracle.state_dict())
    g = torch.Generator().manual_seed(42)
    B, K = 4, 12
    frames, audio = _frames(B, K, fs, g) * 0.5, _spec(B, K, as_, g) * 0.5
    hidden = torch.randn(B * K, 128, generator=g) if attn == "casl" else torch.randn(1, 128, generator=g)
    wgt = torch.randn(B, K, enc.state_dim, generator=g)
    s_o = oracle(frames, audio, hidden)
    oracle.zero_grad()
    (s_o * wgt).sum().backward()
    tol = _autocast_tol(oracle, lambda o: (o(frames, audio, hidden).float() * wgt).sum(), base=3e-2)
    s = enc(frames.to(dev), audio.to(dev), hidden.to(dev))
    assert s.shape == s_o.shape and s.dtype == torch.bfloat16
    assert rel_err(s, s_o) <= 1e-2
    assert rel_err(enc.attn_weights.reshape(-1, 2), oracle.attn_weights) <= 1e-2
    enc.zero_grad(set_to_none=True)
    (s.float() * wgt.to(dev)).sum().backward()
    torch.cuda.synchronize()
    for n, p in oracle.named_parameters():
        e = rel_err(enc.get_parameter(n).grad, p.grad)
        assert e <= tol(n), (n, e, tol(n))


def test_multimodal_dt_with_modal_attention_trains(cuda_device):
    """End to end: towers -> casl attention (no hidden state: the DT has none) -> DT fused step vs the oracle's autograd, and
    get_action on the cached graph."""
    from mtrl_b200.encoder import MultimodalDecisionTransformer, MultimodalStateEncoder
    from oracle.encoders import MultimodalDecisionTransformer as OracleMM, MultimodalStateEncoder as OracleEncoder
    from oracle.trainer import masked_action_mse
    from test_encoder_gpu import _autocast_tol, _frames, _spec
    from util import CONFIGS, build_pair, make_inputs

    dev = cuda_device
    fs, as_ = (3, 64, 64), (1, 64, 64)
    torch.manual_seed(0)
    o_enc = OracleEncoder(fs, as_, "big", "concat", 1.0, attn_type="casl")
    enc = MultimodalStateEncoder(fs, as_, "big", "concat", 1.0, device=dev, attn_type="casl")
    enc.load_state_dict(o_enc.state_dict())
    o_dt, dt, K = build_pair("c2", dev)
    oracle, model = OracleMM(o_enc, o_dt), MultimodalDecisionTransformer(enc, dt)
    B = 4
    g = torch.Generator().manual_seed(7)
    frames, audio = _frames(B, K, fs, g) * 0.5, _spec(B, K, as_, g) * 0.5
    _, a, r, t, m = make_inputs(B, K, 4, CONFIGS["c2"]["act_dim"], 1000, seed=5, onehot_actions=True)
    frames[m == 0] = 0
    audio[m == 0] = 0

    def step(o):
        _, ap, _ = o(frames, audio, a, None, r, t, m)
        return masked_action_mse(ap.float(), a, m)

    oracle.zero_grad()
    loss_o = step(oracle)
    loss_o.backward()
    tol = _autocast_tol(oracle, step)
    d = lambda x: x.to(dev)
    out = model.loss_and_backward(d(frames), d(audio), d(a), d(r), d(t), d(m))
    torch.cuda.synchronize()
    assert abs(out[0].item() - loss_o.item()) <= 2e-2 * abs(loss_o.item()), (out[0].item(), loss_o.item())
    for n, p in oracle.named_parameters():
        if p.grad is None:
            continue
        if n == "encoder.attn.state_fc.weight":  # no hidden state: exactly zero on both sides
            assert bool((model.get_parameter(n).grad == 0).all()) and bool((p.grad == 0).all())
            continue
        e = rel_err(model.get_parameter(n).grad, p.grad)
        assert e <= tol(n), (n, e, tol(n))
    # get_action: last-step action of the B = 1 forward over the first trajectory's (unpadded) history
    model.eval(); oracle.eval()
    b = 1
    with torch.no_grad():
        _, ap, _ = oracle(frames[b:b + 1], audio[b:b + 1], a[b:b + 1], None, r[b:b + 1], t[b:b + 1], torch.ones(1, K))
    act = model.get_action(d(frames[b]), d(audio[b]), d(a[b]), None, d(r[b]), d(t[b]))
    assert rel_err(act, ap[0, -1]) <= 1e-2
