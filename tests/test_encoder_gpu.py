"""Multimodal state encoder (towers + fusion) and the end-to-end multimodal Decision Transformer vs the fp32 oracle.

The oracle (oracle/encoders.py) is numerically the reference's own conv_factory / two towers / fusion at 1x84x84 (pinned against
/root/reference/src/agents.py by tests/golden/encoder_ref_*.pt, see test_oracle_cpu.py) and its (C_in, H, W) generalisation elsewhere.
Tolerances: features rel-err <= 1e-2; gradients <= 2e-2, relaxed for ReLU-mask flips exactly as in util.grad_tolerances
(every tower tensor sits behind a ReLU): max(base, 2 x torch bf16-autocast error of the oracle on the same inputs); the flips are
random, so two bf16 implementations differ from fp32 by independent draws of the same size.  base = 2e-2, or 3e-2 for the
21-sample tower unit test whose random-projection loss is pure gradient noise.
"""
import copy

import pytest
import torch

from util import CONFIGS, make_inputs, rel_err

pytestmark = pytest.mark.gpu


def _frames(B, K, shape, g):
    return (torch.randint(0, 256, (B, K, *shape), generator=g).float() / 255.0) * 2 - 1  # SURVEY.md §8d synthetic frames


def _spec(B, K, shape, g):
    return torch.randn(B, K, *shape, generator=g).clamp_(-1, 1)


def _pair(frame_shape, audio_shape, conv_type, fusion, dev, input_scale=1.0 / 255.0):
    from mtrl_b200.encoder import MultimodalStateEncoder
    from oracle.encoders import MultimodalStateEncoder as OracleEncoder

    torch.manual_seed(0)
    oracle = OracleEncoder(frame_shape, audio_shape, conv_type, fusion, input_scale)
    with torch.no_grad():
        for n, p in oracle.named_parameters():
            if n.endswith(".bias"):
                p.normal_(0, 0.01 * input_scale * 255)
    enc = MultimodalStateEncoder(frame_shape, audio_shape, conv_type, fusion, input_scale, device=dev)
    enc.load_state_dict(oracle.state_dict())
    return oracle, enc


def _autocast_tol(oracle, fn, base=2e-2, factor=2.0):
    ref = {n: p.grad.clone() for n, p in oracle.named_parameters() if p.grad is not None}
    o2 = copy.deepcopy(oracle)
    o2.zero_grad()
    with torch.autocast("cpu", dtype=torch.bfloat16):
        out = fn(o2)
    out.float().backward()
    ac = {n: rel_err(p.grad, ref[n]) for n, p in o2.named_parameters() if p.grad is not None}
    return lambda n: max(base, factor * ac.get(n, 0.0))


@pytest.mark.parametrize("frame_shape,audio_shape,conv_type,fusion,scale", [
    ((1, 84, 84), (1, 84, 84), "big", "concat", 1 / 255.0),   # reference-faithful (Minecraft.py:53-57, agents.py:118-119)
    ((3, 64, 64), (1, 64, 64), "big", "concat", 1 / 255.0),   # BASELINE.json config 2
    ((3, 64, 64), (2, 64, 64), "big", "concat", 1.0),         # config 3: stereo spectrogram
    ((1, 84, 84), (1, 84, 84), "small", "sum", 1 / 255.0),
    ((3, 64, 64), (1, 64, 64), "small", "concat", 1.0),
    ((1, 64, 64), (1, 64, 64), "big", "sum", 1.0),
])
def test_encoder_forward_backward(cuda_device, frame_shape, audio_shape, conv_type, fusion, scale):
    dev = cuda_device
    oracle, enc = _pair(frame_shape, audio_shape, conv_type, fusion, dev, scale)
    g = torch.Generator().manual_seed(42)
    B, K = 3, 7
    frames, audio = _frames(B, K, frame_shape, g), _spec(B, K, audio_shape, g)
    if scale == 1.0:
        frames, audio = frames * 0.5, audio * 0.5
    wgt = torch.randn(B, K, enc.state_dim, generator=g)
    s_o = oracle(frames, audio)
    oracle.zero_grad()
    (s_o * wgt).sum().backward()
    tol = _autocast_tol(oracle, lambda o: (o(frames, audio).float() * wgt).sum(), base=3e-2)
    s = enc(frames.to(dev), audio.to(dev))
    assert s.shape == s_o.shape and s.dtype == torch.bfloat16
    assert rel_err(s, s_o) <= 1e-2
    enc.zero_grad(set_to_none=True)
    (s.float() * wgt.to(dev)).sum().backward()
    torch.cuda.synchronize()
    for n, p in oracle.named_parameters():
        e = rel_err(enc.get_parameter(n).grad, p.grad)
        assert e <= tol(n), (n, e, tol(n))


@pytest.mark.parametrize("cfg_name,frame_shape,audio_shape,B", [("c2", (3, 64, 64), (1, 64, 64), 4), ("c2", (1, 84, 84), (1, 84, 84), 2),
                                                               ("c3", (3, 64, 64), (2, 64, 64), 2),
                                                               # the shapes bench.py times (BASELINE.json configs[1] / [2] at batch 64 per GPU)
                                                               ("c2", (3, 64, 64), (1, 64, 64), 64), ("c3", (3, 64, 64), (2, 64, 64), 64)])
def test_multimodal_dt_end_to_end(cuda_device, cfg_name, frame_shape, audio_shape, B):
    """frames + spectrograms -> towers -> state tokens -> DT -> masked action MSE; fused step vs the oracle's autograd."""
    from mtrl_b200.encoder import MultimodalDecisionTransformer
    from oracle.encoders import MultimodalDecisionTransformer as OracleMM
    from oracle.trainer import masked_action_mse
    from util import build_pair

    dev = cuda_device
    o_enc, enc = _pair(frame_shape, audio_shape, "big", "concat", dev, 1.0)
    o_dt, dt, K = build_pair(cfg_name, dev)
    cfg = CONFIGS[cfg_name]
    oracle, model = OracleMM(o_enc, o_dt), MultimodalDecisionTransformer(enc, dt)
    g = torch.Generator().manual_seed(7)
    frames, audio = _frames(B, K, frame_shape, g) * 0.5, _spec(B, K, audio_shape, g) * 0.5
    _, a, r, t, m = make_inputs(B, K, 4, cfg["act_dim"], 1000, seed=5, onehot_actions=True)
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
    assert out[1].item() == float((m > 0).sum())
    assert abs(out[0].item() - loss_o.item()) <= 2e-2 * abs(loss_o.item()), (out[0].item(), loss_o.item())
    for n, p in oracle.named_parameters():
        if p.grad is None:
            continue
        e = rel_err(model.get_parameter(n).grad, p.grad)
        assert e <= tol(n), (n, e, tol(n))
    # generic autograd path gives the same gradients as the fused one
    fused = {n: p.grad.clone() for n, p in model.named_parameters() if p.grad is not None}
    model.zero_grad(set_to_none=True)
    _, ap, _ = model(d(frames), d(audio), d(a), None, d(r), d(t), d(m))
    keep = d(m) > 0
    torch.mean((ap[keep] - d(a)[keep]) ** 2).backward()
    torch.cuda.synchronize()
    for n, p in model.named_parameters():
        if n in fused and p.grad is not None and fused[n].norm() > 0:
            assert rel_err(p.grad, fused[n]) <= 1e-2, n


@pytest.mark.parametrize("frame_shape,audio_shape,conv_type", [((3, 64, 64), (1, 64, 64), "big"), ((1, 84, 84), (1, 84, 84), "small")])
def test_uint8_pixel_frames_bit_exact(frame_shape, audio_shape, conv_type):
    """uint8 pixels decoded in-kernel (img * 2 / 255 - 1, REF Minecraft.py:333-335) == the same frames decoded on the host to fp32:
    states and every gradient bit-identical."""
    dev = torch.device("cuda:0")
    g = torch.Generator().manual_seed(7)
    B, K = 3, 5
    _, enc = _pair(frame_shape, audio_shape, conv_type, "concat", dev)
    px = torch.randint(0, 256, (B, K, *frame_shape), generator=g, dtype=torch.uint8)
    px[0, 0].fill_(0); px[0, 1].fill_(255)  # extremes
    f32 = px.float() * 2.0 / 255.0 - 1  # the reference's own expression
    audio = _spec(B, K, audio_shape, g)
    d_states = torch.randn(B, K, enc.state_dim, generator=g).to(torch.bfloat16).to(dev)
    outs = []
    for frames in (f32, px):
        s = enc.forward_raw(frames.to(dev), audio.to(dev)).clone()
        grads = torch.zeros_like(enc.flat_params)
        enc.backward_raw(d_states.clone(), masked=False, grads=grads)
        torch.cuda.synchronize()
        outs.append((s, grads.clone()))
    assert torch.equal(outs[0][0], outs[1][0])
    assert torch.equal(outs[0][1], outs[1][1])
    assert outs[0][1].abs().sum() > 0


def test_multimodal_get_action(cuda_device):
    """Raw frame / spectrogram history -> towers -> DT get_action (uint8 pixels and fp32 frames), vs the oracle on the decoded frames."""
    from mtrl_b200.encoder import MultimodalDecisionTransformer
    from oracle.encoders import MultimodalDecisionTransformer as OracleMM
    from util import build_pair

    dev = cuda_device
    frame_shape, audio_shape = (3, 64, 64), (1, 64, 64)
    o_enc, enc = _pair(frame_shape, audio_shape, "big", "concat", dev, 1.0)
    o_dt, dt, K = build_pair("c2", dev)
    oracle, model = OracleMM(o_enc, o_dt), MultimodalDecisionTransformer(enc, dt)
    g = torch.Generator().manual_seed(11)
    for t in (1, 9, K, K + 4):
        px = torch.randint(0, 256, (t, *frame_shape), generator=g, dtype=torch.uint8)
        f32 = px.float() * 2.0 / 255.0 - 1
        audio = _spec(1, t, audio_shape, g)[0] * 0.5
        a = torch.nn.functional.one_hot(torch.randint(0, 4, (t,), generator=g), 4).float()
        a[-1] = 0
        r, ts = torch.rand(t, 1, generator=g) * 3, torch.arange(t)
        with torch.no_grad():
            states = oracle.encoder(f32[-K:].unsqueeze(0), audio[-K:].unsqueeze(0))[0]
            ref = o_dt.get_action(states, a, None, r, ts)
        for frames in (px, f32):
            got = model.get_action(frames.to(dev), audio.to(dev), a.to(dev), None, r.to(dev), ts.to(dev))
            assert rel_err(got, ref) <= 1e-2, (t, frames.dtype)
