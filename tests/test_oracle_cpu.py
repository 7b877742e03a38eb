"""Pins the oracle (test infrastructure) on CPU: against the committed golden fixtures produced by the REFERENCE's own code
(tests/golden/make_golden.py: /root/reference/src/agents.py conv_factory / towers / fusion) and by the HuggingFace port of the
un-vendored decision-transformer submodule, and against the live HF model when transformers is importable."""
import os

import pytest
import torch

from util import make_inputs

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _load(name):
    return torch.load(os.path.join(GOLD, name), weights_only=False)


def _sums(module):
    return {k: (float(v.double().sum()), float(v.double().abs().sum())) for k, v in module.state_dict().items()}


def _same_weights(module, sums):
    mine = _sums(module)
    assert mine.keys() == sums.keys()
    for k in sums:
        assert mine[k][0] == pytest.approx(sums[k][0], rel=1e-9, abs=1e-9) and mine[k][1] == pytest.approx(sums[k][1], rel=1e-9), k


@pytest.mark.parametrize("size", ["big", "small"])
def test_oracle_towers_equal_reference_conv_factory(size):
    """oracle conv_factory(size, 1, (84, 84)) == reference agents.conv_factory(size) (agents.py:30-55), `/255` scaling (:133-134), fusion."""
    from oracle.encoders import MultimodalStateEncoder

    fx = _load(f"encoder_ref_{size}.pt")
    torch.manual_seed(fx["seed"])
    enc = MultimodalStateEncoder((1, 84, 84), (1, 84, 84), size, "concat")  # creates video_net then audio_net, like the fixture
    _same_weights(enc.video_net, fx["video_sums"])
    _same_weights(enc.audio_net, fx["audio_sums"])
    x = fx["x"]
    with torch.no_grad():
        out = enc(x[:, 0:1].unsqueeze(0), x[:, 1:2].unsqueeze(0))[0]
        enc.fusion_type, enc.state_dim = "sum", enc.feature_size
        out_sum = enc(x[:, 0:1].unsqueeze(0), x[:, 1:2].unsqueeze(0))[0]
    assert torch.equal(out, fx["concat_dim1"])  # bit-exact: same torch ops as the reference
    assert torch.equal(out_sum, fx["sum"])


def test_oracle_fusion_equals_minecraft_agent_at_batch_one():
    """MinecraftAgent.get_states' dim-0 cat + reshape (agents.py:138,146) equals per-sample [video | audio] at B=1."""
    from oracle.encoders import MultimodalStateEncoder

    fx = _load("encoder_ref_fusion_b1.pt")
    torch.manual_seed(fx["seed"])
    enc = MultimodalStateEncoder((1, 84, 84), (1, 84, 84), "big", "concat")
    _same_weights(enc.video_net, fx["video_sums"])
    x = fx["x"]
    with torch.no_grad():
        out = enc(x[:, 0:1].unsqueeze(0), x[:, 1:2].unsqueeze(0))
    assert torch.equal(out.reshape(-1), fx["fused"].reshape(-1))


def test_oracle_dt_equals_hf_fixture():
    from oracle.decision_transformer import DecisionTransformer

    fx = _load("dt_hf_c1.pt")
    o = DecisionTransformer(11, 3, 64, max_length=10, max_ep_len=100, n_layer=2, n_head=2, activation_function="relu", resid_pdrop=0.0,
                            attn_pdrop=0.0).eval()
    o.load_hf_state_dict(fx["state"])
    s, a, r, t, m = fx["inputs"]
    with torch.no_grad():
        outs = o(s, a, None, r, t, m)
    keep = m > 0
    for mine, ref in zip(outs, fx["outputs"]):
        assert (mine[keep] - ref[keep]).abs().max().item() < 2e-5


def test_oracle_dt_equals_live_hf():
    pytest.importorskip("transformers")
    from oracle.check_vs_hf import check

    for kw in [dict(), dict(act="gelu_new", H=64, nh=4, L=2, K=12), dict(sd=16, ad=4, tanh=False, H=96, nh=3)]:
        assert check(**kw) < 2e-5


def test_oracle_reproduces_its_own_golden():
    from oracle.decision_transformer import DecisionTransformer
    from oracle.trainer import masked_action_mse

    fx = _load("dt_oracle_small.pt")
    o = DecisionTransformer(11, 3, 64, max_length=10, max_ep_len=100, n_layer=2, n_head=2, activation_function="gelu_new", resid_pdrop=0.0,
                            attn_pdrop=0.0)
    o.load_state_dict(fx["state"])
    s, a, r, t, m = fx["inputs"]
    sp, ap, rp = o(s, a, None, r, t, m)
    loss = masked_action_mse(ap, a, m)
    loss.backward()
    assert torch.allclose(ap, fx["outputs"][1], atol=1e-6) and abs(loss.item() - fx["loss"].item()) < 1e-7
    for n, p in o.named_parameters():
        if n in fx["grads"]:
            assert torch.allclose(p.grad, fx["grads"][n], atol=1e-6, rtol=1e-4), n


def test_token_interleave_and_mask_rules():
    """A2/A3 integer rules of the oracle itself: changing inputs under the mask leaves unmasked predictions untouched (exactly 0)."""
    from oracle.decision_transformer import DecisionTransformer

    torch.manual_seed(0)
    o = DecisionTransformer(5, 2, 32, max_length=6, max_ep_len=50, n_layer=1, n_head=1, resid_pdrop=0.0, attn_pdrop=0.0).eval()
    s, a, r, t, m = make_inputs(4, 6, 5, 2, 50, seed=3)
    with torch.no_grad():
        _, ap1, _ = o(s, a, None, r, t, m)
        s2, a2 = s.clone(), a.clone()
        s2[m == 0], a2[m == 0] = 9.0, -3.0
        _, ap2, _ = o(s2, a2, None, r, t, m)
    assert torch.equal(ap1[m > 0], ap2[m > 0])


@pytest.mark.parametrize("t_hist", [1, 5, 6, 11])
def test_get_action_padding_rules(t_hist):
    """A9: last-K window, LEFT zero padding, mask [0]*(K-t)+[1]*t, action of the last state token == explicit padded forward."""
    from oracle.decision_transformer import DecisionTransformer

    torch.manual_seed(1)
    K = 6
    o = DecisionTransformer(5, 2, 32, max_length=K, max_ep_len=50, n_layer=2, n_head=2, resid_pdrop=0.0, attn_pdrop=0.0).eval()
    g = torch.Generator().manual_seed(t_hist)
    s, a, r, ts = torch.randn(t_hist, 5, generator=g), torch.randn(t_hist, 2, generator=g), torch.rand(t_hist, 1, generator=g), torch.arange(t_hist)
    got = o.get_action(s, a, None, r, ts)
    t = min(t_hist, K)
    pad = K - t
    S = torch.cat([torch.zeros(pad, 5), s[-K:]])[None]
    A = torch.cat([torch.zeros(pad, 2), a[-K:]])[None]
    R = torch.cat([torch.zeros(pad, 1), r[-K:]])[None]
    T = torch.cat([torch.zeros(pad, dtype=torch.long), ts[-K:]])[None]
    M = torch.cat([torch.zeros(pad), torch.ones(t)])[None]
    with torch.no_grad():
        _, ap, _ = o(S, A, None, R, T, M)
    assert torch.equal(got, ap[0, -1])


def test_train_step_loss_rule():
    """A10: loss = mean((a_hat - a)^2) over rows with attention_mask > 0; clip 0.25; rtg[:, :-1] when rtg carries K+1 entries."""
    from oracle.decision_transformer import DecisionTransformer
    from oracle.trainer import SequenceTrainer, masked_action_mse

    torch.manual_seed(2)
    o = DecisionTransformer(5, 2, 32, max_length=6, max_ep_len=50, n_layer=1, n_head=1, resid_pdrop=0.0, attn_pdrop=0.0)
    s, a, r, t, m = make_inputs(4, 6, 5, 2, 50, seed=4)
    with torch.no_grad():
        _, ap, _ = o(s, a, None, r, t, m)
    manual = sum(((ap[b, k] - a[b, k]) ** 2).sum() for b in range(4) for k in range(6) if m[b, k] > 0) / (int((m > 0).sum()) * 2)
    assert abs(masked_action_mse(ap, a, m).item() - manual.item()) < 1e-6
    rtg = torch.cat([r, r[:, -1:]], dim=1)
    tr = SequenceTrainer(o, torch.optim.AdamW(o.parameters(), lr=1e-3), 4, lambda bs: (s, a, None, None, rtg, t, m))
    l0 = tr.train_step()
    assert abs(l0 - manual.item()) < 1e-6
    assert tr.train_step() < l0


@pytest.mark.parametrize("kind,F", [("casl", 512), ("new", 256)])
def test_oracle_modal_attention_equals_reference_classes(kind, F):
    """oracle CaslAttention / NewAttention == the REFERENCE's classes (agents.py:58-90) on the committed fixture: bit-exact."""
    from oracle.encoders import CaslAttention, NewAttention

    d = torch.load(os.path.join(GOLD, "modal_attn_ref.pt"))[kind]
    m = (CaslAttention if kind == "casl" else NewAttention)(F)
    assert list(m.state_dict().keys()) == list(d["state"].keys())  # same module names and order as the reference
    m.load_state_dict(d["state"])
    with torch.no_grad():
        v2, a2, w = m(d["v"], d["a"], d["h"][0])
    assert torch.equal(v2, d["v_out"]) and torch.equal(a2, d["a_out"]) and torch.equal(w, d["w"])
    # no hidden state (the Decision Transformer case) == an all-zero hidden state
    with torch.no_grad():
        w0 = m(d["v"], d["a"], None)[2]
        wz = m(d["v"], d["a"], torch.zeros(d["v"].shape[0], 128))[2]
    assert torch.equal(w0, wz)


def test_oracle_encoder_with_attention_param_names():
    """attn.* parameter names follow MinecraftAgent (self.attn = CaslAttention / NewAttention, agents.py:110-113)."""
    from oracle.encoders import MultimodalStateEncoder

    e = MultimodalStateEncoder((1, 84, 84), (1, 84, 84), "small", "sum", attn_type="casl")
    names = [n for n, _ in e.named_parameters() if n.startswith("attn.")]
    assert names == [f"attn.{m}.{k}" for m in ("audio_fc", "video_fc", "state_fc", "attention") for k in ("weight", "bias")]
    x = torch.rand(2, 3, 1, 84, 84)
    s = e(x, x)
    assert s.shape == (2, 3, 256) and e.attn_weights.shape == (6, 2)


def test_oracle_mfcc_restatement_properties():
    """oracle/mfcc.py restates python_speech_features 0.6 (absent from this image: parity unpinned for the cepstra).  What can be
    checked without it: the pieces it delegates to numpy / scipy are called the same way, and its documented invariants hold."""
    import numpy as np
    from scipy.fftpack import dct

    from oracle import mfcc as O

    sr, n = 16000, 4000
    rng = np.random.default_rng(0)
    t = np.arange(n) / sr
    x = np.round(5000 * np.sin(2 * np.pi * 1000 * t) + 300 * rng.standard_normal(n)).astype(np.int16)
    assert O.calculate_nfft(sr, 0.025) == 512 and O.calculate_nfft(44100, 0.025) == 2048 and O.calculate_nfft(24000, 0.025) == 1024
    assert O.frame_geometry(n, sr) == (400, 160, 1 + int(np.ceil((n - 400) / 160)))
    assert O.round_half_up(1102.5) == 1103 and O.round_half_up(2.5) == 3  # not banker's rounding
    fb = O.get_filterbanks(26, 512, sr)
    assert fb.shape == (26, 257) and fb.min() >= 0 and fb.max() <= 1.0
    c = O.psf_mfcc(x, sr)
    assert c.shape == (O.frame_geometry(n, sr)[2], 13) and c.dtype == np.float64
    # c0 is the log frame energy of the pre-emphasised frame (appendEnergy=True); Parseval ties it to the time domain
    y = np.append(x[0], x[1:] - 0.97 * x[:-1])
    # (the one-sided spectrum holds half of the energy, up to the DC / Nyquist bins)
    assert c[3, 0] == pytest.approx(np.log(np.sum(y[480:880] ** 2) / 2), abs=1e-2)
    # the DCT it delegates to: ortho DCT-II of a constant vector has only the DC term
    d = dct(np.ones((1, 26)), type=2, axis=1, norm="ortho")
    assert d[0, 0] == pytest.approx(np.sqrt(26)) and np.abs(d[0, 1:]).max() < 1e-12
    # a 1 kHz tone: the strongest filter contains the tone's bin
    b = O.filterbank_bins(26, 512, sr)
    pspec = np.abs(np.fft.rfft(y[480:880], 512)) ** 2 / 512
    j = int(np.argmax(pspec @ fb.T))
    assert b[j] <= 1000 * 513 / sr <= b[j + 2]
    # stereo input: pre-emphasis per channel, then numpy.append flattens to interleaved samples (as inside the reference's call)
    st = np.stack([x, x[::-1]], axis=1).astype(np.float64)
    pre = np.concatenate([st[:1], st[1:] - 0.97 * st[:-1]]).reshape(-1)
    assert np.array_equal(O.psf_mfcc(st, sr), O.psf_mfcc(pre, sr, preemph=0.0))
    img = O.wav_to_mfcc(sr, x)
    assert img.shape == (84, 84) and img.dtype == np.float32 and img.min() == -1.0 and img.max() == 1.0


def test_oracle_mfcc_reproduces_committed_golden():
    import numpy as np

    from oracle import mfcc as O

    d = np.load(os.path.join(GOLD, "mfcc_ref.npz"))
    assert np.array_equal(O.psf_mfcc(d["wav"], int(d["samplerate"])), d["cepstra"])
    assert np.array_equal(O.wav_to_mfcc(int(d["samplerate"]), d["wav"]), d["image"])


def test_oracle_mfcc_cross_checked_against_torchaudio_blocks():
    """Second source for the unpinned MFCC restatement: torchaudio 2.11 building blocks, stage by
    stage.  Power spectrum and DCT share their definition with python_speech_features and must agree to rounding; the mel filterbank
    differs by design (floored FFT bins vs exact frequencies) — the residual on the reference's three assets is committed beside the
    generator (oracle/check_mfcc_torchaudio.py -> tests/golden/mfcc_torchaudio_residual.json)."""
    import json

    import numpy as np

    pytest.importorskip("torchaudio")
    from oracle.check_mfcc_torchaudio import stages

    rng = np.random.default_rng(3)
    t = np.arange(24000) / 24000.0
    sig = (3000 * np.sin(2 * np.pi * 440 * t) + 500 * rng.standard_normal(t.size)).astype(np.int16)
    r = stages(sig, 24000)
    assert r["powspec_vs_torch_stft_max_rel"] < 1e-12
    assert r["dct_vs_torchaudio_create_dct_max_rel"] < 1e-5
    assert 0.0 < r["filterbank_vs_torchaudio_melscale_fbanks_rel_l2"] < 0.1      # the documented design difference, bounded
    assert r["cepstra_with_torchaudio_filterbank_rel_l2"] < 0.1
    fx = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "mfcc_torchaudio_residual.json")))
    for name in ("noise", "target_good", "target_bad"):
        assert fx[name]["powspec_vs_torch_stft_max_rel"] < 1e-12 and fx[name]["dct_vs_torchaudio_create_dct_max_rel"] < 1e-5
        assert fx[name]["cepstra_with_torchaudio_filterbank_rel_l2"] < 0.06
