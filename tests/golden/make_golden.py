"""Generates the committed golden fixtures (run in the build container, where /root/reference and transformers exist):

  encoder_ref_{big,small}.pt : the REFERENCE's own conv_factory / two towers / fusion (imported from /root/reference/src/agents.py
                               with stub modules for its absent deps) on seeded inputs -> pins oracle/encoders.py
  modal_attn_ref.pt          : the REFERENCE's CaslAttention / NewAttention on seeded features -> pins the oracle's restatement
  mfcc_ref.npz               : 1 s excerpt of the reference's asset target_good.wav + the oracle's MFCC cepstra / image of it
  dt_hf_c1.pt                : HuggingFace DecisionTransformerModel (wpe zeroed = upstream semantics) on seeded inputs -> pins
                               oracle/decision_transformer.py
  dt_oracle_small.pt         : oracle forward / loss / gradient norms on a tiny config -> fixture the GPU path is compared with
Usage: python tests/golden/make_golden.py
"""
import os
import sys
from unittest.mock import MagicMock

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def import_reference_agents():
    for m in ["gym", "gym.spaces", "gym.spaces.box", "gym.spaces.discrete", "stable_baselines3", "stable_baselines3.common",
              "stable_baselines3.common.atari_wrappers", "Minecraft", "python_speech_features"]:
        sys.modules.setdefault(m, MagicMock())
    sys.path.insert(0, "/root/reference/src")
    import agents  # noqa: E402

    return agents


def checksums(module):
    return {k: (float(v.double().sum()), float(v.double().abs().sum())) for k, v in module.state_dict().items()}


def encoder_fixture(agents, size):
    torch.manual_seed(7)
    video, audio = agents.conv_factory(size), agents.conv_factory(size)
    g = torch.Generator().manual_seed(11)
    x = torch.rand(5, 2, 84, 84, generator=g) * 2 - 1  # obs layout (2, 84, 84) = [frame, MFCC image] (Minecraft.py:128-132)
    with torch.no_grad():
        v = video(torch.index_select(x, 1, torch.tensor([0])) / 255.0)  # agents.py:133
        a = audio(torch.index_select(x, 1, torch.tensor([1])) / 255.0)  # agents.py:134
    # weights are NOT stored (13 MB): both towers are re-created from the same seed by the oracle's conv_factory (same init call
    # order as the reference's) and verified against these per-tensor checksums before the outputs are compared
    return {"seed": 7, "x": x, "video_sums": checksums(video), "audio_sums": checksums(audio), "video_out": v, "audio_out": a,
            "concat_dim1": torch.cat([v, a], dim=1), "sum": v + a}


def fusion_b1_fixture(agents):
    """MinecraftAgent.get_states features at B=1 (where the reference's dim-0 cat + reshape equals per-sample concat)."""
    torch.manual_seed(3)
    env = MagicMock()
    env.single_action_space.n = 4
    ag = agents.MinecraftAgent(env, torch.device("cpu"), "big", None, "concat")
    x = torch.rand(1, 2, 84, 84, generator=torch.Generator().manual_seed(5)) * 2 - 1
    with torch.no_grad():
        v = ag.video_net(torch.index_select(x, 1, torch.tensor([0])) / 255.0)
        a = ag.audio_net(torch.index_select(x, 1, torch.tensor([1])) / 255.0)
        fused = torch.cat([v, a]).reshape(-1, 1, ag.lstm.input_size)  # agents.py:138,146
    return {"seed": 3, "x": x, "video_sums": checksums(ag.video_net), "audio_sums": checksums(ag.audio_net), "fused": fused}


def modal_attn_fixture(agents):
    """The REFERENCE's CaslAttention / NewAttention (agents.py:58-90) on seeded features -> pins oracle/encoders.py's restatement.
    casl takes lstm_state[0] of shape [1, N, 128] (per-sample rows); new repeats ONE row [1, 1, 128] over the samples."""
    out = {}
    g = torch.Generator().manual_seed(21)
    for name, cls, F in [("casl", agents.CaslAttention, 512), ("new", agents.NewAttention, 256)]:
        torch.manual_seed(13)
        m = cls(F, torch.device("cpu"))
        n = 6
        v, a = torch.relu(torch.randn(n, F, generator=g)), torch.relu(torch.randn(n, F, generator=g))
        h = torch.randn(1, n if name == "casl" else 1, 128, generator=g)
        with torch.no_grad():
            v2, a2, w = m(v, a, (h, torch.zeros_like(h)))
        out[name] = {"state": m.state_dict(), "v": v, "a": a, "h": h, "v_out": v2, "a_out": a2, "w": w}
    return out


def mfcc_fixture():
    """1 s of the reference's stereo asset target_good.wav (24 kHz int16) -> oracle cepstra and image (oracle/mfcc.py)."""
    import numpy as np
    import scipy.io.wavfile as wav

    from oracle import mfcc as O

    sr, a = wav.read("/root/reference/environments/Minecraft/assets/target_good.wav")
    a = np.ascontiguousarray(a[24000:48000])
    return {"samplerate": np.int64(sr), "wav": a, "cepstra": O.psf_mfcc(a, sr), "image": O.wav_to_mfcc(sr, a)}


def hf_fixture():
    from transformers import DecisionTransformerConfig, DecisionTransformerModel
    from util import make_inputs

    cfg = DecisionTransformerConfig(state_dim=11, act_dim=3, hidden_size=64, max_ep_len=100, action_tanh=True, n_layer=2, n_head=2,
                                    n_inner=256, activation_function="relu", n_positions=1024, resid_pdrop=0.0, attn_pdrop=0.0, embd_pdrop=0.0)
    torch.manual_seed(1)
    hf = DecisionTransformerModel(cfg).eval()
    hf.encoder.wpe.weight.data.zero_()
    s, a, r, t, m = make_inputs(4, 10, 11, 3, 100, seed=2)
    with torch.no_grad():
        sp, ap, rp = hf(states=s, actions=a, rewards=None, returns_to_go=r, timesteps=t, attention_mask=m, return_dict=False)[:3]
    sd = {k: v for k, v in hf.state_dict().items() if ".wte." not in k and ".wpe." not in k and not k.endswith(".attn.bias")}
    return {"state": sd, "inputs": (s, a, r, t, m), "outputs": (sp, ap, rp)}


def oracle_small_fixture():
    from oracle.decision_transformer import DecisionTransformer
    from oracle.trainer import masked_action_mse
    from util import make_inputs

    torch.manual_seed(5)
    o = DecisionTransformer(11, 3, 64, max_length=10, max_ep_len=100, n_layer=2, n_head=2, activation_function="gelu_new", resid_pdrop=0.0,
                            attn_pdrop=0.0)
    s, a, r, t, m = make_inputs(4, 10, 11, 3, 100, seed=6)
    sp, ap, rp = o(s, a, None, r, t, m)
    loss = masked_action_mse(ap, a, m)
    loss.backward()
    return {"state": o.state_dict(), "inputs": (s, a, r, t, m), "outputs": (sp.detach(), ap.detach(), rp.detach()), "loss": loss.detach(),
            "grads": {n: p.grad.clone() for n, p in o.named_parameters() if p.grad is not None}}


if __name__ == "__main__":
    agents = import_reference_agents()
    torch.save(encoder_fixture(agents, "big"), os.path.join(HERE, "encoder_ref_big.pt"))
    torch.save(encoder_fixture(agents, "small"), os.path.join(HERE, "encoder_ref_small.pt"))
    torch.save(fusion_b1_fixture(agents), os.path.join(HERE, "encoder_ref_fusion_b1.pt"))
    torch.save(modal_attn_fixture(agents), os.path.join(HERE, "modal_attn_ref.pt"))
    import numpy as np
    np.savez_compressed(os.path.join(HERE, "mfcc_ref.npz"), **mfcc_fixture())
    torch.save(hf_fixture(), os.path.join(HERE, "dt_hf_c1.pt"))
    torch.save(oracle_small_fixture(), os.path.join(HERE, "dt_oracle_small.pt"))
    for f in sorted(os.listdir(HERE)):
        print(f, os.path.getsize(os.path.join(HERE, f)))
