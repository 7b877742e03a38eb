"""ORACLE (test infrastructure): multimodal state encoders of the reference, restated in fp32 PyTorch.

Follows /root/reference/src/agents.py:
  * conv_factory('big'|'small')                         agents.py:30-55  (Nature-CNN towers, no padding, bias, ReLU)
  * layer_init: orthogonal(gain sqrt 2), zero bias      utils.py:146-149
  * two towers video_net / audio_net                    agents.py:118-119
  * channel select + `/ 255.0` input scaling            agents.py:133-134
  * fusion 'concat' | 'sum'                             agents.py:137-143
Generalisation: the reference hard-codes in_channels=1 and an 84x84 input (flatten 64*7*7 / 32*9*9);
BASELINE.json names 3x64x64 frames and 1- or 2-channel 64x64 spectrograms, so C_in/H/W are parameters here and the
flatten size is derived.  At (1, 84, 84) this module is numerically the reference's conv_factory
(pinned by oracle/check_vs_reference.py and tests/golden/encoder_ref_*.pt).
Fusion quirk: agents.py:138 concatenates along dim 0 and relies on the reshape at :146, which equals the intended
per-sample [video | audio] only for batch 1; sibling scripts use dim=1 explicitly (src/minecraft_lstm_google.py:83).
The oracle implements the per-sample (dim=1) semantics.
"""
from __future__ import annotations

import numpy as np
import torch
import torch.nn as nn


def layer_init(layer, std=np.sqrt(2), bias_const=0.0):  # utils.py:146-149
    torch.nn.init.orthogonal_(layer.weight, std)
    torch.nn.init.constant_(layer.bias, bias_const)
    return layer


def conv_out(n: int, k: int, s: int) -> int:
    return (n - k) // s + 1


def tower_dims(size: str, h: int, w: int):
    """(conv specs [(C_out, k, stride)], feature size, flatten dims) for conv_factory(size) on an h x w input."""
    if size == "big":
        specs, feat = [(32, 8, 4), (64, 4, 2), (64, 3, 1)], 512
    elif size == "small":
        specs, feat = [(16, 8, 4), (32, 4, 2)], 256
    else:
        raise NotImplementedError(size)
    for _, k, s in specs:
        h, w = conv_out(h, k, s), conv_out(w, k, s)
    return specs, feat, (specs[-1][0], h, w)


def conv_factory(size="big", in_channels=1, in_hw=(84, 84)):
    specs, feat, (c, h, w) = tower_dims(size, *in_hw)
    layers, cin = [], in_channels
    for cout, k, s in specs:
        layers += [layer_init(nn.Conv2d(cin, cout, k, stride=s)), nn.ReLU()]
        cin = cout
    layers += [nn.Flatten(), layer_init(nn.Linear(c * h * w, feat)), nn.ReLU()]
    return nn.Sequential(*layers)


class MultimodalStateEncoder(nn.Module):
    """frames [B,K,Cv,H,W], audio [B,K,Ca,H,W] -> states [B,K,state_dim]; mirrors MinecraftAgent's towers + fusion."""

    def __init__(self, frame_shape=(1, 84, 84), audio_shape=(1, 84, 84), conv_type="big", fusion_type="concat", input_scale=1.0 / 255.0):
        super().__init__()
        self.frame_shape, self.audio_shape = tuple(frame_shape), tuple(audio_shape)
        self.conv_type, self.fusion_type, self.input_scale = conv_type, fusion_type, input_scale
        self.video_net = conv_factory(conv_type, frame_shape[0], frame_shape[1:])
        self.audio_net = conv_factory(conv_type, audio_shape[0], audio_shape[1:])
        self.feature_size = 512 if conv_type == "big" else 256
        if fusion_type == "concat":
            self.state_dim = 2 * self.feature_size
        elif fusion_type == "sum":
            self.state_dim = self.feature_size
        else:
            raise NotImplementedError(fusion_type)

    def forward(self, frames, audio):
        B, K = frames.shape[:2]
        inv = 1.0 / self.input_scale
        if abs(inv - round(inv)) < 1e-6:  # the reference DIVIDES (`/ 255.0`, agents.py:133-134): keep the same rounding
            f, au = frames.reshape(B * K, *self.frame_shape) / float(round(inv)), audio.reshape(B * K, *self.audio_shape) / float(round(inv))
        else:
            f, au = frames.reshape(B * K, *self.frame_shape) * self.input_scale, audio.reshape(B * K, *self.audio_shape) * self.input_scale
        v = self.video_net(f)  # agents.py:133
        a = self.audio_net(au)  # agents.py:134
        fused = torch.cat([v, a], dim=1) if self.fusion_type == "concat" else v + a  # agents.py:137-143 (dim=1 semantics)
        return fused.reshape(B, K, self.state_dim)


class MultimodalDecisionTransformer(nn.Module):
    """Encoder towers -> DecisionTransformer state tokens, trained end to end (SURVEY.md §3.4 last paragraph)."""

    def __init__(self, encoder: MultimodalStateEncoder, dt):
        super().__init__()
        self.encoder, self.dt = encoder, dt

    def forward(self, frames, audio, actions, rewards, returns_to_go, timesteps, attention_mask=None):
        states = self.encoder(frames, audio)
        return self.dt(states, actions, rewards, returns_to_go, timesteps, attention_mask)
