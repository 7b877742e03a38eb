"""Pin the DT oracle against the in-container HuggingFace port (wpe zeroed = upstream semantics).  Run on CPU."""
from __future__ import annotations

import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle.decision_transformer import DecisionTransformer  # noqa: E402


def make_inputs(B, K, sd, ad, E, seed, pad=True):
    g = torch.Generator().manual_seed(seed)
    states = torch.randn(B, K, sd, generator=g)
    actions = torch.rand(B, K, ad, generator=g) * 2 - 1
    rtg = torch.rand(B, K, 1, generator=g) * 3
    start = torch.randint(0, E - K, (B, 1), generator=g)
    timesteps = start + torch.arange(K)[None]
    mask = torch.ones(B, K)
    if pad:
        for b in range(0, B, 2):
            p = int(torch.randint(0, K // 2 + 1, (1,), generator=g))
            mask[b, :p] = 0
            states[b, :p] = 0
            actions[b, :p] = -10.0
            rtg[b, :p] = 0
            timesteps[b, :p] = 0
    return states, actions, rtg, timesteps, mask


def check(sd=11, ad=3, H=128, L=3, nh=1, K=20, B=8, E=1000, act="relu", tanh=True, seed=0):
    from transformers import DecisionTransformerConfig, DecisionTransformerModel

    cfg = DecisionTransformerConfig(state_dim=sd, act_dim=ad, hidden_size=H, max_ep_len=E, action_tanh=tanh, n_layer=L, n_head=nh,
                                    n_inner=4 * H, activation_function=act, n_positions=1024, resid_pdrop=0.0, attn_pdrop=0.0,
                                    embd_pdrop=0.0)
    torch.manual_seed(seed)
    hf = DecisionTransformerModel(cfg).eval()
    hf.encoder.wpe.weight.data.zero_()
    mine = DecisionTransformer(sd, ad, H, max_length=K, max_ep_len=E, action_tanh=tanh, n_layer=L, n_head=nh, activation_function=act,
                               resid_pdrop=0.0, attn_pdrop=0.0).eval()
    mine.load_hf_state_dict(hf.state_dict())
    s, a, r, t, m = make_inputs(B, K, sd, ad, E, seed + 1)
    with torch.no_grad():
        o_hf = hf(states=s, actions=a, rewards=None, returns_to_go=r, timesteps=t, attention_mask=m, return_dict=False)
        o_me = mine(s, a, None, r, t, m)
    keep = m > 0
    errs = [(x[keep] - y[keep]).abs().max().item() for x, y in zip(o_hf[:3], o_me)]
    return max(errs)


if __name__ == "__main__":
    for kw in [dict(), dict(act="gelu_new", H=64, nh=4, L=2, K=12), dict(sd=16, ad=4, tanh=False, H=96, nh=3)]:
        e = check(**kw)
        print(kw, "max|oracle - HF| on unmasked =", e)
        assert e < 2e-5, e
    print("oracle == HF port (wpe zeroed) OK")
