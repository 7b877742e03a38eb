"""ORACLE (test infrastructure, not product code): fp32 PyTorch restatement of the Decision Transformer.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this.

PARITY STATUS: "parity unpinned by the reference".  The reference tree's `decision-transformer/` is an
empty, un-vendored git submodule (/root/reference/.SUBMODULES.json:8-16, commit e2d82e68..., named by
/root/reference/README.md:15 as kzl/decision-transformer); no reference file calls or tests it.
This restatement follows the in-container HuggingFace port
  transformers/models/decision_transformer/modeling_decision_transformer.py   ("HF" below)
with the one documented upstream difference: the GPT-2 trunk has NO positional embedding (HF keeps
`wpe` and feeds position_ids=0, HF:425-426,641; upstream removes it, comment kept at HF:514-515).
It is pinned against HF (wpe zeroed) by oracle/check_vs_hf.py and tests/test_oracle_cpu.py.

Semantics restated here:
  * token order R_1,s_1,a_1,R_2,... (HF:623-627), embed_ln (HF:628), mask expansion (HF:631-635)
  * pre-LN GPT-2 blocks, Conv1D weights stored [in, out] (HF:100,179), causal+padding mask (HF:408-414)
  * heads: return/state from action tokens, action from state tokens (HF:650-655)
  * get_action: last-K window, LEFT zero padding, mask = [0]*(K-t)+[1]*t, returns action_preds[0,-1]
  * train_step loss: mean((a_hat - a)^2) over rows with attention_mask > 0 (oracle/trainer.py)
"""
from __future__ import annotations

import math

import torch
import torch.nn as nn
import torch.nn.functional as F


class Conv1D(nn.Module):
    """y = x @ W + b with W stored [in, out] (GPT-2 convention, HF Conv1D used at HF:100-101)."""

    def __init__(self, nf: int, nx: int):
        super().__init__()
        self.weight = nn.Parameter(torch.empty(nx, nf))
        self.bias = nn.Parameter(torch.zeros(nf))
        nn.init.normal_(self.weight, std=0.02)

    def forward(self, x):
        return torch.addmm(self.bias, x.reshape(-1, x.size(-1)), self.weight).view(*x.shape[:-1], -1)


def _act(name: str):
    if name == "relu":
        return F.relu
    if name in ("gelu_new", "gelu_tanh"):
        return lambda x: F.gelu(x, approximate="tanh")
    raise ValueError(f"unsupported activation_function {name!r} (relu | gelu_new)")


class MaskedDropout(nn.Dropout):
    """nn.Dropout, except that a test may pin the keep mask (`keep`: bool tensor shaped like the input) so that this oracle and the
    CUDA path drop the SAME elements: y = x * keep / (1 - p).  Without a pinned mask it is plain nn.Dropout (torch's generator)."""

    keep = None

    def forward(self, x):
        if self.keep is None or not self.training or self.p == 0.0:
            return super().forward(x)
        return x * self.keep.to(x.dtype).view(x.shape) / (1.0 - self.p)


def set_dropout_masks(model, embd=None, attn=None, resid_attn=None, resid_mlp=None):
    """Pin the keep masks of a DecisionTransformer oracle: `embd` [B,T,H]; per-layer lists `attn` [B,nh,T,T], `resid_*` [B,T,H]."""
    tr = model.transformer
    tr.drop.keep = embd
    for l, blk in enumerate(tr.h):
        blk.attn.attn_dropout.keep = None if attn is None else attn[l]
        blk.attn.resid_dropout.keep = None if resid_attn is None else resid_attn[l]
        blk.mlp.dropout.keep = None if resid_mlp is None else resid_mlp[l]


class Attention(nn.Module):
    def __init__(self, hidden: int, n_head: int, attn_pdrop: float, resid_pdrop: float):
        super().__init__()
        assert hidden % n_head == 0
        self.n_head, self.head_dim = n_head, hidden // n_head
        self.c_attn = Conv1D(3 * hidden, hidden)
        self.c_proj = Conv1D(hidden, hidden)
        self.attn_dropout = MaskedDropout(attn_pdrop)
        self.resid_dropout = MaskedDropout(resid_pdrop)

    def forward(self, x, allowed):  # allowed: bool [B, 1, T, T]
        B, T, H = x.shape
        q, k, v = self.c_attn(x).split(H, dim=2)  # HF:179
        q = q.view(B, T, self.n_head, self.head_dim).transpose(1, 2)
        k = k.view(B, T, self.n_head, self.head_dim).transpose(1, 2)
        v = v.view(B, T, self.n_head, self.head_dim).transpose(1, 2)
        w = (q @ k.transpose(-1, -2)) * (1.0 / math.sqrt(self.head_dim))  # HF:47-52, scaling HF:89-94
        w = w.masked_fill(~allowed, torch.finfo(w.dtype).min)
        w = F.softmax(w, dim=-1)
        w = self.attn_dropout(w)
        o = (w @ v).transpose(1, 2).reshape(B, T, H)
        return self.resid_dropout(self.c_proj(o))  # HF:216-218


class MLP(nn.Module):
    def __init__(self, hidden: int, inner: int, act: str, resid_pdrop: float):
        super().__init__()
        self.c_fc = Conv1D(inner, hidden)
        self.c_proj = Conv1D(hidden, inner)
        self.act = _act(act)
        self.dropout = MaskedDropout(resid_pdrop)

    def forward(self, x):
        return self.dropout(self.c_proj(self.act(self.c_fc(x))))  # HF:233-238


class Block(nn.Module):
    def __init__(self, hidden, n_head, inner, act, attn_pdrop, resid_pdrop, eps):
        super().__init__()
        self.ln_1 = nn.LayerNorm(hidden, eps=eps)
        self.attn = Attention(hidden, n_head, attn_pdrop, resid_pdrop)
        self.ln_2 = nn.LayerNorm(hidden, eps=eps)
        self.mlp = MLP(hidden, inner, act, resid_pdrop)

    def forward(self, x, allowed):
        x = x + self.attn(self.ln_1(x), allowed)  # HF:272-281
        x = x + self.mlp(self.ln_2(x))  # HF:303-306
        return x


class GPT2Trunk(nn.Module):
    """GPT-2 stack on input embeddings, WITHOUT positional embeddings (upstream trajectory_gpt2)."""

    def __init__(self, hidden, n_layer, n_head, inner, act, attn_pdrop, resid_pdrop, embd_pdrop, eps):
        super().__init__()
        self.drop = MaskedDropout(embd_pdrop)
        self.h = nn.ModuleList([Block(hidden, n_head, inner, act, attn_pdrop, resid_pdrop, eps) for _ in range(n_layer)])
        self.ln_f = nn.LayerNorm(hidden, eps=eps)

    def forward(self, x, token_mask):  # token_mask: [B, T] (1 = valid)
        B, T, _ = x.shape
        causal = torch.ones(T, T, dtype=torch.bool, device=x.device).tril()
        allowed = causal[None, None] & (token_mask[:, None, None, :] > 0)  # A3: (j <= i) and m3[b, j]
        x = self.drop(x)
        for blk in self.h:
            x = blk(x, allowed)
        return self.ln_f(x)


class DecisionTransformer(nn.Module):
    """DecisionTransformer(state_dim, act_dim, hidden_size, max_length, max_ep_len, action_tanh, **gpt2 kwargs)."""

    def __init__(self, state_dim, act_dim, hidden_size, max_length=None, max_ep_len=4096, action_tanh=True, n_layer=3, n_head=1,
                 n_inner=None, activation_function="relu", n_positions=1024, resid_pdrop=0.1, attn_pdrop=0.1, embd_pdrop=None,
                 layer_norm_epsilon=1e-5, **unused):
        super().__init__()
        self.state_dim, self.act_dim, self.hidden_size = state_dim, act_dim, hidden_size
        self.max_length, self.max_ep_len, self.action_tanh = max_length, max_ep_len, action_tanh
        n_inner = n_inner if n_inner is not None else 4 * hidden_size
        embd_pdrop = resid_pdrop if embd_pdrop is None else embd_pdrop
        self.config = dict(state_dim=state_dim, act_dim=act_dim, hidden_size=hidden_size, max_length=max_length, max_ep_len=max_ep_len,
                           action_tanh=action_tanh, n_layer=n_layer, n_head=n_head, n_inner=n_inner,
                           activation_function=activation_function, resid_pdrop=resid_pdrop, attn_pdrop=attn_pdrop,
                           layer_norm_epsilon=layer_norm_epsilon)
        self.transformer = GPT2Trunk(hidden_size, n_layer, n_head, n_inner, activation_function, attn_pdrop, resid_pdrop, embd_pdrop,
                                     layer_norm_epsilon)
        self.embed_timestep = nn.Embedding(max_ep_len, hidden_size)
        self.embed_return = nn.Linear(1, hidden_size)
        self.embed_state = nn.Linear(state_dim, hidden_size)
        self.embed_action = nn.Linear(act_dim, hidden_size)
        self.embed_ln = nn.LayerNorm(hidden_size, eps=layer_norm_epsilon)
        self.predict_state = nn.Linear(hidden_size, state_dim)
        self.predict_action = nn.Sequential(*([nn.Linear(hidden_size, act_dim)] + ([nn.Tanh()] if action_tanh else [])))
        self.predict_return = nn.Linear(hidden_size, 1)
        self._init_weights(n_layer)

    def _init_weights(self, n_layer):  # HF:327-341 (irrelevant to parity: weights are copied in tests)
        for name, m in self.named_modules():
            if isinstance(m, (nn.Linear, Conv1D)):
                nn.init.normal_(m.weight, std=0.02)
                if m.bias is not None:
                    nn.init.zeros_(m.bias)
            elif isinstance(m, nn.Embedding):
                nn.init.normal_(m.weight, std=0.02)
            elif isinstance(m, nn.LayerNorm):
                nn.init.ones_(m.weight)
                nn.init.zeros_(m.bias)
        for name, p in self.named_parameters():
            if name.endswith("c_proj.weight"):
                nn.init.normal_(p, std=0.02 / math.sqrt(2 * n_layer))

    def forward(self, states, actions, rewards, returns_to_go, timesteps, attention_mask=None):
        B, K = states.shape[0], states.shape[1]
        if attention_mask is None:
            attention_mask = torch.ones((B, K), dtype=torch.long, device=states.device)  # HF:606-608
        te = self.embed_timestep(timesteps)  # HF:614
        s = self.embed_state(states) + te  # HF:611,617
        a = self.embed_action(actions) + te
        r = self.embed_return(returns_to_go) + te
        x = torch.stack((r, s, a), dim=1).permute(0, 2, 1, 3).reshape(B, 3 * K, self.hidden_size)  # HF:623-627
        x = self.embed_ln(x)  # HF:628
        m3 = torch.stack((attention_mask,) * 3, dim=1).permute(0, 2, 1).reshape(B, 3 * K)  # HF:631-635
        x = self.transformer(x, m3)
        x = x.reshape(B, K, 3, self.hidden_size).permute(0, 2, 1, 3)  # HF:650
        return_preds = self.predict_return(x[:, 2])  # HF:653
        state_preds = self.predict_state(x[:, 2])  # HF:654
        action_preds = self.predict_action(x[:, 1])  # HF:655
        return state_preds, action_preds, return_preds  # HF:656-657

    @torch.no_grad()
    def get_action(self, states, actions, rewards, returns_to_go, timesteps, **kwargs):
        """Upstream evaluation entry point — restated; no citable file in this container."""
        states = states.reshape(1, -1, self.state_dim)
        actions = actions.reshape(1, -1, self.act_dim)
        returns_to_go = returns_to_go.reshape(1, -1, 1)
        timesteps = timesteps.reshape(1, -1)
        if self.max_length is not None:
            K = self.max_length
            states, actions = states[:, -K:], actions[:, -K:]
            returns_to_go, timesteps = returns_to_go[:, -K:], timesteps[:, -K:]
            t = states.shape[1]
            dev = states.device
            attention_mask = torch.cat([torch.zeros(K - t), torch.ones(t)]).to(dtype=torch.long, device=dev).reshape(1, -1)
            states = torch.cat([torch.zeros((1, K - t, self.state_dim), device=dev), states], dim=1).to(torch.float32)
            actions = torch.cat([torch.zeros((1, K - t, self.act_dim), device=dev), actions], dim=1).to(torch.float32)
            returns_to_go = torch.cat([torch.zeros((1, K - t, 1), device=dev), returns_to_go], dim=1).to(torch.float32)
            timesteps = torch.cat([torch.zeros((1, K - t), device=dev), timesteps], dim=1).to(torch.long)
        else:
            attention_mask = None
        _, action_preds, _ = self.forward(states, actions, None, returns_to_go, timesteps, attention_mask=attention_mask)
        return action_preds[0, -1]

    # -- weight exchange with the HF port (cross-check) -------------------------------------------
    def load_hf_state_dict(self, sd: dict):
        """Load a transformers DecisionTransformerModel state dict (keys `encoder.*`); wte / wpe / attn.bias ignored."""
        out = {}
        for k, v in sd.items():
            if k.startswith("encoder."):
                k2 = "transformer." + k[len("encoder."):]
            else:
                k2 = k
            if ".wte." in k2 or ".wpe." in k2 or k2.endswith(".attn.bias") or k2.endswith(".attn.masked_bias"):
                continue
            out[k2] = v
        missing, unexpected = self.load_state_dict(out, strict=False)
        assert not missing and not unexpected, (missing, unexpected)
