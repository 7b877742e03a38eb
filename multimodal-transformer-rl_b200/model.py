"""Host-side mirror of the reference ``DecisionTransformer`` over the sm_100a kernels.

API kept from the reference (kzl decision-transformer, restated by HF ``modeling_decision_transformer.py:502-666``):
``DecisionTransformer(state_dim, act_dim, hidden_size, max_length, max_ep_len, action_tanh, **gpt2_kwargs)``,
``.forward(states, actions, rewards, returns_to_go, timesteps, attention_mask) -> (state_preds, action_preds, return_preds)``,
``.get_action(states, actions, rewards, returns_to_go, timesteps)``, ``state_dict`` keys of the upstream model
(``transformer.h.{i}...``; HF's ``encoder.*`` prefix is accepted on load).

All parameters are views into ONE flat fp32 buffer (layout defined by ``dtb_dt_param_layout``) with a bf16 shadow the
tensor-core GEMMs read and a flat gradient buffer the backward writes — the reference's "flatten all grads"
data-parallel idiom (``cleanrl/ppo_atari_multigpu.py:345-359``) without the copies.  Compute is only ever done by
``libdtb200.so``; there is no PyTorch/CPU fallback (calling ``forward`` on a CPU tensor raises).
"""
from __future__ import annotations

import ctypes as C
import warnings

import torch
import torch.nn as nn

from . import _lib

_ACT = {"relu": 1, "gelu_new": 2, "gelu_tanh": 2, "gelu_pytorch_tanh": 2}


def _entry_names(n_layer: int):
    names = ["embed_timestep.weight", "embed_return.weight", "embed_return.bias", "embed_state.weight", "embed_state.bias",
             "embed_action.weight", "embed_action.bias", "embed_ln.weight", "embed_ln.bias"]
    for i in range(n_layer):
        p = f"transformer.h.{i}."
        names += [p + "ln_1.weight", p + "ln_1.bias", p + "attn.c_attn.weight", p + "attn.c_attn.bias", p + "attn.c_proj.weight",
                  p + "attn.c_proj.bias", p + "ln_2.weight", p + "ln_2.bias", p + "mlp.c_fc.weight", p + "mlp.c_fc.bias",
                  p + "mlp.c_proj.weight", p + "mlp.c_proj.bias"]
    names += ["transformer.ln_f.weight", "transformer.ln_f.bias", "predict_state.weight", "predict_return.weight",
              "predict_action.0.weight", "predict_state.bias", "predict_return.bias", "predict_action.0.bias"]
    return names


def _entry_shapes(sd, ad, H, I, E, n_layer):
    shapes = [(E, H), (H, 1), (H,), (H, sd), (H,), (H, ad), (H,), (H,), (H,)]
    for _ in range(n_layer):
        shapes += [(H,), (H,), (H, 3 * H), (3 * H,), (H, H), (H,), (H,), (H,), (H, I), (I,), (I, H), (H,)]
    shapes += [(H,), (H,), (sd, H), (1, H), (ad, H), (sd,), (1,), (ad,)]
    return shapes


class _Node(nn.Module):
    """Parameter container used to reproduce the reference's dotted state_dict names."""


def _set_nested(root: nn.Module, dotted: str, param: nn.Parameter):
    parts = dotted.split(".")
    mod = root
    for p in parts[:-1]:
        if p not in mod._modules:
            mod.add_module(p, _Node())
        mod = mod._modules[p]
    mod.register_parameter(parts[-1], param)


class DecisionTransformer(nn.Module):
    def __init__(self, state_dim, act_dim, hidden_size, max_length=None, max_ep_len=4096, action_tanh=True, n_layer=3, n_head=1,
                 n_inner=None, activation_function="relu", n_positions=1024, resid_pdrop=0.1, attn_pdrop=0.1, embd_pdrop=None,
                 layer_norm_epsilon=1e-5, device=None, **unused):
        super().__init__()
        if activation_function not in _ACT:
            raise ValueError(f"unsupported activation_function {activation_function!r} (relu | gelu_new)")
        self.state_dim, self.act_dim, self.hidden_size = int(state_dim), int(act_dim), int(hidden_size)
        self.max_length, self.max_ep_len, self.action_tanh = max_length, int(max_ep_len), bool(action_tanh)
        n_inner = int(n_inner) if n_inner is not None else 4 * self.hidden_size
        self.config = dict(state_dim=self.state_dim, act_dim=self.act_dim, hidden_size=self.hidden_size, max_length=max_length,
                           max_ep_len=self.max_ep_len, action_tanh=self.action_tanh, n_layer=int(n_layer), n_head=int(n_head),
                           n_inner=n_inner, activation_function=activation_function, resid_pdrop=resid_pdrop, attn_pdrop=attn_pdrop,
                           layer_norm_epsilon=layer_norm_epsilon)
        # nn.Dropout sites of the trunk (GPT2Model.drop / attn_dropout / resid_dropout); embd_pdrop follows resid_pdrop when not given
        self._p_resid, self._p_attn = float(resid_pdrop or 0.0), float(attn_pdrop or 0.0)
        self._p_embd = self._p_resid if embd_pdrop is None else float(embd_pdrop)
        # get_action(use_cache=True) / kv_cache = True: append to a per-layer K/V cache while the history fits the window (exact, see _decode).
        # None = automatic: on where the decode step runs on the tensor core (hidden_size 128: c1 / c2 — measured faster than the graph of
        # the full B = 1 forward), off where only the CUDA-core kernel applies (c3 / c4: measured slower).
        self.kv_cache = None
        self._dc = None               # decode state: cache buffer, timesteps cached, parameter stamp, output buffer
        self._drop_state = None       # device int32 [3] = {seed_lo, seed_hi, step}, read by the kernels at run time
        self._drop_applied = None     # (p_embd, p_attn, p_resid, state ptr) last handed to the library

        lib = _lib.load()  # raises when libdtb200.so is missing: there is no other implementation
        cfg = _lib.DtConfig(self.state_dim, self.act_dim, self.hidden_size, int(n_layer), int(n_head), n_inner, self.max_ep_len,
                            _ACT[activation_function], int(self.action_tanh), float(layer_norm_epsilon))
        handle = C.c_void_p()
        _lib.check(lib.dtb_dt_create(C.byref(cfg), C.byref(handle)))
        self._h = handle
        n = lib.dtb_dt_param_entries(self._h)
        offs, nums = (C.c_longlong * n)(), (C.c_longlong * n)()
        self._total = int(lib.dtb_dt_param_layout(self._h, offs, nums))
        self._names = _entry_names(int(n_layer))
        self._shapes = _entry_shapes(self.state_dim, self.act_dim, self.hidden_size, n_inner, self.max_ep_len, int(n_layer))
        assert len(self._names) == n == len(self._shapes)
        self._offsets = [int(o) for o in offs]
        for shp, num in zip(self._shapes, nums):
            assert int(torch.Size(shp).numel()) == int(num), (shp, num)

        if device is None:
            device = "cuda" if torch.cuda.is_available() else "cpu"
        self._flat = torch.zeros(self._total, dtype=torch.float32, device=device)
        self._flat_grad = None
        self._flat_bf16 = None
        self._staged_version = -1
        self._ws = {}
        self._ws_gen = 0
        self._bind_views(register=True)
        self.reset_parameters()

    # ---- parameter plumbing --------------------------------------------------------------------------------------
    def _bind_views(self, register: bool):
        for name, shp, off in zip(self._names, self._shapes, self._offsets):
            view = self._flat[off:off + torch.Size(shp).numel()].view(shp)
            if register:
                _set_nested(self, name, nn.Parameter(view))
            else:
                self.get_parameter(name).data = view
        self._plist = [self.get_parameter(n) for n in self._names]
        self._staged_version = -1

    def _version_now(self) -> int:
        """Staleness stamp of the fp32 master.  After ``.to()`` / ``.cuda()`` torch rebinds every Parameter with ``param.data = view``,
        which gives each of them its own version counter: an in-place update through a Parameter (``torch.optim.AdamW``) then no longer
        bumps ``_flat._version``, so the stamp is the sum over the flat buffer AND every Parameter (all counters only ever grow)."""
        return self._flat._version + sum(p._version for p in self._plist)

    def _apply(self, fn, recurse=True):
        # .to()/.cuda() rebuild every Parameter separately; gather them back into one flat buffer afterwards
        super()._apply(fn, recurse)
        ref = self.get_parameter(self._names[0])
        flat = torch.zeros(self._total, dtype=torch.float32, device=ref.device)
        for name, shp, off in zip(self._names, self._shapes, self._offsets):
            flat[off:off + torch.Size(shp).numel()].view(shp).copy_(self.get_parameter(name).data.float())
        self._flat = flat
        self._flat_grad, self._flat_bf16, self._ws = None, None, {}
        self._bind_views(register=False)
        return self

    @torch.no_grad()
    def reset_parameters(self):
        """Initialisation of the upstream model the reference names (kzl decision-transformer): only the GPT-2 trunk (``transformer.*``)
        gets the GPT-2 init (HF modeling_gpt2 ``_init_weights``: N(0, 0.02), c_proj scaled by 1/sqrt(2L), LayerNorm (1, 0), zero biases);
        the embedding / prediction layers outside it keep torch's defaults — ``nn.Linear`` kaiming-uniform(a=sqrt 5) weights and
        U(-1/sqrt(fan_in), 1/sqrt(fan_in)) biases, ``nn.Embedding`` N(0, 1), ``nn.LayerNorm`` (1, 0)."""
        L = self.config["n_layer"]
        for name, p in self.named_parameters():
            if name.startswith("transformer."):
                if name.endswith(".bias"):
                    p.zero_()
                elif ".ln_" in name:
                    p.fill_(1.0)
                elif name.endswith("c_proj.weight"):
                    p.normal_(0.0, 0.02 / (2 * L) ** 0.5)
                else:
                    p.normal_(0.0, 0.02)
            elif name == "embed_timestep.weight":
                p.normal_(0.0, 1.0)
            elif name.startswith("embed_ln."):
                p.fill_(1.0) if name.endswith("weight") else p.zero_()
            elif name.endswith(".weight"):
                w = torch.empty(p.shape, dtype=torch.float32)
                nn.init.kaiming_uniform_(w, a=5 ** 0.5)
                p.copy_(w)
            else:
                fan_in = self.get_parameter(name[:-4] + "weight").shape[1]
                p.copy_(torch.empty(p.shape).uniform_(-1.0 / fan_in ** 0.5, 1.0 / fan_in ** 0.5))

    def load_state_dict(self, state_dict, strict=True, assign=False):
        """Accepts upstream (``transformer.*``) and HF (``encoder.*``) keys; wte / wpe / attn.bias buffers are ignored (A1)."""
        sd = {}
        for k, v in state_dict.items():
            k2 = "transformer." + k[len("encoder."):] if k.startswith("encoder.") else k
            if ".wte." in k2 or ".wpe." in k2 or k2.endswith(".attn.bias") or k2.endswith(".attn.masked_bias"):
                continue
            sd[k2] = v
        out = super().load_state_dict(sd, strict=strict, assign=False)
        self._staged_version = -1
        return out

    @property
    def flat_params(self) -> torch.Tensor:
        return self._flat

    @property
    def flat_grads(self) -> torch.Tensor:
        """Flat fp32 gradient buffer (same layout as flat_params); every ``param.grad`` is a view into it."""
        if self._flat_grad is None or self._flat_grad.device != self._flat.device:
            self._flat_grad = torch.zeros_like(self._flat)
        return self._flat_grad

    def attach_grads(self):
        g = self.flat_grads
        for name, shp, off in zip(self._names, self._shapes, self._offsets):
            p = self.get_parameter(name)
            view = g[off:off + torch.Size(shp).numel()].view(shp)
            if p.grad is None or p.grad.data_ptr() != view.data_ptr():
                p.grad = view
        return g

    def stage_grad_range(self, stage: int):
        b, e = C.c_longlong(), C.c_longlong()
        _lib.check(_lib.load().dtb_dt_stage_grad_range(self._h, stage, C.byref(b), C.byref(e)))
        return int(b.value), int(e.value)

    @property
    def n_stages(self) -> int:
        return self.config["n_layer"] + 2

    # ---- device-side plumbing ------------------------------------------------------------------------------------
    def _require_cuda(self, t: torch.Tensor):
        if not t.is_cuda or not self._flat.is_cuda:
            raise _lib.DtbError("DecisionTransformer runs only on an sm_100a CUDA device through libdtb200.so "
                                "(no CPU / eager-PyTorch fallback exists); move the model and inputs to cuda")

    def _stage(self):
        """Refresh the bf16 shadow of the parameters if any of them changed since the last staging."""
        if self._flat_bf16 is None or self._flat_bf16.device != self._flat.device:
            self._flat_bf16 = torch.empty(self._total, dtype=torch.bfloat16, device=self._flat.device)
            self._staged_version = -1
        ver = self._version_now()
        if self._staged_version != ver:
            _lib.check(_lib.load().dtb_dt_stage_params(self._h, _lib.stream_ptr(), _lib.ptr(self._flat), _lib.ptr(self._flat_bf16)))
            self._staged_version = ver

    def restage(self):
        """Force a refresh of the bf16 shadow (call after writing parameters through an aliasing view torch cannot track)."""
        self._staged_version = -1
        self._stage()

    def _workspace(self, B: int, K: int) -> torch.Tensor:
        key = (B, K)
        ws = self._ws.get(key)
        if ws is None:
            nbytes = int(_lib.load().dtb_dt_workspace_bytes(self._h, B, K))
            if len(self._ws) > 4:
                self._ws.clear()
            ws = torch.empty(nbytes, dtype=torch.uint8, device=self._flat.device)
            self._ws[key] = ws
        return ws

    def _make_batch(self, states, actions, returns_to_go, timesteps, attention_mask, states_bf16=None, states_relu=False):
        B, K = actions.shape[0], actions.shape[1]
        keep = []

        def f32(t):
            t = t.detach().to(torch.float32).contiguous()
            keep.append(t)
            return t

        b = _lib.DtBatch()
        b.B, b.K = B, K
        b.states_relu = int(bool(states_relu))
        if states_bf16 is not None:
            sb = states_bf16.detach().contiguous()
            assert sb.dtype == torch.bfloat16
            keep.append(sb)
            b.states_bf16 = sb.data_ptr()
            b.states = None
        else:
            b.states = f32(states).data_ptr()
            b.states_bf16 = None
        b.actions = f32(actions).data_ptr()
        b.returns_to_go = f32(returns_to_go.reshape(B, K)).data_ptr()
        ts = timesteps.detach().to(torch.long).contiguous()
        keep.append(ts)
        b.timesteps = ts.data_ptr()
        if attention_mask is not None:
            b.attention_mask = f32(attention_mask.reshape(B, K)).data_ptr()
        else:
            b.attention_mask = None
        return b, keep

    def _check_shapes(self, states, actions, returns_to_go, timesteps, attention_mask):
        B, K = actions.shape[0], actions.shape[1]
        if states.shape != (B, K, self.state_dim):
            raise ValueError(f"states must be [B, K, {self.state_dim}], got {tuple(states.shape)}")
        if actions.shape[2] != self.act_dim:
            raise ValueError(f"actions must be [B, K, {self.act_dim}], got {tuple(actions.shape)}")
        if returns_to_go.numel() != B * K or timesteps.numel() != B * K:
            raise ValueError("returns_to_go must be [B, K, 1] and timesteps [B, K]")
        if attention_mask is not None and attention_mask.numel() != B * K:
            raise ValueError("attention_mask must be [B, K]")

    # ---- dropout --------------------------------------------------------------------------------------------------
    def set_dropout_seed(self, seed: int, step: int = 0):
        """(Re)seed the in-kernel dropout generator.  Replicas of a data-parallel job pass different seeds (independent masks)."""
        seed = int(seed) & 0xFFFFFFFFFFFFFFFF
        vals = [seed & 0xFFFFFFFF, seed >> 32, int(step) & 0xFFFFFFFF]
        vals = [v - (1 << 32) if v >= (1 << 31) else v for v in vals]
        t = torch.tensor(vals, dtype=torch.int32)
        if self._drop_state is None or self._drop_state.device != self._flat.device:
            self._drop_state = t.to(self._flat.device)
        else:
            self._drop_state.copy_(t)  # in place: captured graphs keep reading the same buffer

    def dropout_active(self) -> bool:
        return bool(self.training and max(self._p_embd, self._p_attn, self._p_resid) > 0.0)

    def _sync_dropout(self) -> bool:
        """Hand the current mode (train: the three probabilities, eval: off) to the library; True when dropout is on."""
        on = self.dropout_active()
        if on and (self._drop_state is None or self._drop_state.device != self._flat.device):
            self.set_dropout_seed(torch.initial_seed())
        want = (self._p_embd, self._p_attn, self._p_resid, self._drop_state.data_ptr()) if on else (0.0, 0.0, 0.0, 0)
        if want != self._drop_applied:
            _lib.check(_lib.load().dtb_dt_set_dropout(self._h, want[0], want[1], want[2], want[3] or None))
            self._drop_applied = want
        return on

    def dropout_keep_mask(self, site: int, p: float, n: int, groups: int = 0) -> torch.Tensor:
        """Keep bits (uint8) of one dropout site at the generator's CURRENT step, exactly as the kernels draw them (tests / oracle)."""
        total = n * max(groups, 1)
        out = torch.empty(total, dtype=torch.uint8, device=self._flat.device)
        _lib.check(_lib.load().dtb_dropout_mask(_lib.stream_ptr(), _lib.ptr(self._drop_state), int(site), float(p), int(groups), int(n),
                                                _lib.ptr(out)))
        return out

    # ---- reference API -------------------------------------------------------------------------------------------
    def forward(self, states, actions, rewards, returns_to_go, timesteps, attention_mask=None):
        """HF:536-657.  ``rewards`` is accepted and ignored, as in the reference (A8)."""
        self._require_cuda(actions)
        self._check_shapes(states, actions, returns_to_go, timesteps, attention_mask)
        params = [self.get_parameter(n) for n in self._names]
        wide = self.state_dim > 64 and states.dtype == torch.bfloat16
        return _DTFunction.apply(self, states, actions, returns_to_go, timesteps, attention_mask, wide, *params)

    # ---- KV-cached get_action ------------------------------------------------------------------------
    def reset_kv_cache(self):
        """Forget the cached keys / values (call when earlier entries of the history were edited in place; a shorter history, i.e. a
        new episode, and a parameter update reset it automatically)."""
        if self._dc is not None:
            self._dc["t"] = 0

    def _decode_ready(self) -> bool:
        return bool(self.max_length is not None and not self.dropout_active() and self._decode_support())

    def _decode_support(self) -> int:
        """dtb_dt_decode_supported of this configuration (0 no, 1 CUDA-core kernel, 2 tensor-core kernel); a property of the config: asked once."""
        v = self.__dict__.get("_dc_support")
        if v is None:
            v = self.__dict__["_dc_support"] = int(_lib.load().dtb_dt_decode_supported(self._h))
        return v

    def _use_cache(self, use_cache) -> bool:
        """use_cache argument > kv_cache attribute > automatic (on where dtb_dt_decode_supported reports the tensor-core kernel: 2)."""
        choice = self.kv_cache if use_cache is None else use_cache
        if choice is None:
            return bool(self.max_length is not None and not self.dropout_active() and self._decode_support() == 2)
        return bool(choice) and self._decode_ready()

    def _step_runner(self, t: int, encoder=None, frames_dtype=None):
        """The graph-replay runner when this call is the next step of a rollout the cache already holds (t - 1 timesteps cached, weights
        unchanged); None when the eager ``_decode`` path has to (re)build the cache first."""
        dc = self._dc
        if dc is None or dc["ptr"] != self._flat.data_ptr() or dc["stamp"] != self._staged_version or dc["t"] != t - 1 or t > self.max_length:
            return None
        r = self.__dict__.get("_step")
        if r is None or r.dead(self) or r.enc is not encoder or (encoder is not None and r.frames.dtype != frames_dtype):
            r = self.__dict__["_step"] = _StepRunner(self, encoder, frames_dtype)
        r.arm(t)
        return r

    def _decode(self, states, states_bf16, actions, returns_to_go, timesteps, t: int):
        """One ``dtb_dt_decode`` call: the tokens of timesteps ``t_cached .. t-1`` go through a single cluster kernel that attends to the
        cached keys / values of everything before.  Exact while ``t <= max_length``: without a positional embedding and with causal
        attention the hidden states of earlier tokens never change, so appending equals the reference's left-padded recompute."""
        lib, K = _lib.load(), self.max_length
        dc = self._dc
        if dc is None or dc["ptr"] != self._flat.data_ptr():
            nbytes = int(lib.dtb_dt_decode_cache_bytes(self._h, K))
            dc = self._dc = dict(cache=torch.zeros(nbytes, dtype=torch.uint8, device=self._flat.device), t=0, stamp=-1, ptr=self._flat.data_ptr(),
                                 out=torch.zeros(self.act_dim, dtype=torch.float32, device=self._flat.device))
        self._stage()
        if dc["stamp"] != self._staged_version:  # weights changed: rebuild the decode layout, forget the cached keys / values
            _lib.check(lib.dtb_dt_decode_stage(self._h, _lib.stream_ptr(), _lib.ptr(self._flat_bf16), K, _lib.ptr(dc["cache"]), dc["cache"].numel()))
            dc["t"], dc["stamp"] = 0, self._staged_version
        if t <= dc["t"]:  # a new episode / repeated call: start over
            dc["t"] = 0
        _lib.check(lib.dtb_dt_decode(self._h, _lib.stream_ptr(), _lib.ptr(self._flat), _lib.ptr(self._flat_bf16), _lib.ptr(states),
                                     _lib.ptr(states_bf16), _lib.ptr(actions), _lib.ptr(returns_to_go), _lib.ptr(timesteps), t, dc["t"], K,
                                     _lib.ptr(dc["cache"]), dc["cache"].numel(), _lib.ptr(dc["out"])))
        dc["t"] = t
        return dc["out"].clone()

    @staticmethod
    def _rows(x, width, dtype):
        x = x.detach()
        if x.dtype != dtype:
            x = x.to(dtype)
        x = x.reshape(-1, width) if width else x.reshape(-1)
        return x if x.is_contiguous() else x.contiguous()

    @torch.no_grad()
    def get_action(self, states, actions, rewards, returns_to_go, timesteps, use_cache=None, **kwargs):
        """Upstream evaluation entry point: last-K window, LEFT zero padding, action of the last state token.

        While the history fits the window (t <= ``max_length``) and ``kv_cache`` is on, the call appends to a per-layer key / value
        cache: ONE kernel launch per rollout step (``_decode``).  Otherwise (sliding window, training-mode dropout, unsupported widths,
        ``use_cache=False``) it is one ``dtb_pad_left`` launch (window + padding + mask, integer-exact) plus the replay of a CUDA graph
        of the B = 1 forward over static history buffers (``_ActionRunner``)."""
        self._require_cuda(states)
        if self.state_dim <= 64 and self._use_cache(use_cache):
            t = states.numel() // self.state_dim
            if t <= self.max_length:
                self._stage()
                r = self._step_runner(t)
                if r is not None:  # (the runner converts dtype / layout only when it has to: no per-call views on the host)
                    return r.step([(states, r.states), (actions, r.actions), (returns_to_go, r.rtg), (timesteps, r.timesteps)], t)
                s = self._rows(states, self.state_dim, torch.float32)
                return self._decode(s, None, self._rows(actions, self.act_dim, torch.float32), self._rows(returns_to_go, 0, torch.float32),
                                    self._rows(timesteps, 0, torch.long), s.shape[0])
        if self.max_length is None:
            states = states.reshape(1, -1, self.state_dim)
            actions = actions.reshape(1, -1, self.act_dim)
            _, action_preds, _ = self._forward_raw(states, actions, returns_to_go.reshape(1, -1, 1), timesteps.reshape(1, -1), None,
                                                   want_all=False)
            return action_preds[0, -1]
        K = self.max_length
        runner = self.__dict__.get("_action_runner")
        if runner is None or runner.dead(self):
            runner = _ActionRunner(self, None)
            self.__dict__["_action_runner"] = runner
        return runner.run([(states, runner.states), (actions, runner.actions), (returns_to_go, runner.rtg), (timesteps, runner.timesteps)])

    # ---- raw paths (no autograd graph) ---------------------------------------------------------------------------
    def _forward_raw(self, states, actions, returns_to_go, timesteps, attention_mask, want_all=True, states_bf16=None, states_relu=False):
        lib = _lib.load()
        B, K = actions.shape[0], actions.shape[1]
        self._stage()
        if self._sync_dropout():  # training mode with p > 0: one fresh set of masks per forward (its backward re-draws the same bits)
            _lib.check(lib.dtb_dropout_advance(_lib.stream_ptr(), _lib.ptr(self._drop_state)))
        batch, keep = self._make_batch(states, actions, returns_to_go, timesteps, attention_mask, states_bf16, states_relu)
        ws = self._workspace(B, K)
        dev = self._flat.device
        ap = torch.empty((B, K, self.act_dim), dtype=torch.float32, device=dev)
        sp = torch.empty((B, K, self.state_dim), dtype=torch.float32, device=dev) if want_all else None
        rp = torch.empty((B, K, 1), dtype=torch.float32, device=dev) if want_all else None
        _lib.check(lib.dtb_dt_forward(self._h, _lib.stream_ptr(), _lib.ptr(self._flat), _lib.ptr(self._flat_bf16), C.byref(batch),
                                      _lib.ptr(ws), ws.numel(), _lib.ptr(ap), _lib.ptr(sp), _lib.ptr(rp)))
        self._ws_gen += 1
        self._last = (batch, keep, ws, self._ws_gen)
        return sp, ap, rp

    def _backward_raw(self, saved, d_action, d_state, d_return, grads, d_states_bf16=None, stage_begin=0, stage_end=None):
        lib = _lib.load()
        batch, keep, ws, gen = saved
        if gen != self._ws_gen:
            raise _lib.DtbError("DecisionTransformer.backward: the activation workspace was overwritten by a later forward; "
                                "run backward before the next forward of the same model and batch shape")
        if stage_end is None:
            stage_end = self.n_stages - 1

        def f32(t):
            if t is None:
                return None
            t = t.detach().to(torch.float32).contiguous()
            keep.append(t)
            return t

        _lib.check(lib.dtb_dt_backward(self._h, _lib.stream_ptr(), _lib.ptr(self._flat), _lib.ptr(self._flat_bf16), C.byref(batch),
                                       _lib.ptr(ws), ws.numel(), _lib.ptr(f32(d_action)), _lib.ptr(f32(d_state)), _lib.ptr(f32(d_return)),
                                       _lib.ptr(grads), _lib.ptr(d_states_bf16), stage_begin, stage_end))

    def loss_and_backward(self, states, actions, returns_to_go, timesteps, attention_mask, action_target=None, grad_scale=1.0,
                          states_bf16=None, d_states_bf16=None, states_relu=False, on_stage_done=None):
        """Fused training step body: forward + masked action MSE (A10) + backward straight into ``flat_grads``.

        Returns a device tensor ``[loss, n_valid_rows]`` (no host sync).  ``on_stage_done(stage, begin, end)`` is called after each
        backward stage is enqueued (data-parallel hosts all-reduce ``flat_grads[begin:end]`` there)."""
        lib = _lib.load()
        self._require_cuda(actions)
        target = (actions if action_target is None else action_target).detach().to(torch.float32).contiguous()
        self._forward_raw(states, actions, returns_to_go, timesteps, attention_mask, want_all=False, states_bf16=states_bf16,
                          states_relu=states_relu)
        batch, keep, ws, _ = self._last
        loss = torch.empty(2, dtype=torch.float32, device=self._flat.device)
        _lib.check(lib.dtb_dt_loss(self._h, _lib.stream_ptr(), C.byref(batch), _lib.ptr(target), float(grad_scale), _lib.ptr(ws), ws.numel(),
                                   _lib.ptr(loss)))
        grads = self.attach_grads()
        if on_stage_done is None:
            self._backward_raw(self._last, None, None, None, grads, d_states_bf16)
        else:
            for s in range(self.n_stages):
                self._backward_raw(self._last, None, None, None, grads, d_states_bf16, s, s)
                on_stage_done(s, *self.stage_grad_range(s))
        return loss

    def __del__(self):
        try:
            if getattr(self, "_h", None) is not None and self._h.value:
                _lib.load().dtb_dt_destroy(self._h)
                self._h = None
        except Exception:
            pass


class _StepRunner:
    """Steady-state rollout step of the KV-cached get_action: ONE ``dtb_pad_left`` launch (the last two rows of every per-timestep
    history -> static 2-row windows; the only launch that reads the caller's tensors) + the replay of ONE CUDA graph holding the towers
    on the newest frame (multimodal model) and ``dtb_dt_decode_step`` (one kernel; the step count lives in device memory and is
    advanced by the kernel, so the graph never changes).  Anything else — first call, new episode, changed weights, a skipped step —
    goes through the eager ``_decode`` path, after which ``arm`` re-synchronises the device-side count."""

    def __init__(self, dt, encoder=None, frames_dtype=None):
        dev = dt._flat.device
        self.dt, self.enc, self.dev = dt, encoder, dev
        self.flat_ptr, self.cache_ptr = dt._flat.data_ptr(), dt._dc["cache"].data_ptr()
        self.actions = torch.zeros((1, 2, dt.act_dim), dtype=torch.float32, device=dev)
        self.rtg = torch.zeros((1, 2, 1), dtype=torch.float32, device=dev)
        self.timesteps = torch.zeros((1, 2), dtype=torch.long, device=dev)
        self.mask = torch.zeros((1, 2), dtype=torch.float32, device=dev)
        self.states = self.frames = self.audio = None
        if encoder is None:
            self.states = torch.zeros((1, 2, dt.state_dim), dtype=torch.float32, device=dev)
        else:
            self.frames = torch.zeros((1, 2, *encoder.frame_shape), dtype=frames_dtype, device=dev)
            self.audio = torch.zeros((1, 2, *encoder.audio_shape), dtype=torch.float32, device=dev)
        self.t_dev = torch.zeros(1, dtype=torch.int32, device=dev)
        self.t_next = -1  # host mirror of *t_dev (the history length the next step expects); -1 = unknown
        self.graph, self._keep = None, None

    def dead(self, dt):
        return self.flat_ptr != dt._flat.data_ptr() or dt._dc is None or self.cache_ptr != dt._dc["cache"].data_ptr()

    def arm(self, t_next: int):
        if self.t_next != t_next:
            self.t_dev.fill_(t_next)
            self.t_next = t_next

    def _body(self):
        dt, lib = self.dt, _lib.load()
        dc, feat = dt._dc, None
        if self.enc is not None:
            feat = self.enc.forward_raw(self.frames[:, 1:2], self.audio[:, 1:2])  # towers on the newest frame only
        _lib.check(lib.dtb_dt_decode_step(dt._h, _lib.stream_ptr(), _lib.ptr(dt._flat), _lib.ptr(dt._flat_bf16), _lib.ptr(self.states),
                                          _lib.ptr(feat), _lib.ptr(self.actions), _lib.ptr(self.rtg), _lib.ptr(self.timesteps),
                                          _lib.ptr(self.t_dev), dt.max_length, _lib.ptr(dc["cache"]), dc["cache"].numel(), _lib.ptr(dc["out"])))
        self._keep = (feat, None if self.enc is None else self.enc._last)

    def step(self, hist, t: int):
        """hist: [(source tensor with one row per timestep of the whole t-step history, static 2-row window)]; requires *t_dev == t."""
        lib = _lib.load()
        n = len(hist)
        if "_dsts" not in self.__dict__:
            self._dsts = (C.c_void_p * n)(*[d.data_ptr() for _, d in hist])
            self._row_elems = [d[0, 0].numel() for _, d in hist]
            self._rows = (C.c_int * n)(*[d[0, 0].numel() * d.element_size() for _, d in hist])
            self._srcs = (C.c_void_p * n)()
        srcs, keep = self._srcs, []
        for i, (src, dst) in enumerate(hist):
            if src.dtype != dst.dtype or not src.is_contiguous() or src.requires_grad:
                src = src.detach().to(dst.dtype).contiguous()
                keep.append(src)
            if src.numel() // self._row_elems[i] != t:
                raise ValueError("get_action: states / actions / returns_to_go / timesteps must have the same history length")
            srcs[i] = src.data_ptr() + (max(0, t - 2) * self._rows[i])
        _lib.check(lib.dtb_pad_left(_lib.stream_ptr(), n, srcs, self._dsts, self._rows, min(t, 2), 2, _lib.ptr(self.mask)))
        if self.graph is None:
            self._body()  # this call's step runs eagerly (it is also the warm-up: lazy allocations, tensor-map cache) ...
            out = self.dt._dc["out"].clone()
            torch.cuda.current_stream(self.dev).synchronize()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):  # ... and is then captured (capture does not execute) for every later step
                self._body()
            self.graph = g
        else:
            self.graph.replay()
            out = self.dt._dc["out"].clone()
        self.t_next = t + 1
        self.dt._dc["t"] = t
        return out


class _ActionRunner:
    """Static K-step history buffers + one CUDA graph of the B = 1 forward (towers included for the multimodal model).

    ``run`` = ``dtb_pad_left`` (the only launch that reads the caller's tensors: keep the last t <= K rows, left-pad with zeros,
    mask = [0]*(K-t) + [1]*t) + graph replay + a clone of the action row.  Weights are re-staged to bf16 outside the graph whenever
    the fp32 master changed, so the graph stays valid across optimizer steps."""

    def __init__(self, dt, encoder=None, frames_dtype=None):
        dev, K = dt._flat.device, dt.max_length
        self.dt, self.enc, self.K, self.dev = dt, encoder, K, dev
        self.flat_ptr = dt._flat.data_ptr()
        self.actions = torch.zeros((1, K, dt.act_dim), dtype=torch.float32, device=dev)
        self.rtg = torch.zeros((1, K, 1), dtype=torch.float32, device=dev)
        self.timesteps = torch.zeros((1, K), dtype=torch.long, device=dev)
        self.mask = torch.zeros((1, K), dtype=torch.float32, device=dev)
        if encoder is None:
            self.states = torch.zeros((1, K, dt.state_dim), dtype=torch.float32, device=dev)
        else:
            self.frames = torch.zeros((1, K, *encoder.frame_shape), dtype=frames_dtype, device=dev)
            self.audio = torch.zeros((1, K, *encoder.audio_shape), dtype=torch.float32, device=dev)
        self.graph, self._keep = None, None
        self.drop_mode = dt.dropout_active()

    def dead(self, dt):
        # the model was moved / re-created, or train() / eval() changed the dropout mode: buffers and graph are stale
        return self.flat_ptr != dt._flat.data_ptr() or self.drop_mode != dt.dropout_active()

    def _body(self):
        dt = self.dt
        if self.enc is None:
            _, ap, _ = dt._forward_raw(self.states, self.actions, self.rtg, self.timesteps, self.mask, want_all=False)
        else:
            sb = self.enc.forward_raw(self.frames, self.audio)
            _, ap, _ = dt._forward_raw(None, self.actions, self.rtg, self.timesteps, self.mask, want_all=False, states_bf16=sb)
        self.ap_last = ap[0, -1]  # allocated from the graph's private pool during capture: stays valid (and is rewritten) on every replay
        self._keep = (dt._last, None if self.enc is None else self.enc._last)  # workspaces stay alive as long as the graph does

    def run(self, hist):
        """hist: [(source tensor with one row per timestep of the WHOLE history, static destination buffer)].  The last-K window is taken
        by pointer offset (no slicing / reshaping on the host: every torch view costs microseconds on a 200 us call)."""
        lib = _lib.load()
        n = len(hist)
        if "_dsts" not in self.__dict__:
            self._dsts = (C.c_void_p * n)(*[d.data_ptr() for _, d in hist])
            self._row_elems = [d[0, 0].numel() for _, d in hist]
            self._rows = (C.c_int * n)(*[d[0, 0].numel() * d.element_size() for _, d in hist])
            self._srcs = (C.c_void_p * n)()
        srcs, keep, t = self._srcs, [], -1
        for i, (src, dst) in enumerate(hist):
            if src.dtype != dst.dtype or not src.is_contiguous() or src.requires_grad:
                src = src.detach().to(dst.dtype).contiguous()
                keep.append(src)
            rows = src.numel() // self._row_elems[i]
            if t < 0:
                t = rows
            elif rows != t:
                raise ValueError("get_action: states / actions / returns_to_go / timesteps must have the same history length")
            srcs[i] = src.data_ptr() + (max(0, rows - self.K) * self._rows[i])
        t = min(t, self.K)
        self.dt._stage()
        if self.enc is not None:
            self.enc._stage()
        _lib.check(lib.dtb_pad_left(_lib.stream_ptr(), n, srcs, self._dsts, self._rows, t, self.K, _lib.ptr(self.mask)))
        if self.graph is None:
            self._body()  # warm-up (lazy allocations, tensor-map cache) outside the capture
            torch.cuda.current_stream(self.dev).synchronize()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                self._body()
            self.graph = g
        self.graph.replay()
        return self.ap_last.clone()  # the caller's copy: the static buffer is overwritten by the next call


class _DTFunction(torch.autograd.Function):
    """Generic autograd bridge: ``loss.backward()`` populates every ``param.grad`` exactly as it does for the oracle."""

    @staticmethod
    def forward(ctx, model, states, actions, returns_to_go, timesteps, attention_mask, wide, *params):
        sb = states if wide else None
        sp, ap, rp = model._forward_raw(None if wide else states, actions, returns_to_go, timesteps, attention_mask, want_all=True,
                                        states_bf16=sb)
        ctx.model, ctx.saved, ctx.wide = model, model._last, wide
        ctx.states_need_grad = bool(ctx.needs_input_grad[1])
        ctx.state_shape = states.shape
        ctx.set_materialize_grads(False)
        return sp, ap, rp

    @staticmethod
    def backward(ctx, d_state, d_action, d_return):
        model = ctx.model
        if ctx.states_need_grad and not ctx.wide and model.state_dim > 64:
            raise _lib.DtbError("a wide state (state_dim > 64) must be handed over as bf16 to get its gradient (the tensor-core path)")
        grads = torch.empty_like(model._flat)
        if d_action is None:  # NULL would mean "use the fused-loss gradient kept in the workspace"
            d_action = torch.zeros(ctx.state_shape[:2] + (model.act_dim,), dtype=torch.float32, device=grads.device)
        # wide states: bf16 gradient from the tensor-core dgrad; vector states (state_dim <= 64): fp32 gradient from embed_dstates
        d_states = torch.empty(ctx.state_shape, dtype=torch.bfloat16 if ctx.wide else torch.float32, device=grads.device) if ctx.states_need_grad else None
        model._backward_raw(ctx.saved, d_action, d_state, d_return, grads, d_states)
        out = [grads[off:off + torch.Size(shp).numel()].view(shp) for shp, off in zip(model._shapes, model._offsets)]
        return (None, d_states, None, None, None, None, None, *out)
