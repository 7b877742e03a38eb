"""Host-side mirror of the reference's multimodal state encoder (``MinecraftAgent`` towers + fusion, REF src/agents.py:93-143)
and of the multimodal Decision Transformer built on it (encoder output = state tokens, trained end to end).

``MultimodalStateEncoder(frame_shape, audio_shape, conv_type='big', fusion_type='concat')(frames, audio) -> states[B,K,state_dim]``
mirrors ``MinecraftAgent(envs, device, conv_type, attn_type=None, fusion_type)``'s ``video_net`` / ``audio_net`` / fusion; the
parameter names (``video_net.0.weight`` ...) and layouts are the reference's ``conv_factory`` nn.Sequential ones, so a
reference state_dict loads directly.  Compute runs only in ``libdtb200.so`` (no fallback).
"""
from __future__ import annotations

import ctypes as C

import torch
import torch.nn as nn

from . import _lib
from .model import DecisionTransformer, _set_nested


def _conv_out(n, k, s):
    return (n - k) // s + 1


def _tower_entries(prefix, conv_type, cin, h, w):
    if conv_type == "big":
        specs, feat, fc_idx = [(32, 8, 4), (64, 4, 2), (64, 3, 1)], 512, 7
    elif conv_type == "small":
        specs, feat, fc_idx = [(16, 8, 4), (32, 4, 2)], 256, 5
    else:
        raise NotImplementedError(conv_type)
    names, shapes = [], []
    for i, (cout, k, s) in enumerate(specs):
        names += [f"{prefix}.{2 * i}.weight", f"{prefix}.{2 * i}.bias"]
        shapes += [(cout, cin, k, k), (cout,)]
        cin, h, w = cout, _conv_out(h, k, s), _conv_out(w, k, s)
    names += [f"{prefix}.{fc_idx}.weight", f"{prefix}.{fc_idx}.bias"]
    shapes += [(feat, cin * h * w), (feat,)]
    return names, shapes, feat


class MultimodalStateEncoder(nn.Module):
    def __init__(self, frame_shape=(1, 84, 84), audio_shape=(1, 84, 84), conv_type="big", fusion_type="concat", input_scale=1.0 / 255.0,
                 device=None, attn_type=None):
        super().__init__()
        self.frame_shape, self.audio_shape = tuple(int(x) for x in frame_shape), tuple(int(x) for x in audio_shape)
        self.conv_type, self.fusion_type, self.input_scale = conv_type, fusion_type, float(input_scale)
        if fusion_type not in ("concat", "sum"):
            raise NotImplementedError(fusion_type)
        if attn_type not in (None, "", "casl", "new"):
            raise NotImplementedError(attn_type)  # as MinecraftAgent.__init__ (agents.py:108-114)
        self.attn_type = attn_type or None
        lib = _lib.load()
        big = 1 if conv_type == "big" else 0
        # with modality attention the towers always write [video | audio] side by side; the attention op does the fusion
        cfg = _lib.EncoderConfig(_lib.TowerConfig(*self.frame_shape, big), _lib.TowerConfig(*self.audio_shape, big),
                                 0 if (fusion_type == "concat" or self.attn_type) else 1, self.input_scale)
        h = C.c_void_p()
        _lib.check(lib.dtb_encoder_create(C.byref(cfg), C.byref(h)))
        self._h = h
        self.state_dim = int(lib.dtb_encoder_state_dim(self._h))
        n = lib.dtb_encoder_param_entries(self._h)
        offs, nums = (C.c_longlong * n)(), (C.c_longlong * n)()
        self._total = int(lib.dtb_encoder_param_layout(self._h, offs, nums))
        vn, vs, self.feature_size = _tower_entries("video_net", conv_type, *self.frame_shape)
        an, as_, _ = _tower_entries("audio_net", conv_type, *self.audio_shape)
        self._names, self._shapes = vn + an, vs + as_
        assert len(self._names) == n
        self._offsets = [int(o) for o in offs]
        for shp, num in zip(self._shapes, nums):
            assert int(torch.Size(shp).numel()) == int(num), (shp, num)
        self._tower_total, self._acfg, self._attn_ws, self.attn_weights = self._total, None, {}, None
        if self.attn_type:  # attn.* parameters (reference names, agents.py:58-90) appended to the same flat buffer
            F = self.feature_size
            self._acfg = _lib.ModalAttnConfig(0 if self.attn_type == "casl" else 1, F, 0 if fusion_type == "concat" else 1)
            na = _lib.check(lib.dtb_modal_attn_param_entries(C.byref(self._acfg)))
            aoffs, anums = (C.c_longlong * na)(), (C.c_longlong * na)()
            atotal = int(lib.dtb_modal_attn_param_layout(C.byref(self._acfg), aoffs, anums))
            if self.attn_type == "casl":
                anames = ["attn.audio_fc", "attn.video_fc", "attn.state_fc", "attn.attention"]
                ashapes = [(32, F), (32,), (32, F), (32,), (32, 128), (32,), (2, 32), (2,)]
            else:
                anames = ["attn.fc", "attn.attention"]
                ashapes = [(128, 2 * F + 128), (128,), (2, 128), (2,)]
            self._names = self._names + [f"{m}.{k}" for m in anames for k in ("weight", "bias")]
            self._shapes = self._shapes + ashapes
            self._offsets = self._offsets + [self._tower_total + int(o) for o in aoffs]
            for shp, num in zip(ashapes, anums):
                assert int(torch.Size(shp).numel()) == int(num), (shp, num)
            self._total = self._tower_total + atotal
            self.state_dim = int(lib.dtb_modal_attn_state_dim(C.byref(self._acfg)))
        if device is None:
            device = "cuda" if torch.cuda.is_available() else "cpu"
        self._flat = torch.zeros(self._total, dtype=torch.float32, device=device)
        self._flat_grad, self._staged, self._staged_version, self._ws = None, None, -1, {}
        for name, shp, off in zip(self._names, self._shapes, self._offsets):
            _set_nested(self, name, nn.Parameter(self._flat[off:off + torch.Size(shp).numel()].view(shp)))
        self._plist = [self.get_parameter(n) for n in self._names]
        self.reset_parameters()

    def _version_now(self) -> int:
        """Staleness stamp of the fp32 master: flat buffer + every Parameter (see DecisionTransformer._version_now)."""
        return self._flat._version + sum(p._version for p in self._plist)

    # parameter plumbing shared in spirit with DecisionTransformer (flat buffer + views)
    def _apply(self, fn, recurse=True):
        super()._apply(fn, recurse)
        ref = self.get_parameter(self._names[0])
        flat = torch.zeros(self._total, dtype=torch.float32, device=ref.device)
        for name, shp, off in zip(self._names, self._shapes, self._offsets):
            flat[off:off + torch.Size(shp).numel()].view(shp).copy_(self.get_parameter(name).data.float())
        self._flat = flat
        self._flat_grad, self._staged, self._staged_version, self._ws = None, None, -1, {}
        self._attn_ws = {}
        for name, shp, off in zip(self._names, self._shapes, self._offsets):
            self.get_parameter(name).data = self._flat[off:off + torch.Size(shp).numel()].view(shp)
        return self

    @torch.no_grad()
    def reset_parameters(self):
        """layer_init of the reference (src/utils.py:146-149): orthogonal(gain sqrt 2) weights, zero biases."""
        for name, p in self.named_parameters():
            if name.startswith("attn."):  # plain nn.Linear default init in the reference (agents.py:61-64, 81-82)
                if name.endswith(".weight"):
                    w = torch.empty(p.shape, dtype=torch.float32)
                    torch.nn.init.kaiming_uniform_(w, a=5 ** 0.5)
                    p.copy_(w)
                else:
                    fan_in = self.get_parameter(name[:-4] + "weight").shape[1]
                    p.copy_(torch.empty(p.shape).uniform_(-1.0 / fan_in ** 0.5, 1.0 / fan_in ** 0.5))
            elif name.endswith(".bias"):
                p.zero_()
            else:
                w = torch.empty(p.shape, dtype=torch.float32)
                torch.nn.init.orthogonal_(w, 2 ** 0.5)
                p.copy_(w)

    def load_state_dict(self, state_dict, strict=True, assign=False):
        out = super().load_state_dict(state_dict, strict=strict, assign=False)
        self._staged_version = -1
        return out

    @property
    def flat_params(self):
        return self._flat

    @property
    def flat_grads(self):
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

    def _stage(self):
        if self._staged is None or self._staged.device != self._flat.device:
            n = max(self._total, int(_lib.load().dtb_encoder_staged_elems(self._h)))  # + data-gradient weights of the implicit convs
            self._staged = torch.zeros(n, dtype=torch.bfloat16, device=self._flat.device)
            self._staged_version = -1
        ver = self._version_now()
        if self._staged_version != ver:
            _lib.check(_lib.load().dtb_encoder_stage_params(self._h, _lib.stream_ptr(), _lib.ptr(self._flat), _lib.ptr(self._staged)))
            self._staged_version = ver

    def _workspace(self, N):
        ws = self._ws.get(N)
        if ws is None:
            if len(self._ws) > 4:
                self._ws.clear()
            ws = torch.empty(int(_lib.load().dtb_encoder_workspace_bytes(self._h, N)), dtype=torch.uint8, device=self._flat.device)
            self._ws[N] = ws
        return ws

    def _prep(self, frames, audio):
        if not frames.is_cuda or not self._flat.is_cuda:
            raise _lib.DtbError("MultimodalStateEncoder runs only on an sm_100a CUDA device through libdtb200.so (no CPU fallback)")
        B, K = frames.shape[0], frames.shape[1]
        if tuple(frames.shape[2:]) != self.frame_shape or tuple(audio.shape[2:]) != self.audio_shape or audio.shape[:2] != frames.shape[:2]:
            raise ValueError(f"expected frames [B,K,{self.frame_shape}] and audio [B,K,{self.audio_shape}], got {tuple(frames.shape)}, "
                             f"{tuple(audio.shape)}")
        # uint8 frames are raw pixels: the library applies the reference's decode (img * 2 / 255 - 1, Minecraft.py:333-335) in-kernel
        f = frames.detach().contiguous() if frames.dtype == torch.uint8 else frames.detach().to(torch.float32).contiguous()
        return B, K, f, audio.detach().to(torch.float32).contiguous()

    def _attn_workspace(self, N):
        ws = self._attn_ws.get(N)
        if ws is None:
            if len(self._attn_ws) > 4:
                self._attn_ws.clear()
            nbytes = int(_lib.load().dtb_modal_attn_workspace_bytes(C.byref(self._acfg), N))
            ws = (torch.empty(nbytes, dtype=torch.uint8, device=self._flat.device),
                  torch.empty((N, 2 * self.feature_size), dtype=torch.bfloat16, device=self._flat.device),   # tower features
                  torch.empty((N, 2 * self.feature_size), dtype=torch.bfloat16, device=self._flat.device),   # their gradient
                  torch.empty((N, 2), dtype=torch.float32, device=self._flat.device))                        # attention weights
            self._attn_ws[N] = ws
        return ws

    def _hidden(self, hidden, N):
        """hidden: None (the Decision Transformer has no recurrent state: zeros), [N,128], [1,N,128] (casl's lstm_state[0]) or one
        row [1,128] / [1,1,128] repeated over the samples (NewAttention, agents.py:85)."""
        if hidden is None:
            return None
        h = hidden.detach().to(device=self._flat.device, dtype=torch.float32).reshape(-1, 128)
        if h.shape[0] == 1 and N > 1:
            h = h.expand(N, 128)
        if h.shape[0] != N:
            raise ValueError(f"hidden must have 1 or {N} rows of 128, got {tuple(hidden.shape)}")
        return h.contiguous()

    def forward_raw(self, frames, audio, hidden=None):
        """frames [B,K,Cv,H,W] (fp32 in the reference's decoded range, or uint8 pixels), audio [B,K,Ca,H,W] (fp32) -> states bf16
        [B,K,state_dim]; keeps what backward_raw needs.  hidden: recurrent state for the modality attention (attn_type only)."""
        B, K, f, a = self._prep(frames, audio)
        N = B * K
        self._stage()
        ws = self._workspace(N)
        states = torch.empty((B, K, self.state_dim), dtype=torch.bfloat16, device=self._flat.device)
        lib = _lib.load()
        fwd = lib.dtb_encoder_forward_u8 if f.dtype == torch.uint8 else lib.dtb_encoder_forward
        if not self.attn_type:
            if hidden is not None:
                raise ValueError("hidden is only used by the modality attention (attn_type='casl' | 'new')")
            _lib.check(fwd(self._h, _lib.stream_ptr(), _lib.ptr(self._flat), _lib.ptr(self._staged), N, _lib.ptr(f),
                           _lib.ptr(a), _lib.ptr(ws), ws.numel(), _lib.ptr(states)))
            self._last = (N, f, a, ws, states)
            return states
        aws, feat, _, weights = self._attn_workspace(N)
        h = self._hidden(hidden, N)
        _lib.check(fwd(self._h, _lib.stream_ptr(), _lib.ptr(self._flat), _lib.ptr(self._staged), N, _lib.ptr(f),
                       _lib.ptr(a), _lib.ptr(ws), ws.numel(), _lib.ptr(feat)))
        _lib.check(lib.dtb_modal_attn_forward(C.byref(self._acfg), _lib.stream_ptr(), _lib.ptr(self._flat[self._tower_total:]), N,
                                              _lib.ptr(feat), _lib.ptr(h), _lib.ptr(aws), aws.numel(), _lib.ptr(states), _lib.ptr(weights)))
        self.attn_weights = weights.view(B, K, 2)
        self._last = (N, f, a, ws, states, h is not None)
        return states

    def backward_raw(self, d_states, masked: bool, grads=None):
        N, f, a, ws, states = self._last[:5]
        grads = self.attach_grads() if grads is None else grads
        lib = _lib.load()
        bwd = lib.dtb_encoder_backward_u8 if f.dtype == torch.uint8 else lib.dtb_encoder_backward
        if self.attn_type:
            # attention backward -> gradient of the tower features, already ReLU-masked; a d_states masked by (states > 0) is fine
            # too: states > 0 exactly where the features are (the weights are positive), and masked entries only meet zero features
            aws, feat, dfeat, _ = self._attn_workspace(N)
            _lib.check(lib.dtb_modal_attn_backward(C.byref(self._acfg), _lib.stream_ptr(), _lib.ptr(self._flat[self._tower_total:]), N,
                                                   _lib.ptr(feat), int(self._last[5]), _lib.ptr(d_states), _lib.ptr(aws), aws.numel(),
                                                   _lib.ptr(dfeat), _lib.ptr(grads[self._tower_total:])))
            states, d_states, masked = feat, dfeat, True
        _lib.check(bwd(self._h, _lib.stream_ptr(), _lib.ptr(self._flat), _lib.ptr(self._staged), N, _lib.ptr(f), _lib.ptr(a), _lib.ptr(states),
                       _lib.ptr(d_states), int(masked), _lib.ptr(ws), ws.numel(), _lib.ptr(grads)))
        return grads

    def fc_grad_range(self):
        """(begin, end) of the gradients that are final first in backward: the two FC layers (+ the modality-attention parameters, whose
        backward runs before the towers'); [0, begin) are the conv layers' gradients, final only when backward_raw's kernels are done."""
        b, e = C.c_longlong(), C.c_longlong()
        _lib.check(_lib.load().dtb_encoder_fc_grad_range(self._h, C.byref(b), C.byref(e)))
        return int(b.value), self._total

    def wait_fc_grads(self, stream):
        """Make ``stream`` (a torch.cuda.Stream) wait until the FC-range gradients of the last backward_raw call are final."""
        _lib.check(_lib.load().dtb_encoder_wait_fc_grads(self._h, C.c_void_p(stream.cuda_stream)))

    def forward(self, frames, audio, hidden=None):
        params = [self.get_parameter(n) for n in self._names]
        return _EncoderFunction.apply(self, frames, audio, hidden, *params)

    def __del__(self):
        try:
            if getattr(self, "_h", None) is not None and self._h.value:
                _lib.load().dtb_encoder_destroy(self._h)
                self._h = None
        except Exception:
            pass


class _EncoderFunction(torch.autograd.Function):
    @staticmethod
    def forward(ctx, enc, frames, audio, hidden, *params):
        states = enc.forward_raw(frames, audio, hidden)
        ctx.enc, ctx.saved = enc, enc._last
        return states

    @staticmethod
    def backward(ctx, d_states):
        enc = ctx.enc
        if enc._last is not ctx.saved:
            raise _lib.DtbError("MultimodalStateEncoder.backward: workspace overwritten by a later forward")
        grads = torch.empty_like(enc._flat)
        d = d_states.detach().to(torch.bfloat16).contiguous().clone()  # masked in place by the library
        enc.backward_raw(d, masked=False, grads=grads)
        out = [grads[off:off + torch.Size(shp).numel()].view(shp) for shp, off in zip(enc._shapes, enc._offsets)]
        return (None, None, None, None, *out)


class MultimodalDecisionTransformer(nn.Module):
    """Encoder towers -> state tokens -> DecisionTransformer, trained end to end (configs c2 / c3 / c5 of BASELINE.json)."""

    def __init__(self, encoder: MultimodalStateEncoder, dt: DecisionTransformer):
        super().__init__()
        if dt.state_dim != encoder.state_dim:
            raise ValueError(f"DecisionTransformer.state_dim ({dt.state_dim}) must equal the encoder's ({encoder.state_dim})")
        self.encoder, self.dt = encoder, dt

    def forward(self, frames, audio, actions, rewards, returns_to_go, timesteps, attention_mask=None):
        states = self.encoder(frames, audio)
        return self.dt(states, actions, rewards, returns_to_go, timesteps, attention_mask)

    @torch.no_grad()
    def get_action(self, frames, audio, actions, rewards, returns_to_go, timesteps, **kwargs):
        """frames [t,Cv,H,W] (fp32 or uint8 pixels), audio [t,Ca,H,W]: window + left padding of the raw history in one launch, then one
        CUDA graph of towers -> state tokens -> B = 1 Decision Transformer forward (model._ActionRunner)."""
        from .model import _ActionRunner

        dt, K = self.dt, self.dt.max_length
        t = frames.shape[0]
        if K is not None and t <= K and not self.encoder.attn_type and dt._use_cache(kwargs.get("use_cache")):
            # KV-cached step: the towers run on the NEW frames only, their features join a K-row history buffer, one decode kernel follows
            hist = self.__dict__.get("_feat_hist")
            if hist is None or hist.device != dt._flat.device:
                hist = self.__dict__["_feat_hist"] = torch.zeros((K, self.encoder.state_dim), dtype=torch.bfloat16, device=dt._flat.device)
            dc = dt._dc
            self.encoder._stage()
            dt._stage()
            stale = dc is None or dc["ptr"] != dt._flat.data_ptr() or dc["stamp"] != dt._staged_version or t <= dc["t"] \
                or self.__dict__.get("_feat_stamp") != self.encoder._staged_version
            if not stale and self.__dict__.get("_feat_stamp") == self.encoder._staged_version:
                fdt = torch.uint8 if frames.dtype == torch.uint8 else torch.float32
                r = dt._step_runner(t, self.encoder, fdt)
                if r is not None:  # the next step of a cached rollout: one pad_left launch + one graph replay (towers on the new frame + decode)
                    return r.step([(frames, r.frames), (audio, r.audio), (actions, r.actions), (returns_to_go, r.rtg), (timesteps, r.timesteps)], t)
            t0 = 0 if stale else dc["t"]
            if stale and dc is not None:
                dc["t"] = 0
            self.__dict__["_feat_stamp"] = self.encoder._staged_version
            hist[t0:t] = self.encoder.forward_raw(frames[t0:t].unsqueeze(0), audio[t0:t].unsqueeze(0))[0]
            return dt._decode(None, hist, dt._rows(actions, dt.act_dim, torch.float32), dt._rows(returns_to_go, 0, torch.float32),
                              dt._rows(timesteps, 0, torch.long), t)
        if K is None:
            states = self.encoder.forward_raw(frames.unsqueeze(0), audio.unsqueeze(0))[0].float()
            return dt.get_action(states, actions, rewards, returns_to_go, timesteps, **kwargs)
        fdt = torch.uint8 if frames.dtype == torch.uint8 else torch.float32
        runner = self.__dict__.get("_action_runner")
        if runner is None or runner.dead(dt) or runner.frames.dtype != fdt:
            runner = _ActionRunner(dt, self.encoder, fdt)
            self.__dict__["_action_runner"] = runner
        return runner.run([(frames, runner.frames), (audio, runner.audio), (actions, runner.actions), (returns_to_go, runner.rtg),
                           (timesteps, runner.timesteps)])

    def loss_and_backward(self, frames, audio, actions, returns_to_go, timesteps, attention_mask, action_target=None, grad_scale=1.0,
                          on_stage_done=None):
        """Fused end-to-end training step body: towers -> DT forward -> masked action MSE -> DT backward -> towers backward."""
        states = self.encoder.forward_raw(frames, audio)
        d_states = torch.empty_like(states)
        # the concatenated state is itself relu(features) (scaled by positive attention weights, if any); a sum is not
        relu_out = self.encoder.fusion_type == "concat"
        loss = self.dt.loss_and_backward(None, actions, returns_to_go, timesteps, attention_mask, action_target=action_target,
                                         grad_scale=grad_scale, states_bf16=states, d_states_bf16=d_states, states_relu=relu_out,
                                         on_stage_done=on_stage_done)
        self.encoder.backward_raw(d_states, masked=relu_out)
        if on_stage_done is not None:
            # two ranges of the encoder's flat gradients: the FC layers (final early: GradAllReducer.callback_for waits on the library's
            # event instead of the whole backward) and the conv layers (final when everything enqueued so far is done)
            fc_begin, fc_end = self.encoder.fc_grad_range()
            on_stage_done("encoder_fc", fc_begin, fc_end)
            on_stage_done("encoder", 0, fc_begin)
        return loss
