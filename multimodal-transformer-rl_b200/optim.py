"""Fused tail of the reference's ``SequenceTrainer.train_step()``:

    torch.nn.utils.clip_grad_norm_(model.parameters(), .25)
    optimizer.step()                                   # torch.optim.AdamW(model.parameters(), lr, weight_decay)

over the flat fp32 parameter / gradient buffers of the sm_100a modules (``DecisionTransformer``, ``MultimodalStateEncoder``): one
global gradient norm across all modules, clip coefficient, bias corrections and the step count computed on the device, then one
pass that clips, decays, applies Adam and rewrites the bf16 shadow the GEMMs read (``dtb_clip_adamw_step``, three launches, no host
synchronisation — the step can be captured into the CUDA graph of the training step).  Arithmetic follows torch's single-tensor
AdamW; parity is checked against ``clip_grad_norm_`` + ``torch.optim.AdamW`` in tests/test_optim_gpu.py.
"""
from __future__ import annotations

import ctypes as C

import torch

from . import _lib


class FusedClipAdamW:
    def __init__(self, modules, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=1e-2, max_norm=0.25, write_clipped_grads=True, warmup_steps=0):
        self.modules = list(modules) if isinstance(modules, (list, tuple)) else [modules]
        if not 1 <= len(self.modules) <= 8:
            raise ValueError("FusedClipAdamW takes 1..8 modules")
        for m in self.modules:
            if not hasattr(m, "flat_params") or not hasattr(m, "flat_grads"):
                raise TypeError("FusedClipAdamW needs modules with flat_params / flat_grads (DecisionTransformer, MultimodalStateEncoder)")
        self.lr, self.betas, self.eps, self.weight_decay, self.max_norm = float(lr), (float(betas[0]), float(betas[1])), float(eps), float(weight_decay), max_norm
        self.write_clipped_grads = bool(write_clipped_grads)
        self.warmup_steps = int(warmup_steps)  # > 0: upstream's LambdaLR warm-up, evaluated on the device from the step count
        self.lr_tensor = None  # optional device scalar read at kernel run time (learning-rate schedules inside a CUDA graph)
        self.param_groups = [{"lr": self.lr}]  # scheduler-style access: step() reads param_groups[0]["lr"]
        self._state = None

    # ---- state ---------------------------------------------------------------------------------------------------
    def _ensure_state(self):
        dev = self.modules[0].flat_params.device
        if self._state is not None and self._state["dev"] == dev and all(
                m.data_ptr() == mod.flat_params.data_ptr() for m, mod in zip(self._state["ptrs"], self.modules)):
            return self._state
        lib = _lib.load()
        old = self._state
        st = dict(dev=dev, ptrs=[mod.flat_params for mod in self.modules],
                  exp_avg=[torch.zeros_like(mod.flat_params) for mod in self.modules],
                  exp_avg_sq=[torch.zeros_like(mod.flat_params) for mod in self.modules],
                  step=torch.zeros(1, dtype=torch.int32, device=dev),
                  scratch=torch.zeros(int(lib.dtb_adamw_scratch_floats()), dtype=torch.float32, device=dev))
        if old is not None:  # the modules were moved (.to()): carry the moments over
            for k in ("exp_avg", "exp_avg_sq"):
                for new, prev in zip(st[k], old[k]):
                    new.copy_(prev)
            st["step"].copy_(old["step"])
        self._state = st
        return st

    @property
    def step_count(self) -> int:
        return 0 if self._state is None else int(self._state["step"].item())

    def grad_norm(self) -> torch.Tensor:
        """Global gradient norm (before clipping) of the last step — device scalar, like clip_grad_norm_'s return value."""
        return self._ensure_state()["scratch"][3]

    def zero_grad(self, set_to_none: bool = False):
        """Point every ``param.grad`` at its view of the module's flat gradient buffer and clear the buffer.

        The fused ``loss_and_backward`` overwrites the flat buffers and needs none of this; the autograd bridge (``model(...)`` +
        ``loss.backward()``, ``SequenceTrainer(loss_fn=...)``) ACCUMULATES into ``param.grad``: with the views attached and zeroed here
        the gradients land exactly where ``step()`` reads them.  ``set_to_none`` is ignored on purpose (a ``None`` grad would make
        autograd install tensors outside the flat buffer)."""
        for mod in self.modules:
            mod.attach_grads().zero_()

    def _check_grads_alias(self):
        """step() reads only ``mod.flat_grads``: a ``param.grad`` that lives elsewhere (autograd installed it because it was None) would
        be silently ignored — refuse instead."""
        for mod in self.modules:
            base = mod.flat_grads.data_ptr()
            for p, off in zip(mod._plist, mod._offsets):
                if p.grad is not None and p.grad.data_ptr() != base + 4 * off:
                    raise RuntimeError("FusedClipAdamW.step(): a param.grad does not alias the module's flat gradient buffer; call "
                                       "optimizer.zero_grad() (or module.attach_grads()) before loss.backward()")

    def state_dict(self):
        st = self._ensure_state()
        return {"step": st["step"].clone(), "exp_avg": [t.clone() for t in st["exp_avg"]], "exp_avg_sq": [t.clone() for t in st["exp_avg_sq"]],
                "hparams": dict(lr=self.param_groups[0]["lr"], betas=self.betas, eps=self.eps, weight_decay=self.weight_decay, max_norm=self.max_norm)}

    def load_state_dict(self, sd):
        st = self._ensure_state()
        st["step"].copy_(sd["step"])
        for k in ("exp_avg", "exp_avg_sq"):
            for dst, src in zip(st[k], sd[k]):
                dst.copy_(src)

    # ---- the step ------------------------------------------------------------------------------------------------
    @torch.no_grad()
    def step(self):
        lib = _lib.load()
        st = self._ensure_state()
        self._check_grads_alias()
        n = len(self.modules)
        segs = (_lib.AdamwSegment * n)()
        for i, mod in enumerate(self.modules):
            p, g = mod.flat_params, mod.flat_grads
            shadow = getattr(mod, "_flat_bf16", None)  # plain bf16 cast of the flat buffer (DecisionTransformer); others restage
            segs[i].params, segs[i].grads = p.data_ptr(), g.data_ptr()
            segs[i].exp_avg, segs[i].exp_avg_sq = st["exp_avg"][i].data_ptr(), st["exp_avg_sq"][i].data_ptr()
            segs[i].params_bf16 = shadow.data_ptr() if shadow is not None and shadow.device == p.device else None
            segs[i].n = p.numel()
        hp = _lib.AdamwHparams(float(self.param_groups[0]["lr"]), self.betas[0], self.betas[1], self.eps, self.weight_decay,
                               float(self.max_norm) if self.max_norm else 0.0,
                               self.lr_tensor.data_ptr() if self.lr_tensor is not None else None, int(self.write_clipped_grads),
                               self.warmup_steps)
        _lib.check(lib.dtb_clip_adamw_step(_lib.stream_ptr(), n, segs, C.byref(hp), _lib.ptr(st["step"]), _lib.ptr(st["scratch"])))
        for i, mod in enumerate(self.modules):
            # the library wrote the fp32 master behind torch's back: shadows it rewrote are current, the others must be re-staged
            if segs[i].params_bf16:
                mod._staged_version = mod._version_now()
            else:  # GEMM-ready bf16 copies in another order than the flat buffer (conv towers): re-staged here, as part of the optimizer
                mod._staged_version = -1  # step, so that the forward / backward step itself launches no weight-staging kernels
                mod._stage()
