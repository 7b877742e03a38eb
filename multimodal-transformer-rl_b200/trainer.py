"""Host-side mirror of the upstream ``SequenceTrainer`` driving the sm_100a path.

``SequenceTrainer(model, optimizer, batch_size, get_batch, loss_fn=None, scheduler=None).train_step() -> float``.
With ``loss_fn=None`` (the upstream loss: mean squared action error over rows with ``attention_mask > 0``) the step is the
fused one: forward + loss + backward run inside ``libdtb200.so`` and write gradients straight into the model's flat
gradient buffer; a custom ``loss_fn(s_hat, a_hat, r_hat, s, a, r)`` goes through the generic autograd bridge instead.
Gradient clipping (0.25) and the optimizer are the caller's torch objects, as upstream — or ``mtrl_b200.optim.FusedClipAdamW``, which
does both on the flat buffers in three launches (then ``clip_grad_norm_`` is not called separately; its threshold is ``grad_clip``).
"""
from __future__ import annotations

import torch


class SequenceTrainer:
    def __init__(self, model, optimizer, batch_size, get_batch, loss_fn=None, scheduler=None, eval_fns=None, grad_clip=0.25):
        self.model, self.optimizer, self.batch_size, self.get_batch = model, optimizer, batch_size, get_batch
        self.loss_fn, self.scheduler, self.grad_clip = loss_fn, scheduler, grad_clip
        self.eval_fns = [] if eval_fns is None else eval_fns
        self.diagnostics = {}
        from .optim import FusedClipAdamW

        self._fused_opt = isinstance(optimizer, FusedClipAdamW)
        if self._fused_opt:
            optimizer.max_norm = grad_clip

    def loss_and_backward(self, batch):
        """forward + masked action MSE + backward (what the tokens/s metric times; optimizer excluded).  Returns a device scalar."""
        states, actions, rewards, dones, rtg, timesteps, attention_mask = batch
        rtg_in = rtg[:, :-1] if rtg.shape[1] == states.shape[1] + 1 else rtg
        if self.loss_fn is None:
            out = self.model.loss_and_backward(states, actions, rtg_in, timesteps, attention_mask, action_target=actions)
            return out[0]
        action_target = torch.clone(actions)
        _, action_preds, _ = self.model.forward(states, actions, rewards, rtg_in, timesteps, attention_mask=attention_mask)
        act_dim = action_preds.shape[2]
        keep = attention_mask.reshape(-1) > 0
        action_preds = action_preds.reshape(-1, act_dim)[keep]
        action_target = action_target.reshape(-1, act_dim)[keep]
        loss = self.loss_fn(None, action_preds, None, None, action_target, None)
        self.optimizer.zero_grad()
        loss.backward()
        return loss.detach()

    def train_step(self):
        batch = self.get_batch(self.batch_size)
        loss = self.loss_and_backward(batch)
        if not self._fused_opt:
            torch.nn.utils.clip_grad_norm_(self.model.parameters(), self.grad_clip)
        self.optimizer.step()
        if self.scheduler is not None:
            self.scheduler.step()
        value = loss.detach().cpu().item()
        self.diagnostics["training/action_error"] = value
        return value
