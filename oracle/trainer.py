"""ORACLE (test infrastructure): upstream SequenceTrainer.train_step restated.

No citable file exists in this container (the decision-transformer submodule is un-vendored); the rules below are the
documented assumptions: loss = mean((a_hat - a)^2) over rows with attention_mask > 0, grad-norm clip 0.25, AdamW,
`rtg[:, :-1]` is passed to the model when rtg carries K+1 entries.
"""
from __future__ import annotations

import torch


def masked_action_mse(action_preds, action_target, attention_mask):
    act_dim = action_preds.shape[-1]
    keep = attention_mask.reshape(-1) > 0
    ap = action_preds.reshape(-1, act_dim)[keep]
    at = action_target.reshape(-1, act_dim)[keep]
    return torch.mean((ap - at) ** 2)


class SequenceTrainer:
    def __init__(self, model, optimizer, batch_size, get_batch, loss_fn=None, scheduler=None, grad_clip=0.25):
        self.model, self.optimizer, self.batch_size, self.get_batch = model, optimizer, batch_size, get_batch
        self.loss_fn, self.scheduler, self.grad_clip = loss_fn, scheduler, grad_clip
        self.diagnostics = {}

    def loss_and_backward(self, batch):
        """forward + masked MSE + backward (the fwd+bwd the throughput metric times; optimizer excluded)."""
        states, actions, rewards, dones, rtg, timesteps, attention_mask = batch
        action_target = actions.clone()
        rtg_in = rtg[:, :-1] if rtg.shape[1] == states.shape[1] + 1 else rtg
        _, action_preds, _ = self.model.forward(states, actions, rewards, rtg_in, timesteps, attention_mask=attention_mask)
        loss = masked_action_mse(action_preds, action_target, attention_mask)
        self.optimizer.zero_grad()
        loss.backward()
        return loss

    def train_step(self):
        batch = self.get_batch(self.batch_size)
        loss = self.loss_and_backward(batch)
        torch.nn.utils.clip_grad_norm_(self.model.parameters(), self.grad_clip)
        self.optimizer.step()
        if self.scheduler is not None:
            self.scheduler.step()
        self.diagnostics["training/action_error"] = loss.detach().cpu().item()
        return loss.detach().cpu().item()
