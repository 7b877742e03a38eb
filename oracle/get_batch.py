"""ORACLE (test infrastructure, not product code): numpy restatement of upstream's `get_batch` closure.

PARITY STATUS: "parity unpinned by the reference" — the function lives in kzl/decision-transformer `gym/experiment.py`, the empty
un-vendored submodule of /root/reference (.SUBMODULES.json:8-16); nothing under /root/reference calls or tests it.  This file is
the documented restatement of SURVEY.md §3.1 (sampling proportional to trajectory length over the `pct_traj` best trajectories,
window slice, LEFT padding with states 0 / actions -10 / rewards 0 / dones 2 / timesteps 0 / mask 0, state normalisation AFTER the
padding, `discount_cumsum(rewards[si:], gamma=1)` with K + 1 entries divided by `scale`, timestep clamp to max_ep_len - 1) and is the
definition the CUDA path (csrc/get_batch.cu) is compared against, bit for bit, in tests/test_get_batch_gpu.py.
"""
from __future__ import annotations

import random

import numpy as np
import torch


def discount_cumsum(x, gamma):
    out = np.zeros_like(x)
    out[-1] = x[-1]
    for t in reversed(range(x.shape[0] - 1)):
        out[t] = x[t] + gamma * out[t + 1]
    return out


def select_trajectories(trajectories, pct_traj=1.0):
    """Upstream's selection: keep the highest-return trajectories covering pct_traj of all timesteps; sample proportionally to length."""
    traj_lens = np.array([len(t["rewards"]) for t in trajectories])
    returns = np.array([t["rewards"].sum() for t in trajectories])
    num_timesteps = max(int(pct_traj * traj_lens.sum()), 1)
    sorted_inds = np.argsort(returns)  # lowest to highest
    num_trajectories = 1
    timesteps = traj_lens[sorted_inds[-1]]
    ind = len(trajectories) - 2
    while ind >= 0 and timesteps + traj_lens[sorted_inds[ind]] <= num_timesteps:
        timesteps += traj_lens[sorted_inds[ind]]
        num_trajectories += 1
        ind -= 1
    sorted_inds = sorted_inds[-num_trajectories:]
    p_sample = traj_lens[sorted_inds] / sum(traj_lens[sorted_inds])
    return sorted_inds, p_sample


def sample_indices(trajectories, sorted_inds, p_sample, batch_size):
    """The host RNG calls of upstream, in upstream's order: np.random.choice for the trajectories, random.randint for each start."""
    batch_inds = np.random.choice(np.arange(len(sorted_inds)), size=batch_size, replace=True, p=p_sample)
    traj_idx = np.array([int(sorted_inds[i]) for i in batch_inds], dtype=np.int64)
    start = np.array([random.randint(0, trajectories[int(t)]["rewards"].shape[0] - 1) for t in traj_idx], dtype=np.int64)
    return traj_idx, start


def make_get_batch(trajectories, state_dim, act_dim, max_len, max_ep_len, scale, state_mean, state_std, pct_traj=1.0, device="cpu"):
    sorted_inds, p_sample = select_trajectories(trajectories, pct_traj)

    def get_batch(batch_size=256):
        traj_idx, start = sample_indices(trajectories, sorted_inds, p_sample, batch_size)
        s, a, r, d, rtg, timesteps, mask = [], [], [], [], [], [], []
        for i in range(batch_size):
            traj = trajectories[int(traj_idx[i])]
            si = int(start[i])
            s.append(traj["observations"][si:si + max_len].reshape(1, -1, state_dim))
            a.append(traj["actions"][si:si + max_len].reshape(1, -1, act_dim))
            r.append(traj["rewards"][si:si + max_len].reshape(1, -1, 1))
            d.append((traj["terminals"] if "terminals" in traj else traj["dones"])[si:si + max_len].reshape(1, -1))
            timesteps.append(np.arange(si, si + s[-1].shape[1]).reshape(1, -1))
            timesteps[-1][timesteps[-1] >= max_ep_len] = max_ep_len - 1  # padding cutoff
            rtg.append(discount_cumsum(traj["rewards"][si:], gamma=1.)[:s[-1].shape[1] + 1].reshape(1, -1, 1))
            if rtg[-1].shape[1] <= s[-1].shape[1]:
                rtg[-1] = np.concatenate([rtg[-1], np.zeros((1, 1, 1))], axis=1)
            tlen = s[-1].shape[1]
            s[-1] = np.concatenate([np.zeros((1, max_len - tlen, state_dim)), s[-1]], axis=1)
            s[-1] = (s[-1] - state_mean) / state_std
            a[-1] = np.concatenate([np.ones((1, max_len - tlen, act_dim)) * -10., a[-1]], axis=1)
            r[-1] = np.concatenate([np.zeros((1, max_len - tlen, 1)), r[-1]], axis=1)
            d[-1] = np.concatenate([np.ones((1, max_len - tlen)) * 2, d[-1]], axis=1)
            rtg[-1] = np.concatenate([np.zeros((1, max_len - tlen, 1)), rtg[-1]], axis=1) / scale
            timesteps[-1] = np.concatenate([np.zeros((1, max_len - tlen)), timesteps[-1]], axis=1)
            mask.append(np.concatenate([np.zeros((1, max_len - tlen)), np.ones((1, tlen))], axis=1))
        s = torch.from_numpy(np.concatenate(s, axis=0)).to(dtype=torch.float32, device=device)
        a = torch.from_numpy(np.concatenate(a, axis=0)).to(dtype=torch.float32, device=device)
        r = torch.from_numpy(np.concatenate(r, axis=0)).to(dtype=torch.float32, device=device)
        d = torch.from_numpy(np.concatenate(d, axis=0)).to(dtype=torch.long, device=device)
        rtg = torch.from_numpy(np.concatenate(rtg, axis=0)).to(dtype=torch.float32, device=device)
        timesteps = torch.from_numpy(np.concatenate(timesteps, axis=0)).to(dtype=torch.long, device=device)
        mask = torch.from_numpy(np.concatenate(mask, axis=0)).to(device=device)
        return s, a, r, d, rtg, timesteps, mask

    return get_batch
