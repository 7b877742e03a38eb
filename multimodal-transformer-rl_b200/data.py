"""GPU-side ``get_batch``: upstream's ``get_batch(batch_size)`` closure with the trajectories resident in HBM.

``GpuTrajectoryStore(trajectories, ...)`` flattens the upstream trajectory dicts (``observations`` / ``actions`` / ``rewards`` /
``terminals`` or ``dones``; optional raw per-timestep arrays such as uint8 frames) into per-timestep rows on the device once, together
with the reward suffix sums (upstream's ``discount_cumsum(rewards, gamma=1)``, evaluated by the same loop in the rewards' dtype).
``store.make_get_batch(K, max_ep_len, scale)`` returns a closure with upstream's signature and return tuple
``(states, actions, rewards, dones, rtg, timesteps, mask)``: the (trajectory, start) pairs are drawn on the host with upstream's own
RNG calls (``np.random.choice(p=p_sample)``, ``random.randint``), 2·B int64 cross PCIe, and one ``dtb_get_batch`` launch builds the
batch — bit-identical to the numpy closure (tests/test_get_batch_gpu.py) except that ``mask`` is fp32 instead of float64.
"""
from __future__ import annotations

import ctypes as C
import random

import numpy as np
import torch

from . import _lib


def _suffix_sums(x):
    out = np.zeros_like(x)
    out[-1] = x[-1]
    for t in reversed(range(x.shape[0] - 1)):
        out[t] = x[t] + 1.0 * out[t + 1]
    return out


class GpuTrajectoryStore:
    def __init__(self, trajectories, state_mean=None, state_std=None, pct_traj=1.0, raw_keys=(), device="cuda"):
        self.device = torch.device(device)
        lens = np.array([len(t["rewards"]) for t in trajectories])
        self.traj_lens = lens
        self.offsets_host = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
        dev = self.device
        f32 = lambda k: torch.from_numpy(np.concatenate([np.asarray(t[k], dtype=np.float32).reshape(len(t["rewards"]), -1) for t in trajectories])).to(dev)
        self.has_obs = "observations" in trajectories[0]
        if self.has_obs:
            self.obs = f32("observations").contiguous()
            self.state_dim = self.obs.shape[1]
            all_obs = np.concatenate([t["observations"] for t in trajectories], axis=0)
            self.state_mean_host = np.mean(all_obs, axis=0) if state_mean is None else np.asarray(state_mean)
            self.state_std_host = (np.std(all_obs, axis=0) + 1e-6) if state_std is None else np.asarray(state_std)
            # kept in double on the device: numpy evaluates (s - mean) / std in float64 whatever dtype mean / std arrive in
            self.state_mean = torch.from_numpy(self.state_mean_host.astype(np.float64)).to(dev)
            self.state_std = torch.from_numpy(self.state_std_host.astype(np.float64)).to(dev)
        self.act = f32("actions").contiguous()
        self.act_dim = self.act.shape[1]
        self.rew = f32("rewards").reshape(-1).contiguous()
        dk = "terminals" if "terminals" in trajectories[0] else "dones"
        self.done = f32(dk).reshape(-1).contiguous()
        self.rtg_suffix = torch.from_numpy(np.concatenate([_suffix_sums(np.asarray(t["rewards"])).astype(np.float64) for t in trajectories])).to(dev)
        self.offsets = torch.from_numpy(self.offsets_host).to(dev)
        self.raw = {}
        for k in raw_keys:
            arr = np.concatenate([np.ascontiguousarray(t[k]).reshape(len(t["rewards"]), -1) for t in trajectories])
            if (arr.shape[1] * arr.itemsize) % 16:
                raise ValueError(f"raw field {k!r}: per-timestep rows must be a multiple of 16 bytes")
            self.raw[k] = (torch.from_numpy(arr).to(dev).contiguous(), trajectories[0][k].shape[1:])
        # upstream's trajectory selection / sampling weights
        returns = np.array([np.asarray(t["rewards"]).sum() for t in trajectories])
        num_timesteps = max(int(pct_traj * lens.sum()), 1)
        sorted_inds = np.argsort(returns)
        n, ts, ind = 1, lens[sorted_inds[-1]], len(trajectories) - 2
        while ind >= 0 and ts + lens[sorted_inds[ind]] <= num_timesteps:
            ts += lens[sorted_inds[ind]]
            n += 1
            ind -= 1
        self.sorted_inds = sorted_inds[-n:]
        self.p_sample = lens[self.sorted_inds] / sum(lens[self.sorted_inds])

    def sample_indices(self, batch_size):
        batch_inds = np.random.choice(np.arange(len(self.sorted_inds)), size=batch_size, replace=True, p=self.p_sample)
        traj_idx = np.array([int(self.sorted_inds[i]) for i in batch_inds], dtype=np.int64)
        start = np.array([random.randint(0, int(self.traj_lens[t]) - 1) for t in traj_idx], dtype=np.int64)
        return traj_idx, start

    def gather(self, traj_idx, start, K, max_ep_len, scale):
        lib, dev = _lib.load(), self.device
        B = int(len(traj_idx))
        idx = torch.from_numpy(np.stack([np.asarray(traj_idx, dtype=np.int64), np.asarray(start, dtype=np.int64)])).to(dev, non_blocking=True)
        st, out = _lib.TrajStore(), _lib.BatchOut()
        keep = []
        if self.has_obs:
            st.obs, st.state_dim, st.state_mean, st.state_std = self.obs.data_ptr(), self.state_dim, self.state_mean.data_ptr(), self.state_std.data_ptr()
            s = torch.empty((B, K, self.state_dim), dtype=torch.float32, device=dev)
            out.states = s.data_ptr()
        else:
            s = None
        st.act, st.act_dim, st.rew, st.done, st.rtg_suffix = self.act.data_ptr(), self.act_dim, self.rew.data_ptr(), self.done.data_ptr(), self.rtg_suffix.data_ptr()
        st.traj_offset = self.offsets.data_ptr()
        a = torch.empty((B, K, self.act_dim), dtype=torch.float32, device=dev)
        r = torch.empty((B, K, 1), dtype=torch.float32, device=dev)
        d = torch.empty((B, K), dtype=torch.long, device=dev)
        rtg = torch.empty((B, K + 1, 1), dtype=torch.float32, device=dev)
        ts = torch.empty((B, K), dtype=torch.long, device=dev)
        m = torch.empty((B, K), dtype=torch.float32, device=dev)
        out.actions, out.rewards, out.dones, out.rtg, out.timesteps, out.mask = (a.data_ptr(), r.data_ptr(), d.data_ptr(), rtg.data_ptr(),
                                                                                 ts.data_ptr(), m.data_ptr())
        raws = {}
        st.n_raw = len(self.raw)
        for i, (k, (flat, shape)) in enumerate(self.raw.items()):
            dst = torch.empty((B, K, *shape), dtype=flat.dtype, device=dev)
            st.raw_src[i], st.raw_row_bytes[i], out.raw_dst[i] = flat.data_ptr(), flat.shape[1] * flat.element_size(), dst.data_ptr()
            raws[k] = dst
        keep.append(idx)
        _lib.check(lib.dtb_get_batch(_lib.stream_ptr(), C.byref(st), idx[0].data_ptr(), idx[1].data_ptr(), B, K, int(max_ep_len), float(scale),
                                     C.byref(out)))
        return (s, a, r, d, rtg, ts, m), raws

    def make_get_batch(self, K, max_ep_len, scale, with_raw=False):
        def get_batch(batch_size=256):
            traj_idx, start = self.sample_indices(batch_size)
            tup, raws = self.gather(traj_idx, start, K, max_ep_len, scale)
            return (tup, raws) if with_raw else tup
        return get_batch
