"""GPU-side get_batch (dtb_get_batch) against the numpy restatement of upstream's closure (oracle/get_batch.py): same host RNG calls,
then every tensor of the returned tuple must be bit-identical."""
import random

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _trajectories(n=40, sd=11, ad=3, seed=0, frames=False):
    rng = np.random.default_rng(seed)
    out = []
    for i in range(n):
        L = int(rng.integers(1, 90)) if i else 1  # ragged lengths, including a 1-step trajectory
        t = dict(observations=rng.normal(size=(L, sd)).astype(np.float32), actions=rng.uniform(-1, 1, size=(L, ad)).astype(np.float32),
                 rewards=rng.uniform(0, 5, size=L).astype(np.float32), terminals=np.zeros(L, dtype=bool))
        t["terminals"][-1] = True
        if frames:
            t["frames"] = rng.integers(0, 256, size=(L, 3, 16, 16), dtype=np.uint8)
        out.append(t)
    return out


@pytest.mark.parametrize("K,max_ep_len,pct", [(20, 1000, 1.0), (64, 50, 0.5), (1, 1000, 1.0)])
def test_get_batch_bit_exact(cuda_device, K, max_ep_len, pct):
    from mtrl_b200.data import GpuTrajectoryStore
    from oracle.get_batch import make_get_batch

    trajs = _trajectories()
    all_obs = np.concatenate([t["observations"] for t in trajs], axis=0)
    mean, std = np.mean(all_obs, axis=0), np.std(all_obs, axis=0) + 1e-6
    ref = make_get_batch(trajs, 11, 3, K, max_ep_len, 1000.0, mean, std, pct_traj=pct)
    store = GpuTrajectoryStore(trajs, pct_traj=pct, device=cuda_device)
    ours = store.make_get_batch(K, max_ep_len, 1000.0)
    for it in range(3):
        np.random.seed(7 + it); random.seed(7 + it)
        want = ref(37)
        np.random.seed(7 + it); random.seed(7 + it)
        got = ours(37)
        names = ["states", "actions", "rewards", "dones", "rtg", "timesteps", "mask"]
        for n, w, g in zip(names, want, got):
            g = g.cpu()
            assert g.shape == w.shape, (n, g.shape, w.shape)
            if n == "mask":
                assert g.dtype == torch.float32 and torch.equal(g.double(), w.double())
            else:
                assert g.dtype == w.dtype, (n, g.dtype, w.dtype)
                assert torch.equal(g, w), (n, (g != w).sum().item())
    # windows that run past the end of a trajectory really occur (left padding, rtg tail 0) and so do clamped timesteps
    if K > 1:
        assert (want[6] == 0).any()
    if max_ep_len == 50:
        assert (want[5] == max_ep_len - 1).any()


def test_get_batch_raw_rows_and_train_step(cuda_device):
    """Raw per-timestep rows (uint8 frames) follow the same window / left-padding rule, and the closure feeds SequenceTrainer directly."""
    from mtrl_b200.data import GpuTrajectoryStore
    from mtrl_b200.model import DecisionTransformer
    from mtrl_b200.optim import FusedClipAdamW
    from mtrl_b200.trainer import SequenceTrainer

    trajs = _trajectories(n=12, frames=True, seed=3)
    K = 8
    store = GpuTrajectoryStore(trajs, raw_keys=("frames",), device=cuda_device)
    np.random.seed(1); random.seed(1)
    traj_idx, start = store.sample_indices(9)
    (s, a, r, d, rtg, ts, m), raws = store.gather(traj_idx, start, K, 1000, 1000.0)
    fr = raws["frames"].cpu()
    assert fr.shape == (9, K, 3, 16, 16) and fr.dtype == torch.uint8
    for b in range(9):
        t, si = trajs[int(traj_idx[b])], int(start[b])
        win = torch.from_numpy(t["frames"][si:si + K])
        pad = K - win.shape[0]
        assert torch.equal(fr[b, pad:], win) and (fr[b, :pad] == 0).all()
    model = DecisionTransformer(11, 3, 128, max_length=K, max_ep_len=1000, n_layer=1, n_head=1, resid_pdrop=0.0, attn_pdrop=0.0, device=cuda_device)
    tr = SequenceTrainer(model, FusedClipAdamW(model, lr=1e-3, weight_decay=1e-4), 16, store.make_get_batch(K, 1000, 1000.0))
    losses = [tr.train_step() for _ in range(3)]
    assert all(np.isfinite(losses))
