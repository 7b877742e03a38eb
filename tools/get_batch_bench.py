"""Developer probe: upstream-style numpy get_batch (+ host->device copy) vs the GPU-side one, hopper-shaped synthetic trajectories."""
import os
import random
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from mtrl_b200.data import GpuTrajectoryStore  # noqa: E402
from oracle.get_batch import make_get_batch  # noqa: E402

B, K, sd, ad = int(sys.argv[1]) if len(sys.argv) > 1 else 64, int(sys.argv[2]) if len(sys.argv) > 2 else 20, 11, 3
rng = np.random.default_rng(0)
trajs = []
for _ in range(2000):
    L = int(rng.integers(200, 1000))
    trajs.append(dict(observations=rng.normal(size=(L, sd)).astype(np.float32), actions=rng.uniform(-1, 1, size=(L, ad)).astype(np.float32),
                      rewards=rng.uniform(0, 5, size=L).astype(np.float32), terminals=np.zeros(L, dtype=bool)))
obs = np.concatenate([t["observations"] for t in trajs])
ref = make_get_batch(trajs, sd, ad, K, 1000, 1000.0, obs.mean(0), obs.std(0) + 1e-6, device="cuda")
store = GpuTrajectoryStore(trajs, device="cuda")
ours = store.make_get_batch(K, 1000, 1000.0)
for name, fn in (("numpy closure + H2D", ref), ("dtb_get_batch", ours)):
    np.random.seed(0); random.seed(0)
    for _ in range(5):
        fn(B)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    n = 50
    for _ in range(n):
        fn(B)
    torch.cuda.synchronize()
    print(f"{name:22s} B={B} K={K}: {(time.perf_counter() - t0) / n * 1e6:9.1f} us per batch")
