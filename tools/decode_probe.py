"""Developer probe: GPU time of one KV-cached get_action step (dtb_dt_decode at t = K, cache holding K - 1 timesteps) vs the wall clock of
the Python call, per workload."""
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import bench  # noqa: E402

dev = torch.device("cuda:0")
for name in sys.argv[1:] or ["c1", "c3", "c4"]:
    wl = dict(bench.WORKLOADS[name])
    if wl["enc"] is not None:
        wl["enc"] = None  # trunk only: wide states are fed as bf16 features below
    model, parts = bench.build_ours(wl, dev)
    dt = parts[0]
    dt.eval()
    K = wl["K"]
    b = bench.make_batch(wl, 1, seed=1)
    sd = wl["dt"]["state_dim"]
    s = torch.randn(K, sd, device=dev)
    a, r, ts = b["actions"][0].to(dev), b["rtg"][0].to(dev).reshape(-1), b["timesteps"][0].to(dev)
    if sd > 64:
        call = lambda: dt._decode(None, s.to(torch.bfloat16), a, r, ts, K)  # noqa: E731
        sb = s.to(torch.bfloat16)
        call = lambda: dt._decode(None, sb, a, r, ts, K)  # noqa: E731
    else:
        call = lambda: dt._decode(s, None, a, r, ts, K)  # noqa: E731
    call()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n = 200
    gpu, wall = [], []
    for _ in range(n):
        dt._dc["t"] = K - 1
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        e0.record()
        call()
        e1.record()
        torch.cuda.synchronize()
        wall.append((time.perf_counter() - t0) * 1e6)
        gpu.append(e0.elapsed_time(e1) * 1e3)
    gpu.sort(); wall.sort()
    print(f"{name}: decode step GPU {gpu[n // 2]:.1f} us (event span incl. launch + clone), wall {wall[n // 2]:.1f} us", flush=True)
