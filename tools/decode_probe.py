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
    # the kernel alone: the same call captured in a CUDA graph and replayed back to back (no host time between launches)
    dt._dc["t"] = K - 1
    torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        call()
    torch.cuda.synchronize()
    reps = 200
    e0.record()
    for _ in range(reps):
        graph.replay()
    e1.record()
    torch.cuda.synchronize()
    kern = e0.elapsed_time(e1) * 1e3 / reps
    if os.environ.get("DTB_DECODE_STAMP") == "1":
        cfg = wl["dt"]
        H, L = cfg["hidden_size"], cfg["n_layer"]
        kv = (2 * L * 3 * K * H * 2 + 255) & ~255
        base = (dt._dc["cache"].data_ptr() + 255) & ~255
        off = base - dt._dc["cache"].data_ptr() + kv
        st = dt._dc["cache"][off:off + 8 * (4 + 5 * L)].view(torch.int64).cpu().tolist()
        names = ["start", "embedded"] + [f"L{l}.{ph}" for l in range(L) for ph in ("qkv", "attn", "proj", "fc", "proj2")] + ["layers", "head"]
        print(name, "phase stamps (us since kernel start):", ", ".join(f"{nm} {(v - st[0]) / 1e3:.1f}" for nm, v in zip(names, st)), flush=True)
    print(f"{name}: decode step kernel {kern:.1f} us (graph replays back to back), event span incl. launch + clone {gpu[n // 2]:.1f} us, wall {wall[n // 2]:.1f} us "
          f"[{'FFMA' if os.environ.get('DTB_DECODE_FFMA') == '1' else 'tensor-core where supported'}]", flush=True)
