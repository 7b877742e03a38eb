"""Where does get_action's wall-clock latency go?  GPU time of the cached graph (events) vs host time of each step of _ActionRunner.run.
Usage: python tools/get_action_probe.py [c1|c2]"""
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "c2"
    wl = dict(bench.WORKLOADS[name])
    dev = torch.device("cuda:0")
    model, parts = bench.build_ours(wl, dev, 0.0, 0)
    model.eval()
    K = wl["K"]
    b = bench.make_batch(wl, 1, seed=4321)
    args = [b["states"][0].to(dev)] if wl["enc"] is None else [b["frames"][0].to(dev), b["audio"][0].to(dev)]
    args += [b["actions"][0].to(dev), torch.zeros(K, device=dev), b["rtg"][0].to(dev), b["timesteps"][0].to(dev)]
    for _ in range(20):
        model.get_action(*args)
    torch.cuda.synchronize()
    runner = model.__dict__["_action_runner"]
    # GPU time of the graph alone
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(100):
        runner.graph.replay()
    e1.record()
    torch.cuda.synchronize()
    gpu_us = e0.elapsed_time(e1) * 10
    # wall time of the whole call and of the call without the final sync (host-side cost of enqueueing)
    tot, host = [], []
    for _ in range(200):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        model.get_action(*args)
        t1 = time.perf_counter()
        torch.cuda.current_stream().synchronize()
        t2 = time.perf_counter()
        host.append((t1 - t0) * 1e6)
        tot.append((t2 - t0) * 1e6)
    tot.sort(); host.sort()
    # per-kernel-family time INSIDE the graph (the library's event nodes, as bench.py's roofline pass uses them): a second runner captured
    # with profiling on; the events add a few microseconds per launch, so read shares
    import ctypes as C

    from mtrl_b200 import _lib

    lib = _lib.load()
    lib.dtb_profile_enable(1)
    model.__dict__.pop("_action_runner", None)
    for _ in range(3):
        model.get_action(*args)
        torch.cuda.synchronize()
    fam = {}
    for i, (fname, _) in enumerate(bench.PROF_FAMILIES):
        ms, work, n = C.c_double(), C.c_double(), C.c_longlong()
        _lib.check(lib.dtb_profile_report(i, C.byref(ms), C.byref(work), C.byref(n)))
        if n.value:
            fam[fname.split(",")[0]] = (round(ms.value * 1e3 / 3, 1), n.value // 3)
    lib.dtb_profile_enable(0)
    model.__dict__.pop("_action_runner", None)
    print(f"{name}: in-graph family times (us per call, launches): {fam}")
    print(f"{name}: graph GPU time {gpu_us:.1f} us (back-to-back replays), call wall p50 {tot[100]:.1f} us, host part (until get_action returns) p50 {host[100]:.1f} us")


if __name__ == "__main__":
    main()
