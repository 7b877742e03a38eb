"""Developer probe: timeline of one graph-replayed training step (library event spans), to find the critical path and idle gaps."""
import ctypes as C
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import bench  # noqa: E402
from mtrl_b200 import _lib  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "c2"
wl = dict(bench.WORKLOADS[name])
dev = torch.device("cuda:0")
lib = _lib.load()
model, parts = bench.build_ours(wl, dev)
host = bench.make_batch(wl, wl["B"], seed=1234)
dev_in = {k: v.to(dev) for k, v in host.items()}
step = bench.make_step(wl, model, parts, dev_in, None, 1.0)
for _ in range(3):
    step()
torch.cuda.synchronize()
lib.dtb_profile_enable(1)
g = torch.cuda.CUDAGraph()
with torch.cuda.graph(g):
    step()
for _ in range(3):
    g.replay()
torch.cuda.synchronize()
fams = [f[0] for f in bench.PROF_FAMILIES]
n = lib.dtb_profile_spans()
spans = []
for i in range(n):
    f, a, b = C.c_int(), C.c_double(), C.c_double()
    _lib.check(lib.dtb_profile_span(i, C.byref(f), C.byref(a), C.byref(b)))
    spans.append((a.value * 1e3, b.value * 1e3, fams[f.value], i))
lib.dtb_profile_enable(0)
spans.sort()
t_end = max(s[1] for s in spans)
print(f"# {name}: {n} profiled launches, first start -> last end = {t_end:.1f} us")
busy_until = 0.0
for a, b, fam, i in spans:
    gap = a - busy_until
    flag = f"   <-- {gap:.1f} us with no profiled kernel running" if gap > 3 and busy_until > 0 else ""
    print(f"{a:8.1f} {b:8.1f} {b - a:7.1f}  {fam:28s} #{i}{flag}")
    busy_until = max(busy_until, b)
