import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
name = sys.argv[1]
wl = dict(bench.WORKLOADS[name]); dev = torch.device("cuda:0")
model, parts = bench.build_ours(wl, dev, 0.0, 0); model.eval(); K = wl["K"]
b = bench.make_batch(wl, 1, seed=4321)
args = [b["states"][0].to(dev)] if wl["enc"] is None else [b["frames"][0].to(dev), b["audio"][0].to(dev)]
args += [b["actions"][0].to(dev), torch.zeros(K, device=dev), b["rtg"][0].to(dev), b["timesteps"][0].to(dev)]
for _ in range(4): model.get_action(*args)
torch.cuda.synchronize()
