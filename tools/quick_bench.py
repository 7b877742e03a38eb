"""Developer timing probe (not the bench contract): fused fwd+loss+bwd of the DT trunk per config, CUDA-event timed."""
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from util import CONFIGS, make_inputs  # noqa: E402
from mtrl_b200.model import DecisionTransformer  # noqa: E402


def run(name, B, steps=10, warmup=3, graph=False):
    cfg = dict(CONFIGS[name])
    K = cfg.pop("K")
    dev = torch.device("cuda:0")
    model = DecisionTransformer(max_length=K, max_ep_len=1000, resid_pdrop=0.0, attn_pdrop=0.0, device=dev, **cfg)
    s, a, r, t, m = (x.to(dev) for x in make_inputs(B, K, cfg["state_dim"], cfg["act_dim"], 1000, seed=1, onehot_actions=not cfg["action_tanh"]))
    fn = lambda: model.loss_and_backward(s, a, r, t, m)
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    if graph:
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            out = fn()
        fn = g.replay
        fn()
        torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    wall = (time.perf_counter() - t0) / steps * 1e3
    ms = e0.elapsed_time(e1) / steps
    tok = B * 3 * K
    print(f"{name} B={B} graph={graph}: {ms:.3f} ms/step (wall {wall:.3f}) -> {tok / ms * 1e3:.3e} tokens/s; mem {torch.cuda.max_memory_allocated() / 2**30:.1f} GiB", flush=True)


if __name__ == "__main__":
    run("c1", 64)
    run("c1", 64, graph=True)
    run("c2", 64)
    run("c3", 64)
    run("c4", 32, steps=5)
    run("c4", 256, steps=3, warmup=2)
