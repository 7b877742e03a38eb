"""Developer probe: per-tensor gradient rel-err of the fused step vs the oracle for each config."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from util import CONFIGS, build_pair, make_inputs, rel_err  # noqa: E402
from oracle.trainer import masked_action_mse  # noqa: E402

dev = torch.device("cuda:0")
for name, B in [("c2", 8), ("c2", 64), ("c3", 4), ("c4", 2)]:
    oracle, model, K = build_pair(name, dev)
    cfg = CONFIGS[name]
    s, a, r, t, m = make_inputs(B, K, cfg["state_dim"], cfg["act_dim"], 1000, seed=1234 + B, onehot_actions=not cfg["action_tanh"])
    oracle.zero_grad()
    _, ap, _ = oracle(s, a, None, r, t, m)
    loss = masked_action_mse(ap, a, m)
    loss.backward()
    out = model.loss_and_backward(s.to(dev), a.to(dev), r.to(dev), t.to(dev), m.to(dev))
    torch.cuda.synchronize()
    print(f"== {name} B={B}: loss {out[0].item():.6f} vs {loss.item():.6f}")
    errs = []
    for n, p in oracle.named_parameters():
        if p.grad is None:
            continue
        errs.append((rel_err(model.get_parameter(n).grad, p.grad), n, p.grad.norm().item()))
    errs.sort(reverse=True)
    for e, n, nr in errs[:12]:
        print(f"   {e:.4f}  {n}  |g|={nr:.3e}")
