"""Per-tensor parity table: for every benchmarked configuration at a small and at the benchmarked batch,
  * ours (bf16 tensor-core compute through libdtb200.so) vs the fp32 oracle, and
  * torch's own bf16 autocast of the oracle vs the fp32 oracle (what "any bf16 matmul" does on the same inputs),
as L2-relative errors of the loss and of every gradient tensor.  Writes gpurun_out/parity_table.json (copied to profiles/r02_parity_table.json).

    python tools/parity_table.py [c1:64 c2:8 c2:64 c3:4 c3:64 c4:2 c4:8]
"""
import copy
import json
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from util import CONFIGS, build_pair, make_inputs, rel_err  # noqa: E402

from oracle.trainer import masked_action_mse  # noqa: E402

ENC = {"c2": ((3, 64, 64), (1, 64, 64)), "c3": ((3, 64, 64), (2, 64, 64))}


def one(name, B, dev):
    cfg = CONFIGS[name]
    o_dt, dt, K = build_pair(name, dev)
    g = torch.Generator().manual_seed(7)
    if name in ENC:
        from mtrl_b200.encoder import MultimodalDecisionTransformer, MultimodalStateEncoder
        from oracle.encoders import MultimodalDecisionTransformer as OracleMM
        from oracle.encoders import MultimodalStateEncoder as OracleEncoder

        fs, as_ = ENC[name]
        torch.manual_seed(0)
        o_enc = OracleEncoder(fs, as_, "big", "concat", 1.0)
        enc = MultimodalStateEncoder(fs, as_, "big", "concat", 1.0, device=dev)
        enc.load_state_dict(o_enc.state_dict())
        oracle, model = OracleMM(o_enc, o_dt), MultimodalDecisionTransformer(enc, dt)
        frames = ((torch.randint(0, 256, (B, K, *fs), generator=g).float() / 255.0) * 2 - 1) * 0.5
        audio = torch.randn(B, K, *as_, generator=g).clamp_(-1, 1) * 0.5
        _, a, r, t, m = make_inputs(B, K, 4, cfg["act_dim"], 1000, seed=5, onehot_actions=True)
        frames[m == 0] = 0
        audio[m == 0] = 0
        step = lambda o: masked_action_mse(o(frames, audio, a, None, r, t, m)[1].float(), a, m)  # noqa: E731
        d = lambda x: x.to(dev)  # noqa: E731
        run = lambda: model.loss_and_backward(d(frames), d(audio), d(a), d(r), d(t), d(m))  # noqa: E731
    else:
        oracle, model = o_dt, dt
        s, a, r, t, m = make_inputs(B, K, cfg["state_dim"], cfg["act_dim"], 1000, seed=1234 + B, onehot_actions=not cfg["action_tanh"])
        step = lambda o: masked_action_mse(o(s, a, None, r, t, m)[1].float(), a, m)  # noqa: E731
        d = lambda x: x.to(dev)  # noqa: E731
        run = lambda: model.loss_and_backward(d(s), d(a), d(r), d(t), d(m))  # noqa: E731
    t0 = time.perf_counter()
    oracle.zero_grad()
    loss_o = step(oracle)
    loss_o.backward()
    ref = {n: p.grad.clone() for n, p in oracle.named_parameters() if p.grad is not None}
    o2 = copy.deepcopy(oracle)
    o2.zero_grad()
    with torch.autocast("cpu", dtype=torch.bfloat16):
        loss_ac = step(o2)
    loss_ac.float().backward()
    ac = {n: rel_err(p.grad, ref[n]) for n, p in o2.named_parameters() if p.grad is not None and n in ref}
    out = run()
    torch.cuda.synchronize()
    ours = {n: rel_err(model.get_parameter(n).grad, g_) for n, g_ in ref.items()}
    over = sorted(((n, e) for n, e in ours.items() if e > 2e-2), key=lambda kv: -kv[1])
    row = {"config": name, "batch": B, "tokens": B * 3 * K, "loss_oracle": float(loss_o), "loss_ours_rel_err": abs(float(out[0]) - float(loss_o)) / abs(float(loss_o)),
           "loss_autocast_rel_err": abs(float(loss_ac) - float(loss_o)) / abs(float(loss_o)), "n_tensors": len(ours),
           "max_ours": max(ours.values()), "max_autocast": max(ac.values()), "median_ours": sorted(ours.values())[len(ours) // 2],
           "n_over_2e-2_ours": len(over), "n_over_2e-2_autocast": sum(1 for e in ac.values() if e > 2e-2),
           "over_2e-2": [{"tensor": n, "ours": round(e, 4), "autocast": round(ac.get(n, float("nan")), 4)} for n, e in over],
           "tensors": {n: {"ours": ours[n], "autocast": ac.get(n)} for n in sorted(ours)}, "seconds": time.perf_counter() - t0}
    print(f"{name} B={B}: loss err {row['loss_ours_rel_err']:.2e}; grads max ours {row['max_ours']:.3f} / autocast {row['max_autocast']:.3f}; "
          f"{len(over)} of {len(ours)} tensors over 2e-2 (autocast: {row['n_over_2e-2_autocast']})", flush=True)
    return row


def main():
    dev = torch.device("cuda:0")
    cases = sys.argv[1:] or ["c1:64", "c2:8", "c2:64", "c3:4", "c3:64", "c4:2", "c4:8"]
    rows = []
    for c in cases:
        name, B = c.split(":")
        rows.append(one(name, int(B), dev))
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    with open(os.path.join(ROOT, "gpurun_out", "parity_table.json"), "w") as fh:
        json.dump({"what": "L2 relative error per gradient tensor vs the fp32 oracle: `ours` = libdtb200.so (bf16 tensor cores), `autocast` = "
                           "torch.autocast(bfloat16) of the oracle itself on the same inputs; north-star bound 2e-2", "rows": rows}, fh, indent=1)


if __name__ == "__main__":
    main()
