"""Benchmark contract: `python bench.py --gpus N --steps K --warmup W [--impl reference] [--workload c1|c2|c3|c4]`.

Metric (BASELINE.json): Decision-Transformer training tokens/s (forward + masked action MSE + backward; + gradient all-reduce for N>1;
optimizer excluded) on synthetic trajectories.  Default workload = BASELINE.json configs[1], the Minecraft2d
multimodal DT (3x64x64 frames + 1x64x64 spectrogram -> two conv towers -> K=30, H=128, 3 layers, batch 64 per GPU): weak scaling.
One JSON line on rank 0.  `value` = device-resident inputs; `e2e` = through the public Python API / C ABI with pinned HOST buffers,
host->device copies and the loss read-back inside the timed region.  `--impl reference` times the fp32 PyTorch oracle (for the
towers numerically the reference's own conv_factory) on the host cores.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import torch  # noqa: E402

WORKLOADS = {
    # name: DT config + per-GPU batch + encoder shapes (None = vector states)      — SURVEY.md §8 config table
    "c1": dict(dt=dict(state_dim=11, act_dim=3, hidden_size=128, n_layer=3, n_head=1, activation_function="relu", action_tanh=True), K=20, B=64,
               enc=None, label="gym DT hopper-shaped: state_dim=11 act_dim=3 K=20 embed=128 3 layers 1 head batch=64"),
    "c2": dict(dt=dict(state_dim=1024, act_dim=4, hidden_size=128, n_layer=3, n_head=1, activation_function="relu", action_tanh=False), K=30, B=64,
               enc=dict(frame_shape=(3, 64, 64), audio_shape=(1, 64, 64)),
               label="Minecraft2d multimodal DT: 64x64 RGB frame CNN + audio spectrogram encoder, K=30 embed=128 3 layers, batch=64/GPU"),
    "c3": dict(dt=dict(state_dim=1024, act_dim=18, hidden_size=256, n_layer=6, n_head=8, activation_function="relu", action_tanh=False), K=50, B=64,
               enc=dict(frame_shape=(3, 64, 64), audio_shape=(2, 64, 64)),
               label="Skeleton+ stereo-audio DT: 2-channel spectrogram + video encoder, K=50 embed=256 6 layers 8 heads, batch=64/GPU"),
    "c4": dict(dt=dict(state_dim=11, act_dim=3, hidden_size=768, n_layer=12, n_head=12, activation_function="gelu_new", action_tanh=True), K=200,
               B=256, enc=None, label="scaled DT: K=200 embed=768 12 layers 12 heads bf16 training batch=256"),
}
MAX_EP_LEN = 1000


def make_batch(wl, B, seed, with_index=False):
    """Seeded synthetic trajectories of SURVEY.md §8d, on the host.  Mono-audio workloads (Minecraft2d: c2 / c5) draw every timestep's
    spectrogram from a table of three images, as the reference environment does (Minecraft.py:145-159: mfcc_target_good / _bad /
    _no_listen, built once by _init_mfcc :222-225); `audio` is always the expanded fp32 tensor, with_index=True adds `audio_index`
    (int64 [B, K]) and `audio_table` (fp32 [3, 1, H, W]) for feeders that ship indices instead of images."""
    from util import make_inputs

    dt = wl["dt"]
    onehot = not dt["action_tanh"]
    s, a, r, t, m = make_inputs(B, wl["K"], dt["state_dim"] if wl["enc"] is None else 4, dt["act_dim"], MAX_EP_LEN, seed=seed, onehot_actions=onehot)
    out = dict(actions=a, rtg=r, timesteps=t, mask=m)
    if wl["enc"] is None:
        out["states"] = s
    else:
        g = torch.Generator().manual_seed(seed + 99)
        K = wl["K"]
        # uint8 pixels, the format the frames exist in before the reference's decode img * 2 / 255 - 1 (Minecraft.py:333-335); our arm
        # ships the pixels and decodes in-kernel (bit-identical, tests/test_encoder_gpu.py), the CPU arm decodes on the host
        out["frames"] = torch.randint(0, 256, (B, K, *wl["enc"]["frame_shape"]), generator=g, dtype=torch.uint8)
        ashape = tuple(wl["enc"]["audio_shape"])
        if ashape[0] == 1:
            table = torch.rand(3, *ashape, generator=g) * 2 - 1  # min-max normalised MFCC images in [-1, 1] (Minecraft.py:239-243)
            idx = torch.randint(0, 3, (B, K), generator=g)
            out["audio"] = table[idx]
            if with_index:
                out["audio_index"], out["audio_table"] = idx, table
        else:  # stereo spectrograms (Skeleton+): arbitrary per-timestep images
            out["audio"] = torch.randn(B, K, *ashape, generator=g).clamp_(-1, 1)
    return out


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""

    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.index, self.samples, self._stop, self._t = index, [], threading.Event(), None

    def start(self):
        def run():
            while not self._stop.is_set():
                try:
                    o = subprocess.run(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-i", str(self.index)],
                                       capture_output=True, text=True, timeout=5).stdout.strip()
                    if o:
                        self.samples.append([x.strip() for x in o.split(",")])
                except Exception:
                    pass
                self._stop.wait(0.2)

        self._t = threading.Thread(target=run, daemon=True)
        self._t.start()

    def stop(self):
        self._stop.set()
        if self._t:
            self._t.join(timeout=6)
        mhz = sorted(int(s[0]) for s in self.samples if s and s[0].isdigit())
        mx = [int(s[1]) for s in self.samples if len(s) > 1 and s[1].isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for s in self.samples for n, v in zip(names, s[2:6]) if v.strip().lower() == "active"})
        return {"sm_mhz": mhz[len(mhz) // 2] if mhz else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(self.samples)}


# ------------------------------------------------------------------------------------------------------------------------------
# our arm
# ------------------------------------------------------------------------------------------------------------------------------
ATTN_TYPE = None  # --attn-type: modality attention between the towers and the fusion (agents.py:135-136); None = the reference default


def build_ours(wl, dev, pdrop=0.0, rank=0):
    from mtrl_b200.encoder import MultimodalDecisionTransformer, MultimodalStateEncoder
    from mtrl_b200.model import DecisionTransformer

    torch.manual_seed(0)  # shared seed -> identical replicas (ppo_atari_multigpu.py:200); broadcast below makes it strict
    dt = DecisionTransformer(max_length=wl["K"], max_ep_len=MAX_EP_LEN, resid_pdrop=pdrop, attn_pdrop=pdrop, device=dev, **wl["dt"])
    dt.train()  # training mode: dropout (upstream default 0.1) is drawn inside the kernels, fresh masks on every step / graph replay
    dt.set_dropout_seed(1000 + rank)  # independent masks per replica
    if wl["enc"] is None:
        return dt, [dt]
    enc = MultimodalStateEncoder(wl["enc"]["frame_shape"], wl["enc"]["audio_shape"], "big", "concat", 1.0 / 255.0, device=dev, attn_type=ATTN_TYPE)
    return MultimodalDecisionTransformer(enc, dt), [dt, enc]


def make_step(wl, model, parts, dev_in, reducer, grad_scale):
    """Returns step() -> device tensor [loss, n_valid]: forward + masked action MSE + backward (+ overlapped all-reduce)."""
    d = dev_in
    if wl["enc"] is None:
        dt = parts[0]
        cb = None if reducer is None else reducer.callback_for(model)

        def step():
            out = model.loss_and_backward(d["states"], d["actions"], d["rtg"], d["timesteps"], d["mask"], grad_scale=grad_scale, on_stage_done=cb)
            if reducer is not None:
                reducer.finish()
            return out
    else:
        dt, enc = parts
        cb = None if reducer is None else reducer.callback_for(model)  # trunk buckets + the towers' FC range overlap the backward

        def step():
            out = model.loss_and_backward(d["frames"], d["audio"], d["actions"], d["rtg"], d["timesteps"], d["mask"], grad_scale=grad_scale,
                                          on_stage_done=cb)
            if reducer is not None:
                reducer.finish()
            return out
    return step


def timed_loop(fn, steps, dev, flush=None, dist_on=False):
    """K steps, each bracketed by CUDA events on the launching stream; barrier + synchronize on both sides; max over ranks."""
    import torch.distributed as dist

    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    if dist_on:
        dist.barrier()
    torch.cuda.synchronize(dev)
    t0 = time.perf_counter()
    for i in range(steps):
        if flush is not None:
            flush.zero_()
        ev[i][0].record()
        fn()
        ev[i][1].record()
    torch.cuda.synchronize(dev)
    wall = time.perf_counter() - t0
    if dist_on:
        dist.barrier()
    if flush is None:  # contiguous region: first start -> last end (includes launch gaps)
        total_ms = ev[0][0].elapsed_time(ev[-1][1])
    else:  # flush kernels sit between the per-step brackets and are excluded
        total_ms = sum(a.elapsed_time(b) for a, b in ev)
    if dist_on:
        t = torch.tensor([total_ms], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        total_ms = float(t.item())
    return total_ms / steps, wall / steps * 1e3


def algorithmic_flops_per_step(wl, B):
    """fwd+bwd FLOPs of one step: trunk 3 * tokens * L * (8H^2 + 4HI + 2TH) (causal attention counted half),
    towers 3x forward (conv1 has no dgrad: 2x)."""
    dt, K = wl["dt"], wl["K"]
    H, L, T = dt["hidden_size"], dt["n_layer"], 3 * K
    inner = 4 * H
    trunk = 3.0 * B * T * L * (8 * H * H + 4 * H * inner + 2 * T * H)
    towers = 0.0
    if wl["enc"] is not None:
        for shape in (wl["enc"]["frame_shape"], wl["enc"]["audio_shape"]):
            c, h, w = shape
            specs, f = [(32, 8, 4), (64, 4, 2), (64, 3, 1)], 0.0
            for i, (co, k, s) in enumerate(specs):
                h, w = (h - k) // s + 1, (w - k) // s + 1
                f += (2.0 if i == 0 else 3.0) * 2 * h * w * co * k * k * c
                c = co
            f += 3.0 * 2 * c * h * w * 512
            towers += f * B * K
        towers += 3.0 * 2 * B * K * dt["state_dim"] * H  # embed_state projection of the wide state
    return trunk + towers


def run_ours(args, wl, rank, world, local_rank):
    import torch.distributed as dist

    from mtrl_b200 import _lib
    from mtrl_b200.dp import GradAllReducer, broadcast_parameters, replicas_consistent

    dev = torch.device("cuda", local_rank)
    torch.cuda.set_device(dev)
    lib = _lib.load()
    _lib.check(lib.dtb_check_device())
    B, K = wl["B"], wl["K"]
    model, parts = build_ours(wl, dev, args.dropout, rank)
    broadcast_parameters([p.flat_params for p in parts])
    host = make_batch(wl, B, seed=1234 + rank, with_index=True)  # per-rank data seeds (ppo_atari_multigpu.py:197-199); weak scaling: B per GPU fixed
    # `pinned`: every input as a full tensor (what the step reads; the reference's layouts).  `feed`: what the e2e loop ships per step — the
    # same, except that a table-drawn input (Minecraft2d's audio channel: one of three MFCC images per timestep) travels as its indices
    # and is gathered from the device-resident table on the copy stream (HostBatchPipeline(indexed=...): lossless, the step sees `audio`)
    pinned = {k: v.contiguous().pin_memory() for k, v in host.items() if k not in ("audio_index", "audio_table")}
    indexed, feed = None, pinned
    if "audio_index" in host:
        indexed = {"audio": ("audio_index", host["audio_table"].to(dev))}
        feed = {k: v for k, v in pinned.items() if k != "audio"}
        feed["audio_index"] = host["audio_index"].contiguous().pin_memory()
    reducer = GradAllReducer(bucket_bytes=args.bucket_mb << 20) if world > 1 else None

    def make_step_for(dev_in):
        raw_step = make_step(wl, model, parts, dev_in, reducer, 1.0 / world)
        if not args.restage_each_step:
            # bf16 weight staging is done by whoever changes the fp32 master (FusedClipAdamW.step() rewrites the DT shadow in its own pass
            # and re-stages the towers): it is timed with the optimizer (`train_step_with_optimizer`), not with forward + backward
            return raw_step

        def step():
            for p in parts:
                p._staged_version = -1  # round-1 semantics: fp32 -> bf16 weight staging inside every timed step
            return raw_step()

        return step

    from mtrl_b200.pipeline import HostBatchPipeline

    graphed = not args.no_graph  # NCCL all-reduces on the communication stream are captured into the step's graph too
    try:
        pipe = HostBatchPipeline(make_step_for, pinned, dev, use_graph=graphed, warmup=max(args.warmup, 3), indexed=indexed)
    except Exception as e:  # noqa: BLE001
        if not graphed:
            raise
        print(f"[bench] CUDA graph capture failed ({e}); timing plain launches", file=sys.stderr)
        torch.cuda.synchronize(dev)
        graphed = False
        pipe = HostBatchPipeline(make_step_for, pinned, dev, use_graph=False, warmup=max(args.warmup, 3), indexed=indexed)
    slot = pipe.slots[0]
    dev_in, step = slot["dev"], slot["step"]
    h2d = pipe.bytes_per_batch
    l0 = lib.dtb_launch_count()
    out = step()
    torch.cuda.synchronize(dev)
    launches_per_step = int(lib.dtb_launch_count() - l0)
    fn = slot["graph"].replay if graphed else step
    # working set per step (inputs + saved activations) vs the 126 MB L2
    ws_bytes = h2d + sum(sum(w.numel() for w in p._ws.values()) for p in parts)
    flush = None
    if ws_bytes < 2 * 126e6 or args.flush_l2:
        flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    for _ in range(args.warmup):
        fn()
    clocks = ClockSampler(local_rank)
    if rank == 0:
        clocks.start()
    ms, wall_ms = timed_loop(fn, args.steps, dev, flush, dist_on=world > 1)
    tokens = B * 3 * K * world
    value = tokens / ms * 1e3

    # ---- e2e: pinned host buffers -> H2D (copy stream, double-buffered: batch i+1 copies while step i computes) -> step -> loss D2H + sync,
    # every step, through the package's own feeder (mtrl_b200.pipeline.HostBatchPipeline) ----
    def e2e_loop(batch, deferred=False):
        pipe.deferred_loss = deferred
        pipe.prefetch(batch)
        for _ in range(3):
            pipe.step(batch)
        if deferred:
            pipe.flush()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)
        t0 = time.perf_counter()
        last = None
        for _ in range(args.steps):
            last = pipe.step(batch)
        if deferred:
            last = pipe.flush()
        torch.cuda.synchronize(dev)
        ms_ = (time.perf_counter() - t0) / args.steps * 1e3
        pipe.step(None)  # drain the prefetched slot so that the next loop starts clean
        if deferred:
            pipe.flush()
        torch.cuda.synchronize(dev)
        pipe.deferred_loss = False
        if world > 1:
            t = torch.tensor([ms_], device=dev, dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms_ = float(t.item())
        return ms_, last

    e2e_ms, loss_val = e2e_loop(feed)
    # the same loop with the loss read deferred by one step (HostBatchPipeline(deferred_loss=True) semantics): every loss still crosses to the
    # host inside the timed region, but the host enqueues step i + 1 before it waits for loss i
    e2e_def_ms, loss_deferred = e2e_loop(feed, deferred=True)
    # round 1 / 2 comparison: every input as a full tensor in the reference's layouts (the audio channel as fp32 images)
    e2e_full_ms = None
    if indexed:
        order, pipe._copy_first = pipe._copy_first, True  # 55 MB per step: the copy is the longer pole, enqueue it first (as rounds 1-2 did)
        e2e_full_ms = e2e_loop(pinned)[0]
        pipe._copy_first = order
    h2d_full = sum(v.numel() * v.element_size() for v in pinned.values())
    clock_info = clocks.stop() if rank == 0 else None
    consistent = replicas_consistent([p.flat_params for p in parts])

    # ---- extra (not the headline metric, which excludes the optimizer as BASELINE.json's does): the full train_step() body, i.e. the same
    # step followed by the fused clip_grad_norm_(0.25) + AdamW over the flat buffers (mtrl_b200.optim.FusedClipAdamW), in one graph ----
    with_opt = None
    if not args.no_optimizer_extra:
        try:
            from mtrl_b200.optim import FusedClipAdamW

            opt = FusedClipAdamW(parts, lr=1e-4, weight_decay=1e-4, max_norm=0.25)
            raw = make_step(wl, model, parts, dev_in, reducer, 1.0 / world)  # no forced re-staging: the optimizer rewrites the bf16 shadow

            def full_step():
                out_ = raw()
                opt.step()
                return out_

            for _ in range(3):
                full_step()
            torch.cuda.synchronize(dev)
            fn2 = full_step
            if graphed:
                g2 = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g2):
                    full_step()
                fn2 = g2.replay
            for _ in range(3):
                fn2()
            ms2, _ = timed_loop(fn2, args.steps, dev, flush, dist_on=world > 1)
            with_opt = {"ms_per_step": ms2, "value": tokens / ms2 * 1e3, "unit": "tokens/s", "optimizer_steps": opt.step_count,
                        "what": "fwd + loss + bwd + fused clip_grad_norm_(0.25) + AdamW (3 extra launches), same graph",
                        "replicas_consistent": replicas_consistent([p.flat_params for p in parts])}
        except Exception as e:  # noqa: BLE001
            with_opt = {"error": str(e)[:200]}

    res = {
        "metric": "DT train tokens/s (fwd+bwd)", "value": value, "unit": "tokens/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": {"workload": wl["label"], "name": args.workload, "batch_per_gpu": B, "global_batch": B * world, "seq_len": 3 * K,
                   "parallelism": f"dp{world}", "cuda_graph": graphed, "dropout": args.dropout, "attn_type": ATTN_TYPE,
                   "l2": ("explicit 256 MiB flush between steps" if flush is not None else
                          f"no flush: per-step working set {ws_bytes / 1e6:.0f} MB > 126 MB L2"),
                   "timed": ("weight staging + " if args.restage_each_step else "") + "forward + masked action MSE + backward" +
                            (" + overlapped NCCL all-reduce" if world > 1 else "") +
                            ("" if args.restage_each_step else " (bf16 weight staging belongs to the optimizer step: train_step_with_optimizer)")},
        "e2e": {"value": tokens / e2e_ms * 1e3, "unit": "tokens/s", "ms_per_step": e2e_ms, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 8,
                "host_format": ("pinned host buffers: video frames as uint8 pixels (decoded in-kernel), the audio channel as indices into the environment's "
                                "table of three MFCC images (REF Minecraft.py:145-159, 222-225; gathered on the device, lossless), everything else fp32 / int64 "
                                "in the reference's layouts" if indexed else
                                "pinned host buffers in the reference's layouts (fp32; video frames as uint8 pixels, decoded in-kernel)"),
                "full_tensor_inputs": (None if e2e_full_ms is None else
                                       {"value": tokens / e2e_full_ms * 1e3, "unit": "tokens/s", "ms_per_step": e2e_full_ms, "h2d_bytes_per_step": h2d_full,
                                        "what": "same loop with the audio channel shipped as fp32 images every step (the e2e format of rounds 1-2): bound by the "
                                                "PCIe link, not by the kernels"}),
                "overlap": "H2D of batch i+1 on a copy stream while step i computes (double-buffered device slots)",
                "loss_read": "the step's own loss, one stream sync per step (upstream train_step semantics)",
                "deferred_loss_read": {"value": tokens / e2e_def_ms * 1e3, "unit": "tokens/s", "ms_per_step": e2e_def_ms, "last_loss": loss_deferred,
                                       "what": "same loop, HostBatchPipeline(deferred_loss=True): step() returns the previous step's loss, so the host "
                                               "enqueues step i+1 before waiting for loss i; every loss still crosses to the host in the timed region"}},
        "gpu_launches": launches_per_step * args.steps, "launches_per_step": launches_per_step,
        "clocks": clock_info, "loss": loss_val, "wall_ms_per_step": wall_ms,
        "algorithmic_tflops_per_s": algorithmic_flops_per_step(wl, B * world) / ms / 1e9, "replicas_consistent": consistent,
        "train_step_with_optimizer": with_opt,
    }
    return res, dict(model=model, parts=parts, dev=dev, dev_in=dev_in, step=step, world=world, graph=slot["graph"] if graphed else None,
                     graph_out=slot["out"])


# ------------------------------------------------------------------------------------------------------------------------------
# CPU arm: the fp32 oracle on the host cores (cpu_baseline of our line, and the whole `--impl reference` run)
# ------------------------------------------------------------------------------------------------------------------------------
def build_oracle(wl, pdrop=0.0):
    from oracle.decision_transformer import DecisionTransformer as ODT
    from oracle.encoders import MultimodalDecisionTransformer as OMM
    from oracle.encoders import MultimodalStateEncoder as OEnc

    torch.manual_seed(0)
    dt = ODT(max_length=wl["K"], max_ep_len=MAX_EP_LEN, resid_pdrop=pdrop, attn_pdrop=pdrop, **wl["dt"])
    if wl["enc"] is None:
        return dt
    return OMM(OEnc(wl["enc"]["frame_shape"], wl["enc"]["audio_shape"], "big", "concat", 1.0 / 255.0, attn_type=ATTN_TYPE), dt)


def oracle_step(wl, model, b):
    from oracle.trainer import masked_action_mse

    model.zero_grad(set_to_none=True)
    if wl["enc"] is None:
        _, ap, _ = model(b["states"], b["actions"], None, b["rtg"], b["timesteps"], b["mask"])
    else:
        frames = b["frames"].float() * 2.0 / 255.0 - 1  # the reference's pixel decode (Minecraft.py:333-335), on the host for the CPU arm
        _, ap, _ = model(frames, b["audio"], b["actions"], None, b["rtg"], b["timesteps"], b["mask"])
    loss = masked_action_mse(ap, b["actions"], b["mask"])
    loss.backward()
    return float(loss.detach())


def run_cpu(wl, steps, warmup, budget_s=20.0, pdrop=0.0):
    """Times the oracle on all host threads on a bounded sample: the same workload at a reduced batch chosen so that
    warmup + steps fit in ~budget_s; tokens/s = sample tokens / time (linear in batch)."""
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    model = build_oracle(wl, pdrop)
    model.train()  # nn.Dropout(p) active, as in the reference's train_step
    K = wl["K"]
    Bs = min(wl["B"], 4)
    b = make_batch(wl, Bs, seed=1234)
    t0 = time.perf_counter()
    oracle_step(wl, model, b)
    probe = time.perf_counter() - t0  # includes first-call overheads: conservative
    t0 = time.perf_counter()
    oracle_step(wl, model, b)
    probe = min(probe, time.perf_counter() - t0)
    per_traj = probe / Bs
    want = budget_s / max(1, steps + warmup) / max(per_traj, 1e-6)
    Bs = int(max(1, min(wl["B"], want)))
    b = make_batch(wl, Bs, seed=1234)
    for _ in range(warmup):
        oracle_step(wl, model, b)
    times = []
    for _ in range(steps):
        t0 = time.perf_counter()
        loss = oracle_step(wl, model, b)
        times.append(time.perf_counter() - t0)
    dt_s = sum(times) / steps
    tokens = Bs * 3 * K
    times.sort()
    return {"value": tokens / dt_s, "unit": "tokens/s", "cores": cores, "kind": "port", "ms_per_step": dt_s * 1e3,
            "ms_per_step_min": times[0] * 1e3, "ms_per_step_median": times[len(times) // 2] * 1e3,
            "sample": f"fp32 PyTorch oracle (towers = reference conv_factory math), fwd+loss+bwd, batch {Bs} of {wl['B']} trajectories x K={K} "
                      f"({tokens} tokens/step), dropout {pdrop}, {steps} steps after {warmup} warm-up, {cores} threads", "loss": loss}


def load_peaks():
    """Roofline denominators: MEASURED_PEAKS.json (driver-written for this pool's B200s), else the fallback of B200_PROFILING.md."""
    peaks = {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "source": "fallback (B200_PROFILING.md)"}
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as fh:
            mp = json.load(fh)
        peaks = {"hbm_gbs": float(mp["hbm_gbs"]), "bf16_tflops": float(mp["bf16_tflops"]),
                 "bf16_tflops_sustained": float(mp.get("bf16_tflops_sustained", mp["bf16_tflops"])), "source": "measured (MEASURED_PEAKS.json)"}
    except Exception:
        pass
    # machine balance: a GEMM shape with fewer flops per algorithmic byte than this cannot be tensor-bound (sustained peak / copy peak)
    peaks["balance_flop_per_byte"] = peaks["bf16_tflops_sustained"] * 1e12 / (peaks["hbm_gbs"] * 1e9)
    return peaks


def load_traffic(workload):
    """DRAM bytes per launch (dram__bytes_read.sum + dram__bytes_write.sum) per kernel family, produced from the committed ncu CSV by
    tools/ncu_traffic.py -> profiles/r02_traffic.json (a profiler pass cannot run inside the timed bench)."""
    try:
        with open(os.path.join(ROOT, "profiles", "r02_traffic.json")) as fh:
            return json.load(fh).get(workload, {})
    except Exception:
        return {}


# order = the library's ProfFamily enum (csrc/common.cuh)
PROF_FAMILIES = [("gemm_tcgen05_kernel (+pair)", "tensor"), ("conv1_fwd_kernel", "hbm"), ("conv1_wgrad_kernel", "hbm"), ("attn_fwd_tc_kernel", "tensor"),
                 ("attn_bwd_dq+dkv_tc_kernel", "tensor"), ("ln_fwd_kernel", "hbm"), ("ln_bwd_kernel", "hbm"), ("embed_fwd_kernel", "hbm"),
                 ("colsum_kernel", "hbm"), ("im2col_nhwc_kernel", "hbm"), ("col2im_nhwc_relu_kernel", "hbm"),
                 ("gemm_tcgen05_kernel, HBM-bound shapes", "hbm"), ("convn_kernel<fwd>", "hbm"), ("convn_kernel<dgrad>", "hbm"),
                 ("convn_wgrad_kernel", "hbm")]


def measure_roofline(args, wl, ctx, reps=5):
    """Per-kernel-family durations measured live with CUDA events recorded by the library on the stream each kernel is launched on
    (plain launches, same process, right after the timed region).  The dominant family (largest share of the step) is reported:
    achieved = its algorithmic flops (tensor) or bytes (hbm) / its summed launch durations; peak = MEASURED_PEAKS.json."""
    import ctypes as C

    from mtrl_b200 import _lib

    lib = _lib.load()
    peaks = load_peaks()
    lib.dtb_profile_set_balance.argtypes = [C.c_double]
    lib.dtb_profile_set_balance.restype = None
    lib.dtb_profile_set_balance(peaks["balance_flop_per_byte"])
    step = ctx["step"]
    step()
    torch.cuda.synchronize()

    def report():
        out = {}
        for i in range(len(PROF_FAMILIES)):
            ms, work, n = C.c_double(), C.c_double(), C.c_longlong()
            _lib.check(lib.dtb_profile_report(i, C.byref(ms), C.byref(work), C.byref(n)))
            out[i] = (ms.value, work.value, n.value)
        return out

    acc = {i: [0.0, 0.0, 0] for i in range(len(PROF_FAMILIES))}
    mode = "inside the CUDA graph of the step (external event-record nodes)"
    lib.dtb_profile_enable(1)
    try:
        if args.no_graph or ctx.get("world", 1) > 1:
            raise RuntimeError("no graph")
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            step()
        for _ in range(reps):
            g.replay()
            torch.cuda.synchronize()
            for i, (ms, work, n) in report().items():
                acc[i][0] += ms; acc[i][1] += work; acc[i][2] += n
    except Exception:
        torch.cuda.synchronize()
        mode = "plain launches (includes host launch gaps when the step is launch-bound)"
        lib.dtb_profile_enable(1)
        for _ in range(reps):
            step()
        torch.cuda.synchronize()
        for i, (ms, work, n) in report().items():
            acc[i] = [ms, work, n]
    lib.dtb_profile_enable(0)
    traffic = load_traffic(args.workload)
    fams = []
    for i, (name, bound) in enumerate(PROF_FAMILIES):
        ms, work, n = acc[i]
        if n:
            fams.append(dict(kernel=name, bound=bound, ms_per_step=ms / reps, launches_per_step=n // reps, work_per_step=work / reps))
    if not fams:
        return None
    total = sum(f["ms_per_step"] for f in fams)
    for f in fams:
        f["share_of_profiled"] = f["ms_per_step"] / total
        if f["bound"] == "tensor":
            f["achieved"], f["unit"], f["peak"] = f["work_per_step"] / f["ms_per_step"] / 1e9, "TFLOP/s", peaks["bf16_tflops"]
        else:
            f["achieved"], f["unit"], f["peak"] = f["work_per_step"] / f["ms_per_step"] / 1e6, "GB/s", peaks["hbm_gbs"]
        f["frac"] = f["achieved"] / f["peak"]
        f["algorithmic_per_launch"] = f["work_per_step"] / max(1, f["launches_per_step"])
        tr = traffic.get(f["kernel"])
        f["traffic"] = tr["dram_bytes_per_launch"] if tr else None
    if args.families_only:
        return {"families": sorted(fams, key=lambda f: -f["ms_per_step"]), "profiled_ms_per_step": total, "how": mode}
    top = max(fams, key=lambda f: f["ms_per_step"])
    tr = traffic.get(top["kernel"]) or {}
    out = {"kernel": top["kernel"], "bound": top["bound"], "achieved": top["achieved"], "peak": top["peak"], "unit": top["unit"],
           "frac": top["frac"], "traffic": tr.get("dram_bytes_per_launch"), "traffic_source": tr.get("source"),
           "algorithmic_per_launch": top["algorithmic_per_launch"], "peak_source": peaks["source"],
           "peaks": {k: peaks[k] for k in ("hbm_gbs", "bf16_tflops", "bf16_tflops_sustained", "balance_flop_per_byte")},
           "launches_per_step": top["launches_per_step"], "avg_launch_us": top["ms_per_step"] / max(1, top["launches_per_step"]) * 1e3,
           "how": f"CUDA events around every launch of the family on its own stream, {reps} steps, {mode}; algorithmic work per DESIGN.md §3; "
                  "family spans on concurrent streams overlap, so their sum exceeds the step (see families_serialised)",
           "families": sorted(fams, key=lambda f: -f["ms_per_step"])}
    if ctx.get("world", 1) == 1 and not args.no_extras_serial:
        # the same table with the library's side streams off (DTB_NO_SIDE_STREAM=1): every kernel alone on one stream, spans add up
        try:
            cmd = [sys.executable, os.path.abspath(__file__), "--workload", args.workload, "--steps", "5", "--warmup", "3", "--families-only",
                   "--dropout", str(args.dropout)] + (["--batch", str(args.batch)] if args.batch else []) + (["--attn-type", ATTN_TYPE] if ATTN_TYPE else [])
            r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=dict(os.environ, DTB_NO_SIDE_STREAM="1"))
            ser = json.loads(r.stdout.strip().splitlines()[-1])
            out["families_serialised"] = {"what": "DTB_NO_SIDE_STREAM=1: the whole step on one stream, family times add up to the step",
                                          "ms_per_step": ser.get("ms_per_step"), "profiled_ms_per_step": ser["roofline"]["profiled_ms_per_step"],
                                          "families": [{k: f.get(k) for k in ("kernel", "bound", "ms_per_step", "launches_per_step", "achieved", "unit", "frac")}
                                                       for f in ser["roofline"]["families"]]}
        except Exception as e:  # noqa: BLE001
            out["families_serialised"] = {"error": str(e)[:200]}
    return out


def measure_dp_parity(args, wl, ctx, rank, world):
    """Correctness evidence for the N > 1 line (not only `replicas_consistent`): one un-timed step in eval mode (dropout off) on every
    rank's own batch through the same GradAllReducer + grad_scale = 1/world path, then rank 0 compares its all-reduced flat gradients with
    the uniformly averaged per-shard gradients of the fp32 oracle (the reference idiom, cleanrl/ppo_atari_multigpu.py:345-359), tensor
    by tensor.  Every rank runs the step (it contains the collective); only rank 0 runs the oracle."""
    from util import rel_err

    model, parts, dev_in = ctx["model"], ctx["parts"], ctx["dev_in"]
    from mtrl_b200.dp import GradAllReducer

    was_training = [p.training for p in parts]
    for p in parts:
        p.eval()
    red = GradAllReducer(bucket_bytes=args.bucket_mb << 20)
    step = make_step(wl, model, parts, dev_in, red, 1.0 / world)
    host = make_batch(wl, wl["B"], seed=1234 + rank)
    for k, v in host.items():
        dev_in[k].copy_(v)
    out = step()
    torch.cuda.synchronize()
    res = None
    if rank == 0:
        oracle = build_oracle(wl, 0.0)
        oracle.load_state_dict({k: v.detach().float().cpu() for k, v in model.state_dict().items()})
        oracle.eval()
        ref, losses = {}, []
        for r in range(world):
            losses.append(oracle_step(wl, oracle, make_batch(wl, wl["B"], seed=1234 + r)))
            for n, p in oracle.named_parameters():
                if p.grad is not None:
                    ref[n] = ref.get(n, 0) + p.grad.detach() / world
        errs = {n: rel_err(model.get_parameter(n).grad, g) for n, g in ref.items()}
        worst = max(errs.items(), key=lambda kv: kv[1])
        over = {n: round(e, 4) for n, e in errs.items() if e > 2e-2}
        res = {"max_rel_err": worst[1], "worst_tensor": worst[0], "tensors": len(errs), "tensors_over_2e-2": over,
               "loss_rank0": float(out[0]), "oracle_loss_rank0": losses[0],
               "what": f"rank 0's all-reduced flat gradients vs the mean over {world} ranks of the fp32 oracle's per-shard gradients (dropout off, "
                       "bf16 compute vs fp32: tensors behind a ReLU mask may exceed 2e-2 exactly as at N = 1, see BASELINE.md §4)"}
    for p, t in zip(parts, was_training):
        p.train(t)
    return res


def measure_library_bar(wl, dev, steps=10, warmup=3, pdrop=0.0):
    """The 'library bar' of SURVEY.md §6 / §8d(ii): the fp32 oracle (plain PyTorch modules = cuBLAS / cuDNN / eager kernels) on the same
    B200, fp32 and bf16 autocast, same batch, forward + loss + backward, CUDA-event timed."""
    from oracle.trainer import masked_action_mse

    out = {}
    try:
        model = build_oracle(wl, pdrop).to(dev)
        model.train()
        b = {k: v.to(dev) for k, v in make_batch(wl, wl["B"], seed=1234).items()}
        if wl["enc"] is not None:
            b["frames_f"] = b["frames"].float() * 2.0 / 255.0 - 1

        def one(autocast):
            model.zero_grad(set_to_none=True)
            with torch.autocast("cuda", dtype=torch.bfloat16, enabled=autocast):
                if wl["enc"] is None:
                    _, ap, _ = model(b["states"], b["actions"], None, b["rtg"], b["timesteps"], b["mask"])
                else:
                    _, ap, _ = model(b["frames_f"], b["audio"], b["actions"], None, b["rtg"], b["timesteps"], b["mask"])
                loss = masked_action_mse(ap.float(), b["actions"], b["mask"])
            loss.backward()

        tokens = wl["B"] * 3 * wl["K"]
        for name, ac in (("eager_fp32", False), ("eager_bf16_autocast", True)):
            for _ in range(warmup):
                one(ac)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize(dev)
            e0.record()
            for _ in range(steps):
                one(ac)
            e1.record()
            torch.cuda.synchronize(dev)
            ms = e0.elapsed_time(e1) / steps
            out[name] = {"ms_per_step": ms, "value": tokens / ms * 1e3, "unit": "tokens/s"}
        out["what"] = ("oracle/ (plain torch.nn modules: cuBLAS / cuDNN / eager elementwise kernels) on this B200, device-resident inputs, "
                       f"fwd + masked MSE + bwd, {steps} steps after {warmup} warm-up, no CUDA graph (launch-bound like any eager loop)")
    except Exception as e:  # noqa: BLE001
        out["error"] = str(e)[:200]
    return out


def measure_e2e_resident(args, wl, ctx, steps):
    """e2e with the trajectories RESIDENT in HBM (mtrl_b200.data.GpuTrajectoryStore): every step samples (trajectory,
    start) pairs on the host with upstream's RNG calls, ships the 2 x B int64 indices, builds the batch with one dtb_get_batch launch
    (raw uint8 frame rows and spectrogram rows included), runs the same captured step and reads the loss back."""
    import numpy as np

    from mtrl_b200.data import GpuTrajectoryStore

    dev, dev_in, K, B = ctx["dev"], ctx["dev_in"], wl["K"], wl["B"]
    rng = np.random.default_rng(99)
    ad = wl["dt"]["act_dim"]
    trajs = []
    for _ in range(128):
        T = int(rng.integers(K, 2 * K + 1))
        t = {"actions": np.eye(ad, dtype=np.float32)[rng.integers(0, ad, T)] if not wl["dt"]["action_tanh"] else rng.uniform(-1, 1, (T, ad)).astype(np.float32),
             "rewards": rng.uniform(-1, 1, T).astype(np.float32), "terminals": np.zeros(T, dtype=np.float32)}
        if wl["enc"] is None:
            t["observations"] = rng.standard_normal((T, wl["dt"]["state_dim"])).astype(np.float32)
        else:
            t["frames"] = rng.integers(0, 256, (T, *wl["enc"]["frame_shape"]), dtype=np.uint8)
            t["audio"] = np.clip(rng.standard_normal((T, *wl["enc"]["audio_shape"])), -1, 1).astype(np.float32)
        trajs.append(t)
    store = GpuTrajectoryStore(trajs, raw_keys=() if wl["enc"] is None else ("frames", "audio"), device=dev)
    get_batch = store.make_get_batch(K, MAX_EP_LEN, 1000.0, with_raw=True)
    graph = ctx.get("graph")
    loss_host = torch.empty(2, dtype=torch.float32).pin_memory()

    def one():
        (s, a, r, d, rtg, ts, m), raws = get_batch(B)
        dev_in["actions"].copy_(a); dev_in["rtg"].copy_(rtg[:, :-1]); dev_in["timesteps"].copy_(ts); dev_in["mask"].copy_(m)
        if wl["enc"] is None:
            dev_in["states"].copy_(s)
        else:
            dev_in["frames"].copy_(raws["frames"]); dev_in["audio"].copy_(raws["audio"])
        if graph is not None:
            graph.replay()
            out = ctx["graph_out"]
        else:
            out = ctx["step"]()
        loss_host.copy_(out[:2], non_blocking=True)
        torch.cuda.current_stream(dev).synchronize()
        return float(loss_host[0])

    for _ in range(3):
        one()
    torch.cuda.synchronize(dev)
    t0 = time.perf_counter()
    for _ in range(steps):
        loss = one()
    torch.cuda.synchronize(dev)
    ms = (time.perf_counter() - t0) / steps * 1e3
    return {"value": B * 3 * K / ms * 1e3, "unit": "tokens/s", "ms_per_step": ms, "h2d_bytes_per_step": 16 * B, "d2h_bytes_per_step": 8, "loss": loss,
            "what": "trajectories (raw uint8 frames / spectrograms / actions / rewards) resident in HBM; per step: host RNG sampling, 2 x B int64 "
                    "indices over PCIe, one dtb_get_batch launch, device-to-device hand-over to the captured step, loss read-back + sync"}


def measure_get_action(wl, ctx, calls=300, warmup=50):
    """BASELINE.json's second metric: get_action latency at B = 1 with a full K-step history, through the public API (wall clock, one
    stream sync per call, p50 / p90 / p99 in microseconds), model.eval().
      * `p50_us` ...: what get_action() does by default (`default_path`): the KV-cached step where the tensor-core decode kernel applies
        (hidden 128: c1 / c2), else the recompute path; both paths are also reported on their own (`recompute`, `kv_cached`).
      * `kv_cached`: get_action(use_cache=True) at the last step of a rollout (t = K): the K/V cache holds the first K - 1 timesteps, the
        call appends one (towers on the newest frame only + ONE decode kernel).  Between timed calls the cache is rewound to K - 1
        timesteps (`_dc["t"] = K - 1`: its rows for the unchanged prefix stay valid), so every timed call does a rollout step's work."""
    model, dev, K = ctx["model"], ctx["dev"], wl["K"]
    model.eval()  # evaluation rollouts run without dropout (upstream evaluate_episode_rtg: model.eval())
    b = make_batch(wl, 1, seed=4321)
    args = []
    if wl["enc"] is None:
        args.append(b["states"][0].to(dev))
    else:
        args += [b["frames"][0].to(dev), b["audio"][0].to(dev)]
    args += [b["actions"][0].to(dev), torch.zeros(K, device=dev), b["rtg"][0].to(dev), b["timesteps"][0].to(dev)]
    dt = ctx["parts"][0]

    def run(use_cache):
        lat, a = [], None
        for i in range(warmup + calls):
            if use_cache and dt._dc is not None and dt._dc["t"] == K:
                dt._dc["t"] = K - 1  # rewind: the next call appends timestep K - 1 again
                step = dt.__dict__.get("_step")
                if step is not None:
                    step.arm(K)  # a real rollout never rewinds: the device-side step count is put back outside the timed call
            torch.cuda.synchronize(dev)
            t0 = time.perf_counter()
            a = model.get_action(*args, use_cache=use_cache)
            torch.cuda.current_stream(dev).synchronize()
            if i >= warmup:
                lat.append((time.perf_counter() - t0) * 1e6)
        lat.sort()
        pick = lambda q: lat[min(len(lat) - 1, int(q * len(lat)))]
        return {"p50_us": pick(0.5), "p90_us": pick(0.9), "p99_us": pick(0.99)}, a

    cached_ok = bool(dt._decode_ready())
    auto_cached = bool(dt._use_cache(None)) and not (wl["enc"] is not None and model.encoder.attn_type)
    rec, a = run(False)
    rec["path"] = "use_cache=False: recompute (pad_left + CUDA graph of the whole B = 1 forward over the K-step window)"
    kvc = None
    if cached_ok:
        kvc, a2 = run(True)
        kvc["max_abs_diff_vs_recompute"] = float((a2 - a).abs().max())
        kvc["path"] = ("use_cache=True: KV-cached append at t = K (pad_left of the newest rows + one CUDA graph: " +
                       ("towers on the newest frame + " if wl["enc"] is not None else "") + "dtb_dt_decode_step, one kernel)")
    # headline = what get_action() does when the caller does not choose: the cached step where the tensor-core decode kernel applies
    out = dict(kvc if (auto_cached and kvc is not None) else rec)
    out["default_path"] = "kv_cached" if (auto_cached and kvc is not None) else "recompute"
    out["recompute"] = rec
    if kvc is not None:
        out["kv_cached"] = kvc
    out.update({"calls": calls, "warmup": warmup, "history": K, "action_dim": int(a.numel()),
                "how": "wall clock around model.get_action(...) + stream sync, inputs resident on the device, model.eval()"})
    return out


def other_workloads(args, skip):
    """Short device-resident runs of the other single-GPU configs of BASELINE.json (each in a fresh subprocess so its workspace is
    released): tokens/s, algorithmic TFLOP/s and the dominant kernel's roofline line."""
    out = {}
    runs = [(n, n, args.dropout) for n in ("c1", "c3", "c4") if n != skip]
    attn_runs = {"c2_attn_casl": "casl", "c2_attn_new": "new"}  # the headline workload with the reference's --attn-type options
    runs += [(k, "c2", args.dropout) for k in attn_runs]
    if args.dropout > 0:
        runs += [(f"{n}_dropout0", n, 0.0) for n in (skip, "c4") if n]  # the same steps with dropout off (parity mode)
    for key, name, pdrop in runs:
        cmd = [sys.executable, os.path.abspath(__file__), "--workload", name, "--steps", "10" if name == "c4" else "30", "--warmup", "3",
               "--no-cpu-baseline", "--no-extras", "--dropout", str(pdrop)]
        if key in attn_runs:
            cmd += ["--attn-type", attn_runs[key]]
        elif key != name:
            cmd += ["--no-get-action"]
        name = key
        try:
            r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
            d = json.loads(r.stdout.strip().splitlines()[-1])
            rf = d.get("roofline") or {}
            out[name] = {"workload": d["config"]["workload"], "dropout": d["config"].get("dropout"), "value": d["value"], "unit": d["unit"], "ms_per_step": d["ms_per_step"],
                         "e2e": d["e2e"]["value"], "algorithmic_tflops_per_s": d["algorithmic_tflops_per_s"], "launches_per_step": d["launches_per_step"],
                         "get_action": d.get("get_action"),
                         "roofline": {k: rf.get(k) for k in ("kernel", "bound", "achieved", "peak", "unit", "frac")},
                         "families": [{k: f.get(k) for k in ("kernel", "ms_per_step", "achieved", "unit", "frac")} for f in rf.get("families", [])[:6]]}
        except Exception as e:  # noqa: BLE001
            out[name] = {"error": str(e)[:200]}
    return out


def next_rows():
    """SURVEY.md §8(f) rows built beside the headline path, timed by their own tools (fresh subprocess each): the modality-attention op
    (N4, agents.py:58-90) at the c2 sample count and the wav -> MFCC-image front-end (N3, Minecraft.py:226-245) with its numpy / Pillow
    oracle timed on the host beside it."""
    out = {}
    root = os.path.dirname(os.path.abspath(__file__))
    for key, tool in (("modality_attention", "modal_attn_bench.py"), ("mfcc_front_end", "mfcc_bench.py")):
        try:
            r = subprocess.run([sys.executable, os.path.join(root, "tools", tool)], capture_output=True, text=True, timeout=300)
            out[key] = json.loads(r.stdout.strip().splitlines()[-1])
        except Exception as e:  # noqa: BLE001
            out[key] = {"error": str(e)[:200]}
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c2", choices=sorted(WORKLOADS))
    ap.add_argument("--batch", type=int, default=0, help="per-GPU batch override (default: the workload's)")
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--flush-l2", action="store_true")
    ap.add_argument("--bucket-mb", type=int, default=2)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-roofline", action="store_true")
    ap.add_argument("--no-get-action", action="store_true", help="skip the get_action latency measurement (keeps ncu launch lists to the training step)")
    ap.add_argument("--dropout", type=float, default=0.1, help="resid/attn/embd dropout of the training step (upstream default 0.1)")
    ap.add_argument("--no-optimizer-extra", action="store_true", help="skip the extra timing of the step + fused clip/AdamW")
    ap.add_argument("--attn-type", default=None, choices=["casl", "new"], help="modality attention of the encoders (reference --attn-type, src/utils.py:115); multimodal workloads only")
    ap.add_argument("--restage-each-step", action="store_true", help="re-stage the bf16 weights inside every timed step (round-1 semantics)")
    ap.add_argument("--families-only", action="store_true", help="print only the per-kernel-family table of one step (used for the serialised table)")
    ap.add_argument("--no-dp-parity", action="store_true", help="N > 1: skip the gradient parity check against the oracle's averaged per-shard gradients")
    ap.add_argument("--no-extras-serial", action="store_true", help="skip the serialised (DTB_NO_SIDE_STREAM=1) family table sub-run")
    ap.add_argument("--no-extras", action="store_true", help="skip the other-workload summaries (c1 / c3 / c4) appended to the default line")
    args = ap.parse_args()
    global ATTN_TYPE
    ATTN_TYPE = args.attn_type
    wl = dict(WORKLOADS[args.workload])
    if args.batch:
        wl["B"] = args.batch
    rank, world, local_rank = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))

    if args.impl == "reference":
        if rank != 0:
            return
        steps, warmup = args.steps, args.warmup  # exactly what was asked for; the bounded sample (batch) is sized so that they fit ~2 minutes
        r = run_cpu(wl, steps, warmup, budget_s=120.0, pdrop=args.dropout)
        line = {"impl": "reference", "metric": "DT train tokens/s (fwd+bwd)", "value": r["value"], "unit": "tokens/s", "n_gpus": args.gpus,
                "steps": steps, "warmup": warmup, "ms_per_step": r["ms_per_step"], "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": {"workload": wl["label"], "name": args.workload, "seq_len": 3 * wl["K"], "parallelism": "cpu", "dropout": args.dropout},
                "cpu_baseline": r, "e2e": {"value": r["value"], "unit": "tokens/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                "gpu_launches": 0,
                "note": "the reference's own DT source is an un-vendored empty submodule (SURVEY.md F1); this arm is the fp32 PyTorch oracle, "
                        "whose towers are numerically the reference's conv_factory, on the host cores"}
        print(json.dumps(line), flush=True)
        return

    import torch.distributed as dist

    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    if args.families_only:
        args.no_optimizer_extra = True
    res, ctx = run_ours(args, wl, rank, world, local_rank)
    if args.families_only:
        print(json.dumps({"ms_per_step": res["ms_per_step"], "roofline": measure_roofline(args, wl, ctx)}), flush=True)
        return
    if not args.no_roofline:
        # every rank runs it (the profiled steps contain the gradient all-reduce at N > 1: a collective on rank 0 alone would hang)
        try:
            roof = measure_roofline(args, wl, ctx)
        except Exception as e:  # noqa: BLE001
            roof = {"error": str(e)}
        if rank == 0:
            res["roofline"] = roof
    if world > 1 and not args.no_dp_parity:
        try:
            dpp = measure_dp_parity(args, wl, ctx, rank, world)
        except Exception as e:  # noqa: BLE001
            dpp = {"error": str(e)[:300]}
        if rank == 0:
            res["dp_parity"] = dpp
    if rank == 0:
        if world == 1 and not args.no_extras:
            try:
                res["e2e"]["resident"] = measure_e2e_resident(args, wl, ctx, args.steps)
            except Exception as e:  # noqa: BLE001
                res["e2e"]["resident"] = {"error": str(e)[:300]}
            res["library_bar"] = measure_library_bar(wl, ctx["dev"], pdrop=args.dropout)
        if world == 1 and not args.no_get_action:
            try:
                res["get_action"] = measure_get_action(wl, ctx)
            except Exception as e:  # noqa: BLE001
                res["get_action"] = {"error": str(e)[:200]}
        if world == 1 and not args.no_cpu_baseline:
            res["cpu_baseline"] = run_cpu(wl, 3, 1, budget_s=20.0, pdrop=args.dropout)
        if world == 1 and not args.no_extras:
            del ctx  # release this workload's device memory before the sub-runs
            torch.cuda.empty_cache()
            res["other_workloads"] = other_workloads(args, args.workload)
            res["next_rows"] = next_rows()
        print(json.dumps(res), flush=True)
    if world > 1:
        # The line is out; nothing below may keep the job alive.  Tear down in order (graphs holding captured NCCL kernels first, then the
        # process group) under a watchdog: if any step of the teardown blocks, exit 0 after 30 s (stderr says where it stopped).
        import gc

        where = ["start"]

        def _bail():
            print(f"[bench] rank {rank}: teardown did not finish within 30 s (stuck after '{where[0]}'); exiting", file=sys.stderr, flush=True)
            os._exit(0)

        wd = threading.Timer(30.0, _bail)
        wd.daemon = True
        wd.start()
        torch.cuda.synchronize()
        where[0] = "synchronize"
        dist.barrier()
        where[0] = "barrier"
        ctx = None  # noqa: F841  (CUDA graphs with captured collectives, reducers, pipelines)
        gc.collect()
        torch.cuda.synchronize()
        where[0] = "graphs released"
        dist.destroy_process_group()
        where[0] = "destroy_process_group"
        wd.cancel()
        sys.stdout.flush()
        sys.stderr.flush()
        os._exit(0)  # skip interpreter-exit destructors of CUDA / NCCL objects: the results are printed, the group is destroyed


if __name__ == "__main__":
    main()
