"""Times the GPU audio front-end (wav -> MFCC image) against the numpy / Pillow oracle on the host.
Workload: one clip of the size of the reference's assets/noise.wav (768 764 int16 samples at 44.1 kHz -> 1 742 frames of a 2 048-point
FFT) and a batch of 64 one-second 24 kHz clips (per-step audio of 64 environments).  Usage: python tools/mfcc_bench.py"""
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from mtrl_b200 import audio  # noqa: E402
from oracle import mfcc as O  # noqa: E402


def gpu_time(fn, iters=50, warmup=5):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def main():
    rng = np.random.default_rng(0)
    out = {}
    x = rng.integers(-20000, 20000, 768764).astype(np.int16)
    xd = torch.from_numpy(x).cuda()
    ms = gpu_time(lambda: audio.wav_to_mfcc(44100, xd))
    t0 = time.perf_counter(); ref = O.wav_to_mfcc(44100, x); cpu_ms = (time.perf_counter() - t0) * 1e3
    got = audio.wav_to_mfcc(44100, xd).cpu().numpy()
    out["noise_wav_sized_clip"] = {"samples": int(x.size), "frames": int(O.frame_geometry(x.size, 44100)[2]), "gpu_ms": ms, "cpu_oracle_ms": cpu_ms,
                                   "max_abs_err": float(np.abs(got - ref).max()), "frames_per_s": O.frame_geometry(x.size, 44100)[2] / ms * 1e3}
    clips = rng.integers(-20000, 20000, (64, 24000)).astype(np.int16)
    cd = torch.from_numpy(clips).cuda()
    ms = gpu_time(lambda: audio.wav_batch_to_mfcc(24000, cd))
    t0 = time.perf_counter()
    for c in clips[:8]:
        O.wav_to_mfcc(24000, c)
    cpu_ms = (time.perf_counter() - t0) * 1e3 * 8
    out["batch_64_x_1s_24k"] = {"gpu_ms": ms, "cpu_oracle_ms_extrapolated_from_8": cpu_ms, "clips_per_s": 64 / ms * 1e3}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
