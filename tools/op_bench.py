"""Developer probe: per-kernel durations of the library ops at a config's shapes (back-to-back launches inside one CUDA graph)."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from mtrl_b200 import _lib  # noqa: E402

dev = torch.device("cuda:0")


def timeit(name, fn, reps=50, flops=0, bytes_=0):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(reps):
            fn()
    g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    g.replay()
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) / reps * 1e3
    extra = ""
    if flops:
        extra += f"  {flops / us / 1e6:8.1f} TFLOP/s"
    if bytes_:
        extra += f"  {bytes_ / us / 1e3:8.1f} GB/s"
    print(f"{name:44s} {us:8.2f} us{extra}", flush=True)


def bf(*shape):
    return torch.randn(*shape, device=dev).to(torch.bfloat16)


def gemm_suite(M, H, I):
    x, w_qkv, w_fc = bf(M, H), bf(H, 3 * H), bf(H, I)
    qkv, fc, xo = torch.empty(M, 3 * H, device=dev, dtype=torch.bfloat16), torch.empty(M, I, device=dev, dtype=torch.bfloat16), torch.empty(M, H, device=dev)
    bias3, biasI, biasH, res = torch.zeros(3 * H, device=dev), torch.zeros(I, device=dev), torch.zeros(H, device=dev), torch.randn(M, H, device=dev)
    timeit(f"gemm qkv  M={M} N={3*H} K={H}", lambda: _lib.gemm_bf16(x, w_qkv, M, 3 * H, H, b_mn_major=True, bias=bias3, out_bf16=qkv), flops=2 * M * 3 * H * H)
    timeit(f"gemm fc   M={M} N={I} K={H} relu", lambda: _lib.gemm_bf16(x, w_fc, M, I, H, b_mn_major=True, bias=biasI, act=1, out_bf16=fc), flops=2 * M * I * H)
    w_pr = bf(I, H)
    timeit(f"gemm proj M={M} N={H} K={I} +res f32", lambda: _lib.gemm_bf16(fc, w_pr, M, H, I, b_mn_major=True, bias=biasH, residual=res, out_f32=xo), flops=2 * M * I * H)
    dw = torch.empty(H, I, device=dev)
    part = torch.empty(32, H, I, device=dev)
    for split in (1, 4, 8):
        timeit(f"wgrad dW1 [{H}x{I}] Kred={M} split={split}", lambda: _lib.gemm_bf16(x, fc, H, I, M, a_mn_major=True, b_mn_major=True, out_f32=dw, partial=part, split_k=split), flops=2 * M * I * H)
    y = torch.empty(M, H, device=dev, dtype=torch.bfloat16)
    xf, g, b = torch.randn(M, H, device=dev), torch.ones(H, device=dev), torch.zeros(H, device=dev)
    mean, rstd = torch.empty(M, device=dev), torch.empty(M, device=dev)
    timeit(f"ln_fwd M={M} H={H}", lambda: _lib.layernorm_fwd(xf, g, b, 1e-5, y, mean, rstd), bytes_=M * H * 6)
    dx, dxb, dg, db = torch.empty(M, H, device=dev), torch.empty(M, H, device=dev, dtype=torch.bfloat16), torch.empty(H, device=dev), torch.empty(H, device=dev)
    scratch = torch.empty(_lib.load().dtb_layernorm_bwd_scratch_floats(H), device=dev)
    timeit(f"ln_bwd M={M} H={H}", lambda: _lib.layernorm_bwd(y, xf, mean, rstd, g, xf, dx, dxb, dg, db, scratch), bytes_=M * H * 16)
    out = torch.empty(I, device=dev)
    timeit(f"colsum M={M} N={I}", lambda: _lib.colsum_bf16(fc, out), bytes_=M * I * 2)


def attn_suite(B, K, nh, d):
    T, H = 3 * K, nh * d
    qkv, out, lse = bf(B, T, 3 * H), torch.empty(B, T, H, device=dev, dtype=torch.bfloat16), torch.empty(B, nh, T, device=dev)
    mask = torch.ones(B, K, device=dev)
    fl = 4 * B * nh * T * T * d / 2
    timeit(f"attn_fwd B={B} T={T} nh={nh} d={d}", lambda: _lib.attention_fwd(qkv, mask, nh, out, lse), flops=fl)
    dout, dqkv, delta = bf(B, T, H), torch.empty(B, T, 3 * H, device=dev, dtype=torch.bfloat16), torch.empty(B, nh, T, device=dev)
    timeit(f"attn_bwd B={B} T={T} nh={nh} d={d}", lambda: _lib.attention_bwd(qkv, mask, out, dout, lse, nh, dqkv, delta), flops=2.5 * fl)
    st = torch.tensor([1, 2, 3], dtype=torch.int32, device=dev)  # dropout generator state {seed_lo, seed_hi, step}
    timeit(f"attn_fwd p=0.1 B={B} T={T} nh={nh} d={d}", lambda: _lib.attention_fwd(qkv, mask, nh, out, lse, st, 1, 0.1), flops=fl)
    timeit(f"attn_bwd p=0.1 B={B} T={T} nh={nh} d={d}", lambda: _lib.attention_bwd(qkv, mask, out, dout, lse, nh, dqkv, delta, st, 1, 0.1),
           flops=2.5 * fl)


if __name__ == "__main__":
    which = sys.argv[1] if len(sys.argv) > 1 else "c2"
    if which == "c2":
        gemm_suite(5760, 128, 512)
        attn_suite(64, 30, 1, 128)
    elif which == "c4":
        gemm_suite(int(sys.argv[2]) * 600 if len(sys.argv) > 2 else 19200, 768, 3072)
        attn_suite(int(sys.argv[2]) if len(sys.argv) > 2 else 32, 200, 12, 64)
