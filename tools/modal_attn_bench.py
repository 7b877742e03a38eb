"""Times the modality-attention op (forward + backward through the C ABI) at the c2 sample count (N = 64 x 30 = 1 920) and at 16x that.
Usage: python tools/modal_attn_bench.py [--once]   (--once: one forward + backward per case, for an ncu launch list)"""
import ctypes as C
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from mtrl_b200 import _lib  # noqa: E402


def case(kind, F, fusion, N, once):
    lib = _lib.load()
    dev = torch.device("cuda")
    cfg = _lib.ModalAttnConfig(0 if kind == "casl" else 1, F, 0 if fusion == "concat" else 1)
    total = int(lib.dtb_modal_attn_param_layout(C.byref(cfg), None, None))
    g = torch.Generator(device="cpu").manual_seed(0)
    flat = (torch.randn(total, generator=g) * 0.03).to(dev)
    feat = torch.relu(torch.randn(N, 2 * F, generator=g)).to(torch.bfloat16).to(dev)
    hid = torch.randn(N, 128, generator=g).to(dev)
    S = 2 * F if fusion == "concat" else F
    dS = torch.randn(N, S, generator=g).to(torch.bfloat16).to(dev)
    ws = torch.empty(int(lib.dtb_modal_attn_workspace_bytes(C.byref(cfg), N)), dtype=torch.uint8, device=dev)
    states, w = torch.empty(N, S, dtype=torch.bfloat16, device=dev), torch.empty(N, 2, device=dev)
    dfeat, grads = torch.empty(N, 2 * F, dtype=torch.bfloat16, device=dev), torch.empty_like(flat)

    def fwd():
        _lib.check(lib.dtb_modal_attn_forward(C.byref(cfg), _lib.stream_ptr(), _lib.ptr(flat), N, _lib.ptr(feat), _lib.ptr(hid), _lib.ptr(ws),
                                              ws.numel(), _lib.ptr(states), _lib.ptr(w)))

    def bwd():
        _lib.check(lib.dtb_modal_attn_backward(C.byref(cfg), _lib.stream_ptr(), _lib.ptr(flat), N, _lib.ptr(feat), 1, _lib.ptr(dS), _lib.ptr(ws),
                                               ws.numel(), _lib.ptr(dfeat), _lib.ptr(grads)))

    if once:
        fwd(); bwd(); torch.cuda.synchronize()
        return {}
    for _ in range(3):
        fwd(); bwd()
    torch.cuda.synchronize()
    gf, gb = torch.cuda.CUDAGraph(), torch.cuda.CUDAGraph()
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        with torch.cuda.graph(gf):
            fwd()
        with torch.cuda.graph(gb):
            bwd()
    out = {}
    for name, gr in (("fwd_us", gf), ("bwd_us", gb)):
        for _ in range(5):
            gr.replay()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(50):
            gr.replay()
        e1.record()
        torch.cuda.synchronize()
        out[name] = e0.elapsed_time(e1) / 50 * 1e3
    # algorithmic bytes: features read + states written (fwd); features + dS read, d_feat written (bwd); bf16
    out["fwd_algorithmic_GBps"] = (N * 2 * F * 2 + N * S * 2 + N * 128 * 4) / out["fwd_us"] * 1e-3
    out["bwd_algorithmic_GBps"] = (N * 2 * F * 2 * 2 + N * S * 2) / out["bwd_us"] * 1e-3
    return out


if __name__ == "__main__":
    once = "--once" in sys.argv
    res = {}
    for kind in ("casl", "new"):
        for N in (1920, 30720):
            res[f"{kind}_F512_concat_N{N}"] = case(kind, 512, "concat", N, once)
    print(json.dumps(res))
