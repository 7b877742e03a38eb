"""Developer probe: run the attention forward/backward a few times at a c4-like shape (for ncu)."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from mtrl_b200 import _lib  # noqa: E402

dev = torch.device("cuda:0")
B, K, nh, d = int(sys.argv[1]) if len(sys.argv) > 1 else 64, 200, 12, 64
T, H = 3 * K, nh * d
qkv = torch.randn(B, T, 3 * H, device=dev).to(torch.bfloat16)
out, lse = torch.empty(B, T, H, device=dev, dtype=torch.bfloat16), torch.empty(B, nh, T, device=dev)
mask = torch.ones(B, K, device=dev)
dout, dqkv, delta = torch.randn(B, T, H, device=dev).to(torch.bfloat16), torch.empty(B, T, 3 * H, device=dev, dtype=torch.bfloat16), torch.empty(B, nh, T, device=dev)
for _ in range(3):
    _lib.attention_fwd(qkv, mask, nh, out, lse)
    if "bwd" in sys.argv:
        _lib.attention_bwd(qkv, mask, out, dout, lse, nh, dqkv, delta)
torch.cuda.synchronize()
print("ok")
