"""Runs the fused projection MLP (forward + backward) a few times on one shape, for ncu / timing:
python tools/mlp_prof.py [B I Hd O S]   (default: the C4-256 projection, 16 x 32 -> 128 -> 1 at 256^2)"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import resolution_pde_b200 as R
from resolution_pde_b200 import ops

B, I, Hd, O, S = (int(v) for v in sys.argv[1:6]) if len(sys.argv) >= 6 else (16, 32, 128, 1, 256 * 256)
dev = "cuda:0"
torch.manual_seed(0)
x = torch.randn(B, I, S, device=dev, requires_grad=True)
w1 = (torch.randn(Hd, I, 1, device=dev) / I ** 0.5).requires_grad_(True)
b1 = torch.randn(Hd, device=dev, requires_grad=True)
w2 = (torch.randn(O, Hd, 1, device=dev) / Hd ** 0.5).requires_grad_(True)
b2 = torch.randn(O, device=dev, requires_grad=True)
gy = torch.randn(B, O, S, device=dev)
for it in range(5):
    e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
    e[0].record()
    y = ops.mlp_gelu(x, w1, b1, w2, b2)
    e[1].record()
    y.backward(gy)
    e[2].record()
    torch.cuda.synchronize()
    print(f"fwd {e[0].elapsed_time(e[1]) * 1e3:.1f} us  bwd {e[1].elapsed_time(e[2]) * 1e3:.1f} us")
