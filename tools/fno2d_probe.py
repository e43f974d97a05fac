"""One FNOBlock2d forward + backward at the C4 shape (B 16, width 32, 256^2, modes 12): for ncu captures of the fused kernels."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn.functional as F
import resolution_pde_b200 as R
n = int(sys.argv[1]) if len(sys.argv) > 1 else 256
R.set_precision("3xtf32")
torch.manual_seed(0)
blk = R.FNOBlock2d(32, 32, 12, 12, F.gelu).cuda()
x = torch.randn(16, 32, n, n, device="cuda", requires_grad=True)
for _ in range(3):
    y = blk(x)
    y.sum().backward()
torch.cuda.synchronize()
print("ok", float(y.mean()))
