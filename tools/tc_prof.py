"""One launch of a Linear fwd / dgrad case on a -DRPDE_TC_PROFILE build (per-role cycle counters of CTA 0 on stdout).
usage: tc_prof.py <prec> <fwd|dgrad> <in> <out> <flags>   flags: fwd -> act (0, 1, 0x101 with Z); dgrad -> 0 plain, 3 saved, 7 saved + dbias"""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import resolution_pde_b200 as R
R.set_precision(sys.argv[1])
R.ops.matmul(torch.randn(256, 256, device="cuda"), torch.randn(256, 256, device="cuda"))
torch.cuda.synchronize()
print("---- case", sys.argv[1:], flush=True)
lib = R._lib.load(); st = torch.cuda.current_stream().cuda_stream
T = 1048576
kind, i, o, fl = sys.argv[2], int(sys.argv[3]), int(sys.argv[4]), int(sys.argv[5], 0)
if kind == "fwd":
    x = torch.randn(T, i, device="cuda"); w = torch.randn(o, i, device="cuda"); b = torch.randn(o, device="cuda")
    y = torch.empty(T, o, device="cuda"); z = torch.empty(T, o, device="cuda")
    lib.rpde_linear_fwd(x.data_ptr(), w.data_ptr(), b.data_ptr(), y.data_ptr(), z.data_ptr() if fl & 0x100 else None, T, i, o, fl, 0, st)
else:
    dy = torch.randn(T, o, device="cuda"); w = torch.randn(o, i, device="cuda"); sv = torch.rand(T, i, device="cuda")
    dx = torch.empty(T, i, device="cuda"); db = torch.zeros(i, device="cuda")
    lib.rpde_linear_bwd_input(dy.data_ptr(), w.data_ptr(), sv.data_ptr() if fl & 3 else None, fl & 3, dx.data_ptr(), 0,
                              db.data_ptr() if fl & 4 else None, T, i, o, 0, st)
torch.cuda.synchronize()
