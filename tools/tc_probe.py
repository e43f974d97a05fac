#!/usr/bin/env python
"""Bring-up probe of the tcgen05/TMA contraction engine (run on the B200 box): every case in its own process, so a
trap (bounded mbarrier wait) in one case cannot poison the CUDA context of the others.  Prints one line per case."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

CASES = [
    # M, N, K, trans_a, trans_b
    (256, 256, 256, 0, 1), (256, 256, 256, 0, 0), (256, 256, 256, 1, 1), (256, 256, 256, 1, 0),
    (128, 64, 32, 0, 1), (128, 64, 32, 1, 0), (128, 128, 64, 0, 0),
    (200, 136, 100, 0, 1), (200, 136, 100, 1, 0), (1000, 72, 260, 0, 0),
    (8192, 256, 256, 0, 1), (8192, 256, 64, 0, 1), (8192, 64, 256, 0, 1), (64, 256, 8192, 1, 0), (256, 256, 65536, 1, 0),
]


def one(case, prec):
    import torch
    import resolution_pde_b200 as R
    M, N, K, ta, tb = case
    torch.manual_seed(1)
    a = torch.randn((K, M) if ta else (M, K), device="cuda")
    b = torch.randn((N, K) if tb else (K, N), device="cuda")
    R.set_precision(prec)
    lib = R._lib.load()
    c0 = lib.rpde_tc_launch_count()
    out = R.ops.matmul(a, b, bool(ta), bool(tb))
    torch.cuda.synchronize()
    used = lib.rpde_tc_launch_count() - c0
    ad, bd = a.double(), b.double()
    ref = (ad.t() if ta else ad) @ (bd.t() if tb else bd)
    err = ((out.double() - ref).norm() / ref.norm()).item()
    # what a plain one-pass TF32 product with truncated / rounded operands would give
    def trunc(t):
        return (t.view(torch.int32) & -8192).view(torch.float32)
    at, bt = trunc(a).double(), trunc(b).double()
    ref_t = (at.t() if ta else at) @ (bt.t() if tb else bt)
    err_t = ((out.double() - ref_t).norm() / ref.norm()).item()
    return {"case": case, "prec": prec, "tc": int(used), "rel_err": err, "rel_err_vs_truncated_tf32": err_t}


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "--one":
        print(json.dumps(one(tuple(json.loads(sys.argv[2])), sys.argv[3])))
        sys.exit(0)
    for dbg in ("1", "2"):
        for case in [(256, 256, 256, 0, 0), (256, 256, 256, 1, 1)]:
            env = dict(os.environ, RPDE_TC_DBG=dbg)
            r = subprocess.run([sys.executable, os.path.abspath(__file__), "--one", json.dumps(case), "tf32"], capture_output=True,
                               text=True, timeout=120, env=env)
            line = [l for l in r.stdout.splitlines() if l.startswith("{")]
            print("dbg", dbg, line[-1] if line else r.stderr[-300:], flush=True)
    for prec in ("tf32", "3xtf32"):
        for case in CASES:
            r = subprocess.run([sys.executable, os.path.abspath(__file__), "--one", json.dumps(case), prec], capture_output=True,
                               text=True, timeout=120)
            line = [l for l in r.stdout.splitlines() if l.startswith("{")]
            print(line[-1] if line else json.dumps({"case": case, "prec": prec, "rc": r.returncode, "err": r.stderr[-400:]}), flush=True)
