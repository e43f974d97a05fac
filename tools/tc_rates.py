#!/usr/bin/env python
"""Times the contraction engine on compute-bound and FFNO-shaped products (CUDA events, B200)."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import resolution_pde_b200 as R  # noqa: E402

CASES = [(18944, 256, 512, 0, 1), (18944, 256, 512, 0, 0), (18944, 256, 512, 1, 0), (18944, 128, 2048, 0, 1),
         (1048576, 256, 256, 0, 1), (1048576, 256, 64, 0, 1), (1048576, 64, 256, 0, 1), (1048576, 256, 256, 0, 0)]
for prec in ("tf32", "3xtf32"):
    R.set_precision(prec)
    for (M, N, K, ta, tb) in CASES:
        a = torch.randn((K, M) if ta else (M, K), device="cuda")
        b = torch.randn((N, K) if tb else (K, N), device="cuda")
        for _ in range(2):
            R.ops.matmul(a, b, bool(ta), bool(tb))
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        reps = 5
        s.record()
        for _ in range(reps):
            R.ops.matmul(a, b, bool(ta), bool(tb))
        e.record()
        torch.cuda.synchronize()
        ms = s.elapsed_time(e) / reps
        print(json.dumps({"prec": prec, "M": M, "N": N, "K": K, "ta": ta, "tb": tb, "ms": round(ms, 4),
                          "tflops": round(2.0 * M * N * K / ms / 1e9, 1),
                          "gbs": round(4.0 * (M * K + N * K + M * N) / ms / 1e6, 1)}), flush=True)
        del a, b

# weight gradient of the F-FNO hidden Linear (split-K + chunked accumulation), through the public op
for prec in ("tf32", "3xtf32"):
    R.set_precision(prec)
    for (T, i, o) in [(1048576, 256, 256), (1048576, 64, 256), (1048576, 256, 64)]:
        x = torch.randn(T, i, device="cuda")
        g = torch.randn(T, o, device="cuda")
        dw = torch.empty(o, i, device="cuda")
        db = torch.empty(o, device="cuda")
        lib = R._lib.load()
        st = torch.cuda.current_stream().cuda_stream
        for _ in range(2):
            lib.rpde_linear_bwd_params(g.data_ptr(), x.data_ptr(), dw.data_ptr(), None, T, i, o, 0, st)
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        for _ in range(5):
            lib.rpde_linear_bwd_params(g.data_ptr(), x.data_ptr(), dw.data_ptr(), None, T, i, o, 0, st)
        e.record()
        torch.cuda.synchronize()
        ms = s.elapsed_time(e) / 5
        ref = g[:65536].double().T @ x[:65536].double()
        lib.rpde_linear_bwd_params(g.data_ptr(), x.data_ptr(), dw.data_ptr(), None, 65536, i, o, 0, st)
        err = ((dw.double() - ref).norm() / ref.norm()).item()
        print(json.dumps({"prec": prec, "wgrad": [T, i, o], "ms": round(ms, 4), "tflops": round(2.0 * T * i * o / ms / 1e9, 1),
                          "gbs": round(4.0 * T * (i + o) / ms / 1e6, 1), "rel_err_T65536": err}), flush=True)

# FeedForward-shaped Linear calls with their fused epilogues (bias + GELU + pre-activation store; dgrad * gelu'(z))
def _time(fn, reps=5):
    for _ in range(2):
        fn()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(reps):
        fn()
    e.record()
    torch.cuda.synchronize()
    return s.elapsed_time(e) / reps


for prec in ("tf32", "3xtf32"):
    R.set_precision(prec)
    lib = R._lib.load()
    st = torch.cuda.current_stream().cuda_stream
    T = 1048576
    for (i, o) in [(64, 256), (256, 256), (256, 64)]:
        x = torch.randn(T, i, device="cuda")
        w = torch.randn(o, i, device="cuda") / i ** 0.5
        b = torch.randn(o, device="cuda")
        y = torch.empty(T, o, device="cuda")
        z = torch.empty(T, o, device="cuda")
        ms_plain = _time(lambda: lib.rpde_linear_fwd(x.data_ptr(), w.data_ptr(), b.data_ptr(), y.data_ptr(), None, T, i, o, 0, 0, st))
        ms_gelu = _time(lambda: lib.rpde_linear_fwd(x.data_ptr(), w.data_ptr(), b.data_ptr(), y.data_ptr(), z.data_ptr(), T, i, o, 1, 0, st))
        ref = torch.nn.functional.gelu(x[:4096].double() @ w.double().T + b.double())
        err = ((y[:4096].double() - ref).norm() / ref.norm()).item()
        dy = torch.randn(T, o, device="cuda")
        zp = torch.randn(T, i, device="cuda")
        dx = torch.empty(T, i, device="cuda")
        ms_dgrad = _time(lambda: lib.rpde_linear_bwd_input(dy.data_ptr(), w.data_ptr(), None, 0, dx.data_ptr(), 0, None, T, i, o, 0, st))
        ms_dact = _time(lambda: lib.rpde_linear_bwd_input(dy.data_ptr(), w.data_ptr(), zp.data_ptr(), 1, dx.data_ptr(), 0, None, T, i, o, 0, st))
        zd = zp[:4096].double()
        dg = 0.5 * (1 + torch.erf(zd / 2 ** 0.5)) + zd * torch.exp(-0.5 * zd * zd) / (2 * torch.pi) ** 0.5
        ref = (dy[:4096].double() @ w.double()) * dg
        err2 = ((dx[:4096].double() - ref).norm() / ref.norm()).item()
        print(json.dumps({"prec": prec, "linear": [T, i, o], "fwd_bias_ms": round(ms_plain, 4), "fwd_bias_gelu_z_ms": round(ms_gelu, 4),
                          "fwd_err": err, "dgrad_ms": round(ms_dgrad, 4), "dgrad_dact_ms": round(ms_dact, 4), "dgrad_err": err2}), flush=True)
        del x, y, z, dy, zp, dx
