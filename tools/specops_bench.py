"""Spectral data-preparation operators: device (truncated-DFT contractions, 3xTF32 engine) against the CPU restatement of
the reference functions (torch.fft on the host cores).  One JSON line per case; `hbm_frac` = (input + output bytes) /
time / measured copy bandwidth."""
import json, os, sys, time, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
import resolution_pde_b200 as R
import resolution_pde_b200.spectral_ops as so
from oracle import spectral_ops_port as port

R.set_precision("3xtf32")
pk = bench.peaks()
cases = [
    ("lowpass_filter_1d", (64, 20, 1, 1024), lambda f, x: f.lowpass_filter_1d(x, 0.25)),
    ("lowpass_filter_2d", (16, 10, 1, 256, 256), lambda f, x: f.lowpass_filter_2d(x, 0.125)),
    ("resize_1d 2048->256", (64, 20, 2048), lambda f, x: f.resize_1d(x, 256)),
    ("resize 256^2->64^2", (16, 10, 256, 256), lambda f, x: f.resize(x, (64, 64))),
    ("resize 64^2->256^2", (16, 10, 64, 64), lambda f, x: f.resize(x, (256, 256))),
]
for name, shape, call in cases:
    torch.manual_seed(0)
    xh = torch.randn(*shape)
    xd = xh.cuda()
    for _ in range(3): y = call(so, xd)
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); s.record()
    for _ in range(10): y = call(so, xd)
    e.record(); torch.cuda.synchronize()
    gms = s.elapsed_time(e) / 10
    call(port, xh)
    t0 = time.perf_counter(); n = 0
    while time.perf_counter() - t0 < 2.0:
        ref = call(port, xh); n += 1
    cms = (time.perf_counter() - t0) / n * 1e3
    err = ((y.double().cpu() - ref.double()).norm() / ref.double().norm()).item()
    byts = (xd.numel() + y.numel()) * 4
    print(json.dumps({"op": name, "shape": list(shape), "gpu_ms": round(gms, 4), "cpu_ms": round(cms, 3), "cpu_threads": torch.get_num_threads(),
                      "speedup": round(cms / gms, 1), "rel_err_vs_cpu_fp32": float(f"{err:.2e}"),
                      "hbm_frac": round(byts / (gms * 1e-3) / 1e9 / pk["hbm"], 3)}), flush=True)

# frequency-resolved error metrics: Parseval on the device against the reference's mode-by-mode loop
import resolution_pde_b200.frequency_error as fe
for name, shape, call in [
    ("decompose_error_by_frequency_1d (513 modes)", (64, 1, 1024), lambda f, a, b: f.decompose_error_by_frequency_1d(a, b, None)),
    ("decompose_error_by_frequency_2d (64 bins)", (16, 1, 256, 256), lambda f, a, b: f.decompose_error_by_frequency_2d(a, b, 64)),
]:
    torch.manual_seed(1)
    ah, bh = torch.randn(*shape), torch.randn(*shape)
    ad, bd = ah.cuda(), bh.cuda()
    for _ in range(3): got = call(fe, ad, bd)
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(10): got = call(fe, ad, bd)          # returns numpy: the device sync is inside
    gms = (time.perf_counter() - t0) / 10 * 1e3
    t0 = time.perf_counter(); ref = call(port, ah, bh); cms = (time.perf_counter() - t0) * 1e3
    import numpy as np
    err = float(np.linalg.norm(got[0] - ref[0]) / np.linalg.norm(ref[0]))
    print(json.dumps({"op": name, "shape": list(shape), "gpu_ms_wall_incl_sync": round(gms, 4), "cpu_ms": round(cms, 2),
                      "cpu_threads": torch.get_num_threads(), "speedup": round(cms / gms, 1), "rel_err_vs_cpu_fp32": float(f"{err:.2e}")}), flush=True)
