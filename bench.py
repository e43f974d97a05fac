#!/usr/bin/env python
"""Benchmark of the spectral-convolution hot path: fwd + RelativeL2 + bwd samples/s of the FNO / F-FNO models.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload NAME|all] [--precision fp32|3xtf32|tf32|bf16]
                  [--impl reference] [--no-variants] [--no-kernels] [--no-cpu-baseline] [--no-e2e] [--no-graph]

Prints ONE JSON line (rank 0).
  value        whole-job samples/s with the inputs resident in HBM (CUDA-event timing, max over ranks)
  e2e          the same step driven from pinned host buffers (H2D of x and y, D2H of the loss every step)
  roofline     the dominant kernel of the step, timed IN SITU (torch.profiler / CUPTI over one eager step joined with the
               library's launch trace for the shapes): algorithmic FLOPs (or bytes) of its launches / their summed duration
  roofline_step  the whole step against the measured peaks (algorithmic FLOPs of SURVEY 8d / step time, and the HBM floor)
  kernels      the in-situ table (name, shape, launches, average duration, share); the shares sum to ~1
  variants     the same workload under the other precision variants (each with its stated, tested model-level tolerance)
  cpu_baseline the oracle port of the reference (oracle/reference_port.py; the reference is Python and cannot travel to the
               GPU box) on the host cores at the SAME batch, with a check of this run's loss against the oracle's loss
`--impl reference` times that CPU arm alone with the same config / metric.  Multi-GPU (torchrun): batch-sharded weak
scaling, one flat NCCL all-reduce of the gradients inside the timed step.
"""
from __future__ import annotations

import argparse
import collections
import ctypes
import gc
import json
import os
import re
import statistics
import subprocess
import sys
import tempfile
import time

import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

# (kind, ctor kwargs, oracle hyper, input shape (per-GPU batch first), algorithmic FLOPs/step at that batch, HBM floor bytes)
WORKLOADS = {
    # BASELINE.json configs[2] / north_star target: FFNO2D Navier-Stokes 256x256, batch 16 per GPU (conf/model/ffno_2d/ffno_2d.yaml, dropout -> 0)
    "ffno2d_ns256": dict(kind="FFNO2D", batch=16, shape=(1, 256, 256), bound="tensor", flops=3.51e12, bytes=5.39e9,
                         ctor=dict(in_channels=1, out_channels=1, width=64, n_layers=4, n_modes=64, factor=4, ff_weight_norm=True,
                                   n_ff_layers=3, layer_norm=True, dropout=0.0),
                         hyper=dict(n_layers=4, n_modes=64, n_ff_layers=3, layer_norm=True, use_grid=True)),
    # configs[0]: FFNO1D Kuramoto-Sivashinsky 512, batch 32 (conf/model/ffno_1d/ffno_1d.yaml, dropout -> 0)
    "ffno1d_ks512": dict(kind="FFNO1D", batch=32, shape=(1, 512), bound="tensor", flops=170.8e9, bytes=168e6,
                         ctor=dict(in_channels=1, out_channels=1, width=128, n_layers=4, n_modes=64, factor=4, ff_weight_norm=True,
                                   n_ff_layers=3, layer_norm=True, dropout=0.0, activation="gelu"),
                         hyper=dict(n_layers=4, n_modes=64, n_ff_layers=3, layer_norm=True, activation="gelu")),
    # configs[1]: FNO1d Burgers 1024, modes 16, width 64
    "fno1d_burgers1024": dict(kind="FNO1d", batch=16, shape=(1, 1024), bound="hbm", flops=4.97e9, bytes=84.3e6,
                              ctor=dict(in_channels=1, out_channels=1, modes=16, width=64),
                              hyper=dict(activation="relu", n_blocks=4)),
    # configs[3]: FNO2d Navier-Stokes, modes 12, width 32 at 64/128/256
    "fno2d_ns64": dict(kind="FNO2d", batch=16, shape=(1, 64, 64), bound="hbm", flops=7.99e9, bytes=169e6,
                       ctor=dict(in_channels=1, out_channels=1, modes1=12, modes2=12, width=32), hyper=dict(activation="gelu", n_blocks=4)),
    "fno2d_ns128": dict(kind="FNO2d", batch=16, shape=(1, 128, 128), bound="hbm", flops=27.0e9, bytes=677e6,
                        ctor=dict(in_channels=1, out_channels=1, modes1=12, modes2=12, width=32), hyper=dict(activation="gelu", n_blocks=4)),
    "fno2d_ns256": dict(kind="FNO2d", batch=16, shape=(1, 256, 256), bound="hbm", flops=99.3e9, bytes=2.71e9,
                        ctor=dict(in_channels=1, out_channels=1, modes1=12, modes2=12, width=32), hyper=dict(activation="gelu", n_blocks=4)),
}
# BASELINE.json configs[3] as it is meant: ONE run over a true multi-resolution mix, batches grouped by resolution the way
# ResolutionGroupedDataLoader does (train/mres_training.py:75-131), ragged tail included.  Per GPU and "step" (= one pass):
# 48 fields at 64^2, 32 at 128^2, 24 at 256^2, batch 16 -> batches of 16,16,16 / 16,16 / 16,8.
MIX = dict(kind="FNO2d", batch=16, groups=((64, 48), (128, 32), (256, 24)),
           ctor=dict(in_channels=1, out_channels=1, modes1=12, modes2=12, width=32), hyper=dict(activation="gelu", n_blocks=4),
           bytes_per_sample={64: 169e6 / 16, 128: 677e6 / 16, 256: 2.71e9 / 16}, flops_per_sample={64: 7.99e9 / 16, 128: 27.0e9 / 16, 256: 99.3e9 / 16})
DEFAULT = "ffno2d_ns256"
# stated model-level tolerances (relative L2 per tensor vs the fp64 oracle: outputs, gradients), tests/test_gpu_baseline_configs.py
TOLERANCE = {"fp32": (1e-5, 1e-5), "3xtf32": (1e-5, 1e-5), "tf32": (5e-3, 1.5e-2), "bf16": (1e-2, 4e-2)}
# tensor peak of a variant relative to the measured dense BF16 peak: kind::tf32 UMMA issues at half the kind::f16 rate,
# 3xTF32 issues three UMMAs per credited product
VARIANT_DIV = {"bf16": 1.0, "tf32": 2.0, "3xtf32": 6.0}


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return dict(hbm=d["hbm_gbs"], tensor=d["bf16_tflops_sustained"], tensor_burst=d["bf16_tflops"], src="measured (MEASURED_PEAKS.json)")
    return dict(hbm=6650.0, tensor=1400.0, tensor_burst=1590.0, src="fallback (B200_PROFILING.md)")


def synth_fields(batch, shape, seed):
    """Band-limited Gaussian random fields, standardised: NS-like (4 pi^2 |k|^2 + 49)^-1.25 in 2-D,
    exp(-(k/32)^2) (KS) / k^-1 (Burgers) in 1-D.  Built on the CPU with a seeded generator."""
    g = torch.Generator().manual_seed(seed)
    out = []
    for _ in range(2):
        if len(shape) == 3:
            _, m, n = shape
            kx = torch.fft.fftfreq(m, 1.0 / m).reshape(m, 1)
            ky = torch.fft.fftfreq(n, 1.0 / n).reshape(1, n)
            amp = (4 * torch.pi ** 2 * (kx ** 2 + ky ** 2) + 49.0) ** (-1.25)
            noise = torch.complex(torch.randn(batch, 1, m, n, generator=g), torch.randn(batch, 1, m, n, generator=g))
            f = torch.fft.ifft2(noise * amp).real
        else:
            _, n = shape
            k = torch.arange(n // 2 + 1, dtype=torch.float32)
            amp = torch.exp(-(k / 32.0) ** 2) if n <= 512 else torch.where((k >= 1) & (k <= n // 3), 1.0 / k.clamp(min=1), torch.zeros(()))
            noise = torch.complex(torch.randn(batch, 1, n // 2 + 1, generator=g), torch.randn(batch, 1, n // 2 + 1, generator=g))
            f = torch.fft.irfft(noise * amp, n=n)
        f = (f - f.mean()) / (f.std() + 1e-8)
        out.append(f.float().contiguous())
    return out[0], out[1]


class ClockSampler:
    QUERY = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.gpu, self.proc, self.path = gpu_index, None, None

    def start(self):
        try:
            fd, self.path = tempfile.mkstemp(suffix=".csv")
            os.close(fd)
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu), f"--query-gpu={self.QUERY}", "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=open(self.path, "w"), stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, smax, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in open(self.path):
            f = [s.strip() for s in line.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                smax = float(f[2])
            except ValueError:
                continue
            for nm, val in zip(names, f[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(nm)
        os.unlink(self.path)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": smax, "reasons": sorted(reasons), "samples": len(sm)}


# ------------------------------------------------------------------------------------------------------------ CPU arm
def _oracle_setup(wl):
    from oracle import reference_port as rp
    import resolution_pde_b200 as R
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    torch.manual_seed(0)
    model = getattr(R, wl["kind"])(**wl["ctor"])            # parameters only (the reference initialisers' shapes); never run
    sd = {k: v.detach().clone() for k, v in model.state_dict().items()}
    x, y = synth_fields(wl["batch"], wl["shape"], 1234)     # the GPU arm's rank-0 batch: same fields, same weights
    return rp, sd, x, y, cores


def run_reference(args, wl, name):
    """CPU arm: the oracle port of the reference on all host cores, every step the workload's full per-GPU batch."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    rp, sd, x, y, cores = _oracle_setup(wl)
    times = []
    for i in range(args.warmup + args.steps):
        t0 = time.perf_counter()
        rp.step_loss_and_grads(wl["kind"], sd, x, y, dtype=torch.float32, **wl["hyper"])
        dt = time.perf_counter() - t0
        if i >= args.warmup:
            times.append(dt)
    total = sum(times)
    B = wl["batch"]
    value = B * len(times) / total
    line = {"impl": "reference", "metric": "fwd+bwd samples/s", "value": value, "unit": "samples/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * total / len(times), "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": name, "model": wl["kind"], "per_gpu_batch": B, "global_batch": B, "grid": list(wl["shape"][1:]),
                       "step": "zero_grad + forward + RelativeL2 + backward"},
            "cpu_baseline": {"value": value, "unit": "samples/s", "cores": cores, "kind": "port",
                             "sample": f"{len(times)} step(s) of the full batch of {B} field(s), oracle port of the reference, torch {torch.__version__} CPU, {cores} threads"},
            "e2e": {"value": value, "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


def cpu_baseline(wl, budget_s=25.0):
    rp, sd, x, y, cores = _oracle_setup(wl)
    _, loss, _ = rp.step_loss_and_grads(wl["kind"], sd, x, y, dtype=torch.float32, **wl["hyper"])  # warm-up
    t0, n = time.perf_counter(), 0
    while True:
        rp.step_loss_and_grads(wl["kind"], sd, x, y, dtype=torch.float32, **wl["hyper"])
        n += 1
        el = time.perf_counter() - t0
        if el > budget_s or n >= 20:
            break
    B = wl["batch"]
    return {"value": B * n / el, "unit": "samples/s", "cores": cores, "kind": "port", "loss": float(loss),
            "sample": f"{n} fwd+bwd step(s) of the full batch of {B} field(s) ({el:.1f} s), oracle port of the reference, torch CPU {cores} threads"}


# ------------------------------------------------------------------------------------------------------------ GPU arm
def _short(name):
    name = re.sub(r"^void ", "", name)
    name = name.replace("rpde::", "").replace("tc::", "").replace("ff::", "")
    m = re.match(r"([A-Za-z0-9_:]+(?:<[^>(]*>)?)", name)
    return (m.group(1) if m else name)[:60]


def insitu_kernels(step_fn, lib, T_rows):
    """One eager step under torch.profiler (CUPTI kernel records, no serialisation) joined with the library's launch trace:
    per (kernel, shape) launches, summed duration and algorithmic FLOPs."""
    from torch.profiler import ProfilerActivity, profile
    for _ in range(2):
        step_fn()
    torch.cuda.synchronize()
    lib.rpde_trace_enable(1)
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        step_fn()
        torch.cuda.synchronize()
    n = lib.rpde_trace_count()
    buf = (ctypes.c_longlong * 8)()
    trace = []
    for i in range(n):
        lib.rpde_trace_get(i, buf)
        trace.append(tuple(buf))
    lib.rpde_trace_enable(0)
    evs = sorted((e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA), key=lambda e: e.time_range.start)
    gemm_evs = [e for e in evs if "gemm_tc_kernel" in e.name or "gemm_simt_kernel" in e.name]
    joined = len(gemm_evs) == len(trace)
    ti = 0
    agg = collections.OrderedDict()
    total = 0.0
    for e in evs:
        nm, shape, flops = _short(e.name), "", 0.0
        if e in gemm_evs and joined:
            M, N, K, b, sk, eng, fl, _ = trace[ti]
            ti += 1
            shape = f"M={M} N={N} K={K} b={b}" + (f" sk={sk}" if sk > 1 else "")
            flops = 2.0 * M * N * K * b
        elif "ff3_kernel" in e.name:
            shape, flops = f"T={T_rows} C=64 H=256", 2.0 * T_rows * (64 * 256 + 256 * 256 + 256 * 64)   # fwd: 3 layers; bwd: the dgrad chain
        elif "ff_wgrad_kernel<0>" in e.name or "ff_wgrad_kernel<(int)0>" in e.name:
            shape, flops = f"T={T_rows} dW2", 2.0 * T_rows * 256 * 256
        elif "ff_wgrad_kernel" in e.name:
            shape, flops = f"T={T_rows} dW1,dW3", 2.0 * T_rows * 2 * 64 * 256
        a = agg.setdefault((nm, shape), [0, 0.0, 0.0])
        a[0] += 1
        a[1] += e.device_time
        a[2] += flops
        total += e.device_time
    rows = [{"name": k[0], "shape": k[1], "n": v[0], "avg_us": v[1] / v[0], "total_ms": v[1] / 1e3, "share": v[1] / total,
             "tflops": (v[2] / (v[1] * 1e-6) / 1e12) if v[2] else None} for k, v in agg.items()]
    rows.sort(key=lambda r: -r["total_ms"])
    return rows, total / 1e3, len(evs), joined


def ncu_traffic(kernel_name, shape, prec):
    """DRAM bytes per launch of a kernel from the committed ncu --set full summaries (profiles/r2_ncu_traffic.json), or None"""
    p = os.path.join(ROOT, "profiles", "r2_ncu_traffic.json")
    if not os.path.exists(p):
        return None, None
    for rec in json.load(open(p)):
        if rec["kernel"] in kernel_name and rec.get("precision", prec) == prec and rec.get("shape", shape) == shape:
            return rec["dram_bytes_per_launch"], rec["source"]
    return None, None


def measure_tf32_peak(dev):
    """torch.matmul 8192^3 with allow_tf32 (cuBLAS), best of 10 and a 2 s back-to-back loop"""
    old = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = True
    try:
        a, b = torch.randn(8192, 8192, device=dev), torch.randn(8192, 8192, device=dev)
        for _ in range(3):
            a @ b
        best = 1e9
        for _ in range(10):
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record()
            a @ b
            e.record()
            torch.cuda.synchronize()
            best = min(best, s.elapsed_time(e))
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        n, t0 = 0, time.perf_counter()
        s.record()
        while time.perf_counter() - t0 < 2.0:
            for _ in range(20):
                a @ b
            n += 20
            torch.cuda.synchronize()
        e.record()
        torch.cuda.synchronize()
        fl = 2.0 * 8192 ** 3
        return {"tf32_tflops_burst": fl / (best * 1e-3) / 1e12, "tf32_tflops_sustained": fl * n / (s.elapsed_time(e) * 1e-3) / 1e12,
                "how": "torch.matmul fp32 8192^3 with allow_tf32=True (cuBLAS): best of 10, and back to back for 2 s"}
    finally:
        torch.backends.cuda.matmul.allow_tf32 = old


def run_workload(name, wl, args, prec, dev, world, rank, local, full):
    """Times one workload under one precision.  full: also e2e, clocks, in-situ kernels."""
    import torch.distributed as dist
    import resolution_pde_b200 as R
    from resolution_pde_b200.distributed import GradBucket
    from resolution_pde_b200.training import GraphedTrainStep

    R.set_precision(prec)
    torch.manual_seed(0)                                   # identical weights on every rank
    model = getattr(R, wl["kind"])(**wl["ctor"]).to(dev)
    loss_fn = R.RelativeL2Loss()
    B = wl["batch"]
    xh, yh = synth_fields(B, wl["shape"], 1234 + rank)
    xh, yh = xh.pin_memory(), yh.pin_memory()
    xd, yd = xh.to(dev), yh.to(dev)
    bucket = GradBucket(model.parameters())
    lib = R._lib.load()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev) if wl["bytes"] / 4 < 200e6 else None

    in_graph_reduce = world > 1 and not args.no_graph and os.environ.get("RPDE_BENCH_EAGER_ALLREDUCE", "0") == "0"
    gstep = None if args.no_graph else GraphedTrainStep(model, loss_fn, bucket=bucket if in_graph_reduce else None)
    launches_per_step = None
    if gstep is not None:
        c0 = lib.rpde_launch_count()
        gstep(xd, yd)                                       # warm-up + capture (one captured step = one count)
        launches_per_step = (lib.rpde_launch_count() - c0) // (gstep.warmup + 1)
        xs, ys = gstep.static_inputs(xd.shape, yd.shape)    # device-resident arm feeds the graph's own buffers
    else:
        xs, ys = xd, yd

    def eager_step():
        model.zero_grad(set_to_none=True)
        loss = loss_fn(model(xd), yd)
        loss.backward()
        return loss

    def step(x, y):
        if gstep is not None:
            loss = gstep(x, y)
        else:
            model.zero_grad(set_to_none=True)
            loss = loss_fn(model(x), y)
            loss.backward()
        if world > 1 and not in_graph_reduce:
            bucket.allreduce_mean()
        return loss

    def timed(n, from_host):
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(n)]
        last = None
        for s, e in ev:
            if flush is not None:
                flush.zero_()
            s.record()
            if from_host:
                if gstep is not None:
                    last = step(xh, yh).item()             # H2D into the graph's input buffers, replay, D2H of the loss
                else:
                    last = step(xh.to(dev, non_blocking=True), yh.to(dev, non_blocking=True)).item()
            else:
                last = step(xs, ys)
            e.record()
        torch.cuda.synchronize()
        return sum(s.elapsed_time(e) for s, e in ev), last

    for _ in range(args.warmup):
        step(xs, ys)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    sampler = ClockSampler(local)
    if rank == 0 and full:
        sampler.start()
    l0 = lib.rpde_launch_count()
    ms, loss = timed(args.steps, False)
    launches = lib.rpde_launch_count() - l0
    if launches_per_step is not None:
        launches = launches_per_step * args.steps           # replayed graph nodes (counted once, at capture)
    loss_val = float(loss.detach()) if torch.is_tensor(loss) else float(loss)
    extra = 0
    clocks = None
    if rank == 0 and full:
        if world == 1 and ms < 1500.0:
            # nvidia-smi samples every 100 ms: a timed region of a few ms would yield none.  Keep the same step running
            # (untimed) under the sampler until 1.5 s of load have been observed, and say so.
            t_end = time.perf_counter() + 1.5 - ms / 1e3
            while time.perf_counter() < t_end:
                step(xs, ys)
                extra += 1
                if extra % 16 == 0:
                    torch.cuda.synchronize()
            torch.cuda.synchronize()
        clocks = sampler.stop()
        if extra:
            clocks["note"] = (f"timed region {ms:.1f} ms is short against the 100 ms sampling period: sampled over it and over {extra} "
                              "untimed repetitions of the same step right after it (1.5 s of load in total)")
    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.barrier()
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = t.item()
    res = {"ms": ms, "loss": loss_val, "launches": launches, "clocks": clocks, "B": B, "grad_bytes": bucket.nbytes,
           "flush": flush is not None, "graph": gstep is not None, "in_graph_reduce": in_graph_reduce}
    if full and not args.no_e2e:
        timed(1, True)
        if world > 1:
            dist.barrier()
        ms_e, _ = timed(args.steps, True)
        t = torch.tensor([ms_e], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        res["e2e"] = {"value": world * B * args.steps / (t.item() / 1e3), "unit": "samples/s",
                      "h2d_bytes_per_step": int(xh.numel() * 4 + yh.numel() * 4), "d2h_bytes_per_step": 4}
    if full and rank == 0 and not args.no_kernels:
        T_rows = B * int(torch.tensor(wl["shape"][1:]).prod())
        try:
            res["kernels"] = insitu_kernels(eager_step, lib, T_rows)
        except Exception as ex:  # the profiler is diagnostics: never lose the timing line over it
            res["kernels_error"] = repr(ex)
    if gstep is not None:
        gstep.release()     # captured NCCL kernels must be gone before the process group is destroyed
    del gstep, model, bucket, flush
    gc.collect()
    torch.cuda.empty_cache()
    return res


def mix_dataset(rank, scale=1.0):
    data = []
    for n, cnt in MIX["groups"]:
        x, y = synth_fields(max(1, int(cnt * scale)), (1, n, n), 4321 + n + 1000 * rank)
        data += [(x[i], y[i]) for i in range(x.shape[0])]
    return data


def run_mix(args, prec, dev, world, rank):
    """one pass over the mixed-resolution set per step, every batch through GraphedTrainStep (one graph per (shape, batch))"""
    import torch.distributed as dist
    import resolution_pde_b200 as R
    from resolution_pde_b200.distributed import GradBucket
    R.set_precision(prec)
    torch.manual_seed(0)
    model = R.FNO2d(**MIX["ctor"]).to(dev)
    loss_fn = R.RelativeL2Loss()
    batches = list(R.ResolutionGroupedBatches(mix_dataset(rank), MIX["batch"], shuffle=False, device=dev))
    bucket = GradBucket(model.parameters())
    step = R.GraphedTrainStep(model, loss_fn, bucket=bucket if world > 1 else None)
    lib = R._lib.load()

    def epoch():
        tot = None
        for x, y in batches:
            l = step(x, y)
            tot = l if tot is None else tot + l
        return tot

    l0 = lib.rpde_launch_count()
    epoch()
    per_epoch = (lib.rpde_launch_count() - l0) // (step.warmup + 1)
    for _ in range(max(args.warmup - 1, 0)):
        epoch()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(args.steps):
        tot = epoch()
    e.record()
    torch.cuda.synchronize()
    t = torch.tensor([s.elapsed_time(e)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.barrier()
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = t.item()
    n_samples = sum(c for _, c in MIX["groups"])
    nbytes = sum(MIX["bytes_per_sample"][n] * c for n, c in MIX["groups"])
    loss = float(tot.detach()) / len(batches)
    step.release()
    if rank != 0:
        return None
    pk = peaks()
    step_s = ms / args.steps * 1e-3
    ach = nbytes / step_s / 1e9
    return {"metric": "fwd+bwd samples/s", "value": world * n_samples / step_s, "unit": "samples/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": step_s * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32 (3xTF32)" if prec == "3xtf32" else prec, "data": "synthetic",
            "config": {"workload": "fno2d_ns_mix", "model": "FNO2d", "per_gpu_batch": MIX["batch"], "precision_engine": prec,
                       "mix": {str(n): c for n, c in MIX["groups"]}, "batches_per_step": [int(x.shape[0]) for x, _ in batches],
                       "step": "one pass over the resolution-grouped mix: zero_grad + forward + RelativeL2 + backward per batch"
                               + (" + flat NCCL grad all-reduce in the graph" if world > 1 else ""),
                       "launch": "CUDA graph replay, one graph per (resolution, batch size) pair", "parallelism": f"dp{world}"},
            "loss": loss, "gpu_launches": per_epoch * args.steps,
            "roofline": {"bound": "hbm", "achieved": ach, "peak": pk["hbm"], "unit": "GB/s", "frac": ach / pk["hbm"], "traffic": None,
                         "kernel": "whole pass (ideal-fusion floor bytes of every sample / time): launch-bound", "peak_source": pk["src"]}}


def run_reference_mix(args):
    """CPU arm of the mix: the oracle port over resolution-grouped batches, a bounded sample (4 fields per resolution) per step"""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    from oracle import mres_port, reference_port as rp
    import resolution_pde_b200 as R
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    torch.manual_seed(0)
    model = R.FNO2d(**MIX["ctor"])
    sd = {k: v.detach().clone() for k, v in model.state_dict().items()}
    data = []
    for n, _ in MIX["groups"]:
        x, y = synth_fields(4, (1, n, n), 4321 + n)
        data += [(x[i], y[i]) for i in range(4)]
    batches = mres_port.grouped_batches(data, MIX["batch"])
    times = []
    for i in range(args.warmup + args.steps):
        t0 = time.perf_counter()
        for x, y in batches:
            rp.step_loss_and_grads("FNO2d", sd, x, y, dtype=torch.float32, **MIX["hyper"])
        if i >= args.warmup:
            times.append(time.perf_counter() - t0)
    value = len(data) * len(times) / sum(times)
    print(json.dumps({"impl": "reference", "metric": "fwd+bwd samples/s", "value": value, "unit": "samples/s", "n_gpus": args.gpus,
                      "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * sum(times) / len(times), "higher_is_better": True,
                      "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                      "config": {"workload": "fno2d_ns_mix", "model": "FNO2d", "per_gpu_batch": MIX["batch"],
                                 "mix": {str(n): c for n, c in MIX["groups"]}},
                      "cpu_baseline": {"value": value, "unit": "samples/s", "cores": cores, "kind": "port",
                                       "sample": f"{len(times)} pass(es) over 4 fields per resolution (64^2, 128^2, 256^2), grouped by resolution, oracle port, torch CPU {cores} threads"},
                      "e2e": {"value": value, "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))


def make_line(name, wl, args, prec, world, r, pk, full):
    B = r["B"]
    step_s = r["ms"] / args.steps * 1e-3
    line = {"metric": "fwd+bwd samples/s", "value": world * B / step_s, "unit": "samples/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": step_s * 1e3, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": {"bf16": "bf16 (FeedForward) / tf32", "tf32": "tf32", "3xtf32": "f32 (3xTF32)"}.get(prec, "f32"),
            "data": "synthetic",
            "config": {"workload": name, "model": wl["kind"], "per_gpu_batch": B, "global_batch": B * world,
                       "grid": list(wl["shape"][1:]), "precision_engine": prec,
                       "tolerance": {"outputs": TOLERANCE[prec][0], "gradients": TOLERANCE[prec][1],
                                     "what": "relative L2 per tensor vs the fp64 oracle, tests/test_gpu_baseline_configs.py"},
                       "parallelism": f"dp{world}",
                       "l2": "L2 flushed between timed steps" if r["flush"] else "working set >> L2 (no flush)",
                       "step": "zero_grad + forward + RelativeL2 + backward" + (
                           (" + flat NCCL grad all-reduce" + (" (captured in the step graph)" if r.get("in_graph_reduce") else " (eager)")) if world > 1 else ""),
                       "launch": "CUDA graph replay (one graph per input shape)" if r["graph"] else "eager",
                       "grad_bytes": r["grad_bytes"]},
            "loss": r["loss"], "gpu_launches": r["launches"]}
    vdiv = VARIANT_DIV.get(prec)
    rs = {"flops_achieved_tflops": wl["flops"] / step_s / 1e12, "tensor_peak_tflops": pk["tensor"],
          "frac_tensor": wl["flops"] / step_s / 1e12 / pk["tensor"],
          "hbm_achieved_gbs": wl["bytes"] / step_s / 1e9, "hbm_peak_gbs": pk["hbm"], "frac_hbm": wl["bytes"] / step_s / 1e9 / pk["hbm"],
          "bound": wl["bound"], "peak_source": pk["src"] + ", sustained bf16",
          "note": "algorithmic FLOPs (dense truncated-DFT formulation, SURVEY 8d) and ideal-fusion floor bytes of the whole fwd+bwd step / step time"}
    if vdiv:
        rs["frac_variant_peak"] = wl["flops"] / step_s / 1e12 / (pk["tensor"] / vdiv)
        rs["variant_peak_tflops"] = pk["tensor"] / vdiv
    rs["frac"] = rs["frac_tensor"] if wl["bound"] == "tensor" else rs["frac_hbm"]
    line["roofline_step"] = rs
    if not full:
        return line
    line["clocks"] = r["clocks"]
    line["e2e"] = r.get("e2e")
    roof = None
    if "kernels" in r:
        rows, tot_ms, nk, joined = r["kernels"]
        line["kernels"] = {"source": "torch.profiler (CUPTI) over one eager step + rpde_trace shapes" + ("" if joined else " (trace length mismatch: shapes omitted)"),
                           "launches": nk, "sum_ms": tot_ms,
                           "top": [{k: (round(v, 4) if isinstance(v, float) else v) for k, v in row.items()} for row in rows[:12]]}
        top = rows[0]
        if wl["bound"] == "tensor" and top["tflops"]:
            traffic, src = ncu_traffic(top["name"], top["shape"], prec)
            roof = {"bound": "tensor", "achieved": top["tflops"], "peak": pk["tensor"], "unit": "TFLOP/s", "frac": top["tflops"] / pk["tensor"],
                    "traffic": traffic, "traffic_source": src,
                    "kernel": f"{top['name']} {top['shape']}", "launches_per_step": top["n"], "avg_us_in_situ": top["avg_us"],
                    "share_of_step": top["share"], "peak_source": pk["src"] + ", sustained bf16 (kernel timed inside the step)",
                    "how": "algorithmic FLOPs of the kernel's launches in one step / their summed in-situ duration"}
            if vdiv:
                roof["frac_variant_peak"] = top["tflops"] / (pk["tensor"] / vdiv)
    if roof is None:
        if wl["bound"] == "tensor":
            roof = {"bound": "tensor", "achieved": rs["flops_achieved_tflops"], "peak": pk["tensor"], "unit": "TFLOP/s", "frac": rs["frac_tensor"],
                    "traffic": None, "kernel": "whole fwd+bwd step (no in-situ table)", "peak_source": pk["src"]}
        else:
            roof = {"bound": "hbm", "achieved": rs["hbm_achieved_gbs"], "peak": pk["hbm"], "unit": "GB/s", "frac": rs["frac_hbm"], "traffic": None,
                    "kernel": "whole fwd+bwd step (ideal-fusion floor bytes / step time): launch-bound, no single dominant kernel",
                    "peak_source": pk["src"]}
    roof["step_frac"] = rs["frac"]
    line["roofline"] = roof
    return line


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default=DEFAULT, choices=sorted(WORKLOADS) + ["fno2d_ns_mix", "all"])
    ap.add_argument("--precision", default=os.environ.get("RPDE_PRECISION", "auto"))
    ap.add_argument("--batch", type=int, default=0, help="per-GPU batch override")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-variants", action="store_true", help="skip the tf32 / bf16 variant timings of the same workload")
    ap.add_argument("--no-kernels", action="store_true", help="skip the in-situ kernel table (torch.profiler pass)")
    ap.add_argument("--no-graph", action="store_true", help="launch every op eagerly instead of replaying the captured step")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup
    if args.workload == "fno2d_ns_mix" and args.impl == "reference":
        return run_reference_mix(args)
    name = DEFAULT if args.workload in ("all", "fno2d_ns_mix") else args.workload
    wl = dict(WORKLOADS[name])
    if args.batch:
        wl["flops"] *= args.batch / wl["batch"]
        wl["bytes"] *= args.batch / wl["batch"]
        wl["batch"] = args.batch
    if args.impl == "reference":
        return run_reference(args, wl, name)

    import torch.distributed as dist
    import resolution_pde_b200 as R

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    prec = args.precision
    if prec == "auto":
        prec = "3xtf32" if R._lib.load().rpde_set_precision(1) == 0 else "fp32"
    pk = peaks()
    if args.workload == "fno2d_ns_mix":
        line = run_mix(args, prec, dev, world, rank)
        if rank == 0:
            print(json.dumps(line))
        if world > 1:
            dist.destroy_process_group()
        return

    r = run_workload(name, wl, args, prec, dev, world, rank, local, full=True)
    line = make_line(name, wl, args, prec, world, r, pk, full=True) if rank == 0 else None

    # the other precision variants of the same workload (same steps / warm-up; their own stated tolerances)
    if not args.no_variants and world == 1:
        variants = {}
        for vp in ("tf32", "bf16"):
            if vp == prec:
                continue
            rv = run_workload(name, wl, args, vp, dev, world, rank, local, full=False)
            lv = make_line(name, wl, args, vp, world, rv, pk, full=False)
            variants[vp] = {"value": lv["value"], "ms_per_step": lv["ms_per_step"], "loss": lv["loss"], "gpu_launches": lv["gpu_launches"],
                            "tolerance": lv["config"]["tolerance"], "roofline_step": lv["roofline_step"], "dtype": lv["dtype"]}
        line["variants"] = variants
    if args.workload == "all" and world == 1:
        extra = []
        for other in sorted(WORKLOADS):
            if other == name:
                continue
            wo = dict(WORKLOADS[other])
            ro = run_workload(other, wo, args, prec, dev, world, rank, local, full=False)
            extra.append(make_line(other, wo, args, prec, world, ro, pk, full=False))
        extra.append(run_mix(args, prec, dev, world, rank))
        line["extra"] = extra
    if rank == 0 and world == 1 and wl["bound"] == "tensor" and not args.no_variants:
        try:
            pm = measure_tf32_peak(dev)
            line["peaks_measured_in_run"] = pm

            def against_measured(rs, p):
                # the variant's tensor peak as MEASURED on this box in this run (cuBLAS TF32, sustained; a 3xTF32 product
                # issues three TF32 UMMAs), beside the figure derived from the BF16 peak
                div = {"tf32": 1.0, "3xtf32": 3.0}.get(p)
                if div:
                    rs["variant_peak_measured_tflops"] = pm["tf32_tflops_sustained"] / div
                    rs["frac_variant_peak_measured"] = rs["flops_achieved_tflops"] / rs["variant_peak_measured_tflops"]
            against_measured(line["roofline_step"], prec)
            for vp, v in line.get("variants", {}).items():
                against_measured(v["roofline_step"], vp)
        except Exception as ex:
            line["peaks_measured_in_run"] = {"error": repr(ex)}
    if rank == 0:
        if not args.no_cpu_baseline and world == 1:
            cb = cpu_baseline(wl)
            tol = TOLERANCE[prec][0]
            err = abs(line["loss"] - cb["loss"])
            line["loss_check"] = {"oracle_loss_fp32_cpu": cb["loss"], "abs_err": err, "tolerance": tol * max(1.0, abs(cb["loss"])),
                                  "ok": bool(err < tol * max(1.0, abs(cb["loss"]))),
                                  "what": "this run's loss vs the oracle's loss on the same batch and weights"}
            if not line["loss_check"]["ok"]:
                print(f"bench.py: LOSS CHECK FAILED: {line['loss']} vs oracle {cb['loss']}", file=sys.stderr)
            line["cpu_baseline"] = cb
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
