#!/usr/bin/env python
"""Benchmark of the spectral-convolution hot path: fwd + RelativeL2 + bwd samples/s of the FNO / F-FNO models.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload ffno2d_ns256|fno1d_burgers1024|ffno1d_ks512|fno2d_ns{64,128,256}]
                  [--precision fp32|3xtf32|tf32] [--impl reference]

Prints ONE JSON line (rank 0).  `value` is whole-job samples/s with the inputs resident in HBM; `e2e` is the same
step driven from pinned host buffers (H2D of x and y, D2H of the loss every step).  `--impl reference` times the
CPU restatement of the reference (oracle/reference_port.py; the reference is Python and cannot travel to the GPU
box) on the host cores with the same config/metric.  Multi-GPU (torchrun): batch-sharded weak scaling, one flat
NCCL all-reduce of the gradients inside the timed step.
"""
from __future__ import annotations

import argparse
import json
import os
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
DEFAULT = "ffno2d_ns256"


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return dict(hbm=d["hbm_gbs"], tensor=d["bf16_tflops_sustained"], tensor_burst=d["bf16_tflops"], src="measured")
    return dict(hbm=6650.0, tensor=1400.0, tensor_burst=1590.0, src="fallback")


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


def run_reference(args, wl, name):
    """CPU arm: the oracle port of the reference on the host cores, one bounded sample (1 field) per step."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import reference_port as rp
    import resolution_pde_b200 as R
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    torch.manual_seed(0)
    model = getattr(R, wl["kind"])(**wl["ctor"])            # parameters only (reference initialisers' shapes); never run
    sd = {k: v.detach().clone() for k, v in model.state_dict().items()}
    sample_batch = 1 if len(wl["shape"]) == 3 else 4
    x, y = synth_fields(sample_batch, wl["shape"], 1234)
    times = []
    for i in range(args.warmup + args.steps):
        t0 = time.perf_counter()
        rp.step_loss_and_grads(wl["kind"], sd, x, y, dtype=torch.float32, **wl["hyper"])
        dt = time.perf_counter() - t0
        if i >= args.warmup:
            times.append(dt)
    total = sum(times)
    value = sample_batch * len(times) / total
    line = {"impl": "reference", "metric": "fwd+bwd samples/s", "value": value, "unit": "samples/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * total / len(times), "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": name, "model": wl["kind"], "per_gpu_batch": wl["batch"], "grid": list(wl["shape"][1:]),
                       "sample": f"{sample_batch} field(s) per step"},
            "cpu_baseline": {"value": value, "unit": "samples/s", "cores": cores, "kind": "port",
                             "sample": f"{len(times)} step(s) of {sample_batch} field(s), torch {torch.__version__} CPU, {cores} threads"},
            "e2e": {"value": value, "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


def cpu_baseline(wl, budget_s=25.0):
    from oracle import reference_port as rp
    import resolution_pde_b200 as R
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    torch.manual_seed(0)
    model = getattr(R, wl["kind"])(**wl["ctor"])
    sd = {k: v.detach().clone() for k, v in model.state_dict().items()}
    nb = 1 if len(wl["shape"]) == 3 else 4
    x, y = synth_fields(nb, wl["shape"], 1234)
    rp.step_loss_and_grads(wl["kind"], sd, x, y, dtype=torch.float32, **wl["hyper"])  # warm-up
    t0, n = time.perf_counter(), 0
    while True:
        rp.step_loss_and_grads(wl["kind"], sd, x, y, dtype=torch.float32, **wl["hyper"])
        n += 1
        el = time.perf_counter() - t0
        if el > budget_s or n >= 20:
            break
    return {"value": nb * n / el, "unit": "samples/s", "cores": cores, "kind": "port",
            "sample": f"{n} fwd+bwd step(s) of {nb} field(s) ({el:.1f} s), oracle port of the reference, torch CPU {cores} threads"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default=DEFAULT, choices=sorted(WORKLOADS))
    ap.add_argument("--precision", default=os.environ.get("RPDE_PRECISION", "auto"))
    ap.add_argument("--batch", type=int, default=0, help="per-GPU batch override")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-graph", action="store_true", help="launch every op eagerly instead of replaying the captured step")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup
    name = args.workload
    wl = dict(WORKLOADS[name])
    if args.batch:
        wl["flops"] *= args.batch / wl["batch"]
        wl["bytes"] *= args.batch / wl["batch"]
        wl["batch"] = args.batch
    if args.impl == "reference":
        return run_reference(args, wl, name)

    import torch.distributed as dist
    import resolution_pde_b200 as R
    from resolution_pde_b200.distributed import GradBucket

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

    from resolution_pde_b200.training import GraphedTrainStep
    gstep = None if args.no_graph else GraphedTrainStep(model, loss_fn)
    launches_per_step = None
    if gstep is not None:
        c0 = lib.rpde_launch_count()
        gstep(xd, yd)                                       # warm-up + capture (one captured step = one count)
        launches_per_step = (lib.rpde_launch_count() - c0) // (gstep.warmup + 1)
        xd, yd = gstep.static_inputs(xd.shape, yd.shape)    # device-resident arm feeds the graph's own buffers

    def step(x, y):
        if gstep is not None:
            loss = gstep(x, y)
        else:
            model.zero_grad(set_to_none=True)
            loss = loss_fn(model(x), y)
            loss.backward()
        if world > 1:
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
                last = step(xd, yd)
            e.record()
        torch.cuda.synchronize()
        return sum(s.elapsed_time(e) for s, e in ev), last

    for _ in range(args.warmup):
        step(xd, yd)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    l0 = lib.rpde_launch_count() if hasattr(lib, "rpde_launch_count") else 0
    ms, loss = timed(args.steps, False)
    launches = (lib.rpde_launch_count() - l0) if hasattr(lib, "rpde_launch_count") else None
    if launches_per_step is not None:
        launches = launches_per_step * args.steps           # replayed graph nodes (counted once, at capture)
    loss_val = float(loss.detach()) if torch.is_tensor(loss) else float(loss)
    extra = 0
    if rank == 0 and world == 1 and ms < 1500.0:
        # nvidia-smi samples every 100 ms: a timed region of a few ms would yield none.  Keep the same step running
        # (untimed) under the sampler until 1.5 s of load have been observed, and say so.
        t_end = time.perf_counter() + 1.5 - ms / 1e3
        while time.perf_counter() < t_end:
            step(xd, yd)
            extra += 1
            if extra % 16 == 0:
                torch.cuda.synchronize()
        torch.cuda.synchronize()
    clocks = sampler.stop() if rank == 0 else None
    if clocks is not None and extra:
        clocks["note"] = (f"timed region {ms:.1f} ms is short against the 100 ms sampling period: sampled over it and over {extra} "
                          "untimed repetitions of the same step right after it (1.5 s of load in total)")
    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.barrier()
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = t.item()
    e2e = None
    if not args.no_e2e:
        timed(1, True)
        if world > 1:
            dist.barrier()
        ms_e, _ = timed(args.steps, True)
        t = torch.tensor([ms_e], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e = {"value": world * B * args.steps / (t.item() / 1e3), "unit": "samples/s",
               "h2d_bytes_per_step": int(xh.numel() * 4 + yh.numel() * 4), "d2h_bytes_per_step": 4}

    # dominant kernel, timed alone with CUDA events: the FeedForward hidden GEMM (F-FNO) / the block synthesis GEMM (FNO)
    pk = peaks()
    roof = None
    if rank == 0:
        if wl["bound"] == "tensor":
            T = B * int(torch.tensor(wl["shape"][1:]).prod())
            d = wl["ctor"]["width"] * wl["ctor"]["factor"]
            a = torch.randn(T, d, device=dev)
            w = torch.randn(d, d, device=dev)
            for _ in range(2):
                R.ops.matmul(a, w, False, True)
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record()
            reps = 5
            for _ in range(reps):
                R.ops.matmul(a, w, False, True)
            e.record()
            torch.cuda.synchronize()
            kms = s.elapsed_time(e) / reps
            ach = 2.0 * T * d * d / (kms * 1e-3) / 1e12
            # DRAM bytes per launch of this kernel from the ncu --set full capture (profiles/r1f_ncu_full_ffgemm_3xtf32.csv:
            # dram__bytes_read.sum + dram__bytes_write.sum); only known for the captured shape / engine
            traffic = 2.1020e9 if (name == "ffno2d_ns256" and B == 16 and prec == "3xtf32") else None
            vdiv = {"3xtf32": 6.0, "tf32": 2.0}.get(prec)
            roof = {"bound": "tensor", "achieved": ach, "peak": pk["tensor_burst"], "unit": "TFLOP/s", "frac": ach / pk["tensor_burst"],
                    "traffic": traffic, "kernel": f"contraction engine ({prec}), FeedForward hidden GEMM {T}x{d}x{d}", "kernel_ms": kms,
                    "algorithmic_bytes": 4.0 * (2 * T * d + d * d), "hbm_frac": 4.0 * (2 * T * d + d * d) / (kms * 1e-3) / 1e9 / pk["hbm"],
                    "peak_variant": ({"engine": prec, "peak": pk["tensor_burst"] / vdiv, "frac": ach / (pk["tensor_burst"] / vdiv),
                                      "note": "kind::tf32 UMMA runs at half the bf16 rate; 3xTF32 issues three UMMAs per credited product"}
                                     if vdiv else None),
                    "peak_source": pk["src"] + " bf16 burst (kernel timed alone)",
                    "step": {"achieved": wl["flops"] / (ms / args.steps * 1e-3) / 1e12, "peak": pk["tensor"], "unit": "TFLOP/s",
                             "frac": wl["flops"] / (ms / args.steps * 1e-3) / 1e12 / pk["tensor"],
                             "hbm_frac": wl["bytes"] / (ms / args.steps * 1e-3) / 1e9 / pk["hbm"],
                             "note": "algorithmic FLOPs (dense truncated-DFT formulation) of the whole fwd+bwd step / step time, vs sustained bf16"}}
        else:
            ach = wl["bytes"] / (ms / args.steps * 1e-3) / 1e9
            roof = {"bound": "hbm", "achieved": ach, "peak": pk["hbm"], "unit": "GB/s", "frac": ach / pk["hbm"], "traffic": None,
                    "kernel": "whole fwd+bwd step (ideal-fusion floor bytes / step time)", "peak_source": pk["src"]}
    if rank == 0:
        line = {"metric": "fwd+bwd samples/s", "value": world * B * args.steps / (ms / 1e3), "unit": "samples/s", "n_gpus": world,
                "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": {"workload": name, "model": wl["kind"], "per_gpu_batch": B, "global_batch": B * world,
                           "grid": list(wl["shape"][1:]), "precision_engine": prec, "parallelism": f"dp{world}",
                           "l2": "L2 flushed between timed steps" if flush is not None else "working set >> L2 (no flush)",
                           "step": "zero_grad + forward + RelativeL2 + backward" + (" + flat NCCL grad all-reduce" if world > 1 else ""),
                           "launch": "eager" if gstep is None else "CUDA graph replay (one graph per input shape)",
                           "grad_bytes": bucket.nbytes},
                "loss": loss_val, "clocks": clocks, "e2e": e2e, "gpu_launches": launches, "roofline": roof}
        if not args.no_cpu_baseline and world == 1:
            line["cpu_baseline"] = cpu_baseline(wl)
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
