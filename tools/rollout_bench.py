"""Autoregressive rollout sweep (BASELINE config 5, SURVEY 8d "Rollout (C5): 16 sequential steps, no_grad, with the
decode/encode re-normalisation"): one rollout batch (encode t=0, 16 model steps with re-normalisation, decode, 16 relative-L2
losses) as ONE CUDA-graph replay (`RolloutEvaluator`), against the same batch run eagerly on the device and against the CPU
restatement of the reference loop (oracle/rollout_port.py on the oracle's model port, all host threads, bounded sample).
One JSON line per (model, resolution): rollouts/s = trajectories rolled out per second; `state_hbm_frac` = the rollout
bookkeeping's own algorithmic bytes (per step: read prediction, write trajectory slot, write next state; plus one pass over
predictions and truth for the losses) / time of those kernels alone / measured copy bandwidth."""
import json, os, sys, time, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
import resolution_pde_b200 as R
from resolution_pde_b200.rollout import RolloutEvaluator, SimpleNormalizer
from oracle import reference_port as rp, rollout_port as port

R.set_precision(os.environ.get("RPDE_PRECISION", "3xtf32"))
pk = bench.peaks()
T = 16
xn, yn = SimpleNormalizer(0.05, 1.3), SimpleNormalizer(0.04, 1.25)
CPU_BUDGET = float(os.environ.get("ROLLOUT_CPU_BUDGET", "6"))
cases = []
for n in (32, 64, 128, 256, 512, 1024):
    cases.append(("ffno1d_ks512", (32, T + 1, n)))
for n, b in ((1024, 16),):
    cases.append(("fno1d_burgers1024", (b, T + 1, n)))
for m, b in ((64, 16), (128, 16), (256, 16), (512, 4), (1024, 1)):
    cases.append(("ffno2d_ns256", (b, T + 1, m, m)))
for m, b in ((64, 16), (256, 16), (1024, 1)):
    cases.append(("fno2d_ns256", (b, T + 1, m, m)))
if len(sys.argv) > 1:
    cases = [c for c in cases if c[0] in sys.argv[1:]]


def timed(fn, reps):
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); s.record()
    for _ in range(reps): out = fn()
    e.record(); torch.cuda.synchronize()
    return s.elapsed_time(e) / reps, out


class Ident(torch.nn.Module):
    def __init__(self):
        super().__init__()
        self.p = torch.nn.Parameter(torch.zeros(1))

    def forward(self, x):
        return x


for wname, shape in cases:
    wl = bench.WORKLOADS[wname]
    torch.manual_seed(0)
    model = getattr(R, wl["kind"])(**wl["ctor"])
    sd = {k: v.detach().clone() for k, v in model.state_dict().items()}
    model = model.cuda().eval()
    B = shape[0]
    spatial = shape[2:]
    # a synthetic trajectory: standardised band-limited fields (bench.synth_fields) per time level
    f, _ = bench.synth_fields(B * shape[1], (1,) + tuple(spatial), 77)
    traj_h = (0.05 + 1.3 * f.reshape((B, shape[1]) + tuple(spatial))).contiguous()
    traj = traj_h.cuda()
    lc0 = R._lib.load().rpde_launch_count()
    ev_e = RolloutEvaluator(model, T, xn, yn, use_graph=False)
    ev_e(traj); torch.cuda.synchronize()
    launches = R._lib.load().rpde_launch_count()
    ev_e(traj); torch.cuda.synchronize()
    launches = R._lib.load().rpde_launch_count() - launches
    reps = 20 if traj.numel() < (1 << 24) else 5
    eager_ms, _ = timed(lambda: ev_e(traj), reps)
    ev_g = RolloutEvaluator(model, T, xn, yn, use_graph=True)
    for _ in range(3): ev_g(traj)
    graph_ms, (loss, steps, preds) = timed(lambda: ev_g(traj), reps)
    # bookkeeping kernels alone (identity model): what the rollout adds around the T forwards
    ev_i = RolloutEvaluator(Ident().cuda(), T, xn, yn, use_graph=True)
    for _ in range(3): ev_i(traj)
    book_ms, _ = timed(lambda: ev_i(traj), reps)
    S = traj[0, 0].numel()
    book_bytes = 4 * B * S * (2 + 3 * T + 2 * T)
    # CPU: the reference loop on the oracle's model port, all host threads, bounded
    hyper = dict(wl["hyper"])
    kind = wl["kind"]
    fwd = {"FNO1d": rp.fno1d_forward, "FNO2d": rp.fno2d_forward, "FFNO1D": rp.ffno1d_forward, "FFNO2D": rp.ffno2d_forward}[kind]
    if "activation" in hyper and kind.startswith("FNO"):
        import torch.nn.functional as F
        hyper["activation"] = {"relu": F.relu, "gelu": F.gelu}[hyper["activation"]]
    cpu_B = min(B, 2 if len(spatial) == 2 and spatial[0] >= 256 else B)
    cpu_T = T if S * cpu_B <= (1 << 18) else 2
    with torch.no_grad():
        t0 = time.perf_counter(); n = 0
        while True:
            cl, cs, cp = port.rollout_batch_loss(lambda x: fwd(sd, x, **hyper), traj_h[:cpu_B], cpu_T, (xn.mean, xn.std, xn.eps), (yn.mean, yn.std, yn.eps))
            n += 1
            if time.perf_counter() - t0 > CPU_BUDGET: break
        cpu_s_per_traj_step = (time.perf_counter() - t0) / n / cpu_B / cpu_T
    err = ((preds[:cpu_B, :cpu_T].double().cpu() - cp.double()).norm() / cp.double().norm()).item()
    cpu_rps = 1.0 / (cpu_s_per_traj_step * T)
    print(json.dumps({"workload": f"{wname} rollout T={T}", "traj_shape": list(shape), "engine": R.get_precision(),
                      "graph_ms": round(graph_ms, 4), "eager_ms": round(eager_ms, 4), "launches_per_rollout": int(launches),
                      "rollouts_per_s": round(B / graph_ms * 1e3, 1), "model_steps_per_s": round(B * T / graph_ms * 1e3, 1),
                      "bookkeeping_ms": round(book_ms, 4), "state_hbm_frac": round(book_bytes / (book_ms * 1e-3) / 1e9 / pk["hbm"], 3),
                      "cpu_rollouts_per_s": round(cpu_rps, 3), "cpu_threads": torch.get_num_threads(),
                      "cpu_sample": f"{n} x ({cpu_B} traj x {cpu_T} steps), scaled to {T} steps",
                      "speedup_vs_cpu": round(B / graph_ms * 1e3 / cpu_rps, 1), "batch_loss": round(float(loss), 6),
                      "rel_err_preds_vs_cpu_fp32": float(f"{err:.2e}")}), flush=True)
    del model, ev_e, ev_g, ev_i, traj
    torch.cuda.empty_cache()
