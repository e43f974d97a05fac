"""Training step WITH the optimizer: graph-captured step + FusedAdamW against the reference's loop shape
(eager step + torch.optim.AdamW).  usage: optim_bench.py [workload ...]   -> one JSON line per workload"""
import json, os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
import resolution_pde_b200 as R

R.set_precision("3xtf32")
dev = torch.device("cuda")
loss_fn = R.RelativeL2Loss()


def timed(fn, n=10, warm=3):
    for _ in range(warm): fn()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); s.record()
    for _ in range(n): fn()
    e.record(); torch.cuda.synchronize()
    return s.elapsed_time(e) / n


for name in (sys.argv[1:] or ["ffno2d_ns256", "fno2d_ns128", "ffno1d_ks512", "fno1d_burgers1024"]):
    wl = bench.WORKLOADS[name]
    x, y = bench.synth_fields(wl["batch"], wl["shape"], 1234)
    x, y = x.to(dev), y.to(dev)
    torch.manual_seed(0)
    m1 = getattr(R, wl["kind"])(**wl["ctor"]).to(dev)
    o1 = torch.optim.AdamW(m1.parameters(), lr=1e-3, weight_decay=1e-4)
    def eager_torch():
        m1.zero_grad(set_to_none=True); loss_fn(m1(x), y).backward(); o1.step()
    torch.manual_seed(0)
    m2 = getattr(R, wl["kind"])(**wl["ctor"]).to(dev)
    o2 = R.FusedAdamW(m2.parameters(), lr=1e-3, weight_decay=1e-4)
    step = R.GraphedTrainStep(m2, loss_fn, optimizer=o2)
    step(x, y)
    sx, sy = step.static_inputs(x.shape, y.shape)
    nparam = sum(p.numel() * (2 if p.is_complex() else 1) for p in m2.parameters())
    def opt_only_fused():
        o2.step()
    def opt_only_torch():
        o1.step()
    r = {"workload": name, "param_floats": nparam, "tensors": len(list(m2.parameters())),
         "step_ms_eager_torch_adamw": round(timed(eager_torch), 4),
         "step_ms_graph_fused_adamw": round(timed(lambda: step(sx, sy)), 4),
         "adamw_ms_torch_foreach": round(timed(opt_only_torch, 20), 4), "adamw_ms_fused": round(timed(opt_only_fused, 20), 4)}
    print(json.dumps(r), flush=True)
