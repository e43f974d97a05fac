"""In-situ kernel durations of one eager training step (torch.profiler / CUPTI, no ncu serialisation):
usage: step_kernels.py [workload] [precision]   -> per-kernel-name totals of the last of 3 steps, launch order kept for GEMMs"""
import collections, os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
import resolution_pde_b200 as R
name = sys.argv[1] if len(sys.argv) > 1 else "ffno2d_ns256"
R.set_precision(sys.argv[2] if len(sys.argv) > 2 else "3xtf32")
wl = bench.WORKLOADS[name]
torch.manual_seed(0)
dev = torch.device("cuda")
model = getattr(R, wl["kind"])(**wl["ctor"]).to(dev)
loss_fn = R.RelativeL2Loss()
x, y = bench.synth_fields(wl["batch"], wl["shape"], 1234)
x, y = x.to(dev), y.to(dev)
def step():
    model.zero_grad(set_to_none=True)
    loss_fn(model(x), y).backward()
for _ in range(3): step()
torch.cuda.synchronize()
from torch.profiler import profile, ProfilerActivity
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    step(); torch.cuda.synchronize()
evs = sorted((e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA), key=lambda e: e.time_range.start)
tot = sum(e.device_time for e in evs)
print(f"{len(evs)} kernels, {tot / 1e3:.2f} ms device time")
if os.environ.get("SEQ"):
    for e in evs:
        if e.device_time > float(os.environ.get("SEQ_MIN", "50")): print(f"{e.device_time:9.1f} us  {e.name[:90]}")
else:
    agg = collections.OrderedDict()
    for e in evs:
        a = agg.setdefault(e.name[:100], [0, 0.0]); a[0] += 1; a[1] += e.device_time
    for k, (n, us) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:14]:
        print(f"{us / tot * 100:5.1f}%  n={n:3d}  avg={us / n:8.1f} us  {k}")
