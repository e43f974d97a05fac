"""2-rank data-parallel check (torchrun): gradients after the all-reduce vs rank 0's one-rank gradient at the global batch.
usage: torchrun --nproc-per-node 2 tools/dp_check.py [graph|eager] [case] [engine]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
import resolution_pde_b200 as R
from resolution_pde_b200.distributed import GradBucket, shard_batch
from resolution_pde_b200.training import GraphedTrainStep

mode = sys.argv[1] if len(sys.argv) > 1 else "graph"
case = sys.argv[2] if len(sys.argv) > 2 else "ffno2d"
engine = sys.argv[3] if len(sys.argv) > 3 else "3xtf32"
CASES = {
    "ffno2d": dict(kind="FFNO2D", ctor=dict(in_channels=1, out_channels=1, width=64, n_layers=2, n_modes=16, factor=4, ff_weight_norm=True,
                                            n_ff_layers=3, layer_norm=True, dropout=0.0), shape=(4, 1, 64, 64)),
    "fno1d": dict(kind="FNO1d", ctor=dict(in_channels=1, out_channels=1, modes=16, width=64), shape=(8, 1, 1024)),
}
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
def log(*a):
    print(f"[rank {rank} {time.strftime('%X')}]", *a, flush=True)
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
log("pg up")
c = CASES[case]
R.set_precision(engine)
torch.manual_seed(0)
model = getattr(R, c["kind"])(**c["ctor"]).to(dev)
g = torch.Generator().manual_seed(1)
x, y = torch.randn(*c["shape"], generator=g), torch.randn(*c["shape"], generator=g)
lo, hi = shard_batch(c["shape"][0], rank, world)
bucket = GradBucket(model.parameters())
loss_fn = R.RelativeL2Loss()
if mode == "graph":
    step = GraphedTrainStep(model, loss_fn, bucket=bucket)
    for i in range(3):
        step(x[lo:hi].to(dev), y[lo:hi].to(dev))
        torch.cuda.synchronize()
        log("graph step", i)
else:
    for i in range(2):
        model.zero_grad(set_to_none=True)
        loss_fn(model(x[lo:hi].to(dev)), y[lo:hi].to(dev)).backward()
        bucket.allreduce_mean()
        torch.cuda.synchronize()
        log("eager step", i)
got = {k: p.grad.detach().clone() for k, p in model.named_parameters()}
model.zero_grad(set_to_none=True)
loss_fn(model(x.to(dev)), y.to(dev)).backward()
torch.cuda.synchronize()
worst = max(((got[k] - p.grad).norm() / p.grad.norm().clamp_min(1e-30)).item() for k, p in model.named_parameters() if p.grad.abs().max() > 1e-8)
log(f"mode={mode} case={case} engine={engine} worst relative gradient difference vs one-rank global batch: {worst:.3e}")
t = torch.ones(1, device=dev)
dist.all_reduce(t)
torch.cuda.synchronize()
log("eager collective after the replays ok", t.item())
if mode == "graph":
    step.release()
    log("graph released")
dist.barrier()
torch.cuda.synchronize()
log("barrier ok")
dist.destroy_process_group()
log("process group destroyed")
