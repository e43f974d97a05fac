"""Multi-GPU rollout evaluation check: under torchrun, every rank evaluates its round-robin share of the
trajectory batches (`evaluate_rollout(distributed=True)`: one NCCL all-reduce of (sum, count) at the end) and the result must
equal the single-process evaluation of all batches on the same rank.  Prints one JSON line from rank 0 with device-timed
throughput (max over ranks).   torchrun --nproc-per-node N tools/rollout_dist_check.py"""
import json, os, sys, torch
import torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
import resolution_pde_b200 as R
from resolution_pde_b200.rollout import SimpleNormalizer, evaluate_rollout

rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
R.set_precision("3xtf32")
T, NB = 16, 16
out = {}
for wname, shape in (("fno1d_burgers1024", (16, T + 1, 1024)), ("ffno2d_ns256", (4, T + 1, 128, 128))):
    wl = bench.WORKLOADS[wname]
    torch.manual_seed(0)
    model = getattr(R, wl["kind"])(**wl["ctor"]).cuda().eval()
    xn, yn = SimpleNormalizer(0.05, 1.3), SimpleNormalizer(0.04, 1.25)
    f, _ = bench.synth_fields(NB * shape[0] * shape[1], (1,) + tuple(shape[2:]), 99)
    batches = list((0.05 + 1.3 * f.reshape((NB, shape[0], shape[1]) + tuple(shape[2:]))).cuda().unbind(0))
    ev = R.RolloutEvaluator(model, T, xn, yn)
    single = evaluate_rollout(model, batches, T, xn, yn, distributed=False, evaluator=ev)   # also the warm-up (graph capture)
    sharded = evaluate_rollout(model, batches, T, xn, yn, distributed=True, evaluator=ev)
    assert abs(single - sharded) <= 2e-6 * abs(single), (rank, single, sharded)
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    if world > 1: dist.barrier()
    torch.cuda.synchronize(); s.record()
    for _ in range(3): evaluate_rollout(model, batches, T, xn, yn, distributed=True, evaluator=ev)
    e.record(); torch.cuda.synchronize()
    ms = torch.tensor([s.elapsed_time(e) / 3], device="cuda")
    if world > 1: dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    out[wname] = {"traj_shape": list(shape), "batches": NB, "loss_single": single, "loss_sharded": sharded,
                  "ms_per_sweep_max_over_ranks": round(ms.item(), 3), "rollouts_per_s": round(NB * shape[0] / ms.item() * 1e3, 1)}
if rank == 0:
    print(json.dumps({"n_gpus": world, "rollout_steps": T, **out}))
if world > 1:
    dist.destroy_process_group()
