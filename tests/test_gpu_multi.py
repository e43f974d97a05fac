"""GPU (-m gpu), needs >= 2 devices (skipped otherwise; run with `gpurun --gpus 2`): the data-parallel CUDA path.

1. Two NCCL ranks, real drop-in modules (FFNO2D at the C3 hyper-parameters, FNO1d at C2): each rank computes the
   gradients of its half of the batch through GraphedTrainStep with the all-reduce captured in the step graph
   (distributed.GradBucket.allreduce_mean); the result must equal the one-rank gradient at the global batch (1e-6
   relative; RelativeL2Loss averages over the batch, so the mean of the two half-batch gradients IS the full-batch one).
   (SURVEY section 4 last bullet, section 8c item 3, section 8e.)
2. nn.DataParallel (the reference's own multi-GPU path, main_2d.py:146-149) over R.FFNO2D under the tcgen05 engine: one Python
   thread per device, replicated non-leaf parameters, per-device kernel attributes -- outputs and gradients must match
   the single-device step.
"""
import os
import socket

import pytest
import torch

from helpers import rel

pytestmark = pytest.mark.gpu
NEED2 = pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two CUDA devices")

CASES = {
    "ffno2d": dict(kind="FFNO2D", ctor=dict(in_channels=1, out_channels=1, width=64, n_layers=2, n_modes=16, factor=4, ff_weight_norm=True,
                                            n_ff_layers=3, layer_norm=True, dropout=0.0), shape=(4, 1, 64, 64)),
    "fno1d": dict(kind="FNO1d", ctor=dict(in_channels=1, out_channels=1, modes=16, width=64), shape=(8, 1, 1024)),
}


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, case, engine, out):
    import torch.distributed as dist
    import resolution_pde_b200 as R
    from resolution_pde_b200.distributed import GradBucket, shard_batch
    from resolution_pde_b200.training import GraphedTrainStep
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    try:
        c = CASES[case]
        R.set_precision(engine)
        torch.manual_seed(0)
        model = getattr(R, c["kind"])(**c["ctor"]).to(dev)
        g = torch.Generator().manual_seed(1)
        x, y = torch.randn(*c["shape"], generator=g), torch.randn(*c["shape"], generator=g)
        lo, hi = shard_batch(c["shape"][0], rank, world)
        bucket = GradBucket(model.parameters())
        step = GraphedTrainStep(model, R.RelativeL2Loss(), bucket=bucket)
        for _ in range(2):                                   # second call = pure graph replay (collective included)
            step(x[lo:hi].to(dev), y[lo:hi].to(dev))
        torch.cuda.synchronize()
        got = {k: p.grad.detach().cpu().clone() for k, p in model.named_parameters()}
        if rank == 0:
            model.zero_grad(set_to_none=True)
            R.RelativeL2Loss()(model(x.to(dev)), y.to(dev)).backward()      # one rank, global batch, eager
            torch.cuda.synchronize()
            want = {k: p.grad.detach().cpu().clone() for k, p in model.named_parameters()}
            out.put({k: rel(got[k], want[k]) for k in want if want[k].abs().max() > 1e-8})
        step.release()                                       # captured NCCL kernels must go before the communicator
        dist.barrier()
    finally:
        dist.destroy_process_group()


@NEED2
@pytest.mark.parametrize("case,engine", [("ffno2d", "3xtf32"), ("fno1d", "3xtf32"), ("ffno2d", "bf16")])
def test_two_rank_gradients_equal_one_rank_global_batch(case, engine):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, case, engine, out)) for r in range(2)]
    for p in procs:
        p.start()
    try:
        errs = out.get(timeout=120)
        for p in procs:
            p.join(timeout=60)
            assert p.exitcode == 0
    finally:
        for p in procs:
            if p.is_alive():
                p.kill()
    # the fp32-accurate engine is deterministic up to summation order; the bf16 variant rounds the half batches differently
    tol = 1e-5 if engine == "3xtf32" else 2e-2
    bad = {k: e for k, e in errs.items() if e > tol}
    assert not bad, bad


@NEED2
@pytest.mark.parametrize("engine", ["3xtf32", "bf16"])
def test_dataparallel_matches_single_device(engine):
    import resolution_pde_b200 as R
    c = CASES["ffno2d"]
    torch.manual_seed(0)
    model = getattr(R, c["kind"])(**c["ctor"]).to("cuda:0")
    x, y = torch.randn(*c["shape"], device="cuda:0"), torch.randn(*c["shape"], device="cuda:0")
    R.set_precision(engine)
    try:
        model.zero_grad(set_to_none=True)
        out1 = model(x)
        R.RelativeL2Loss()(out1, y).backward()
        want = {k: p.grad.detach().clone() for k, p in model.named_parameters()}
        dp = torch.nn.DataParallel(model, device_ids=[0, 1])       # main_2d.py:146-149
        model.zero_grad(set_to_none=True)
        out2 = dp(x)
        R.RelativeL2Loss()(out2, y).backward()
        torch.cuda.synchronize()
    finally:
        R.set_precision("fp32")
    tol = 1e-5 if engine == "3xtf32" else 2e-2
    assert rel(out2, out1) < tol, rel(out2, out1)
    for k, p in model.named_parameters():
        if want[k].abs().max() > 1e-8:
            assert rel(p.grad, want[k]) < tol, (k, rel(p.grad, want[k]))
