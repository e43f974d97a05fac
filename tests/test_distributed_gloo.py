"""Host-side data-parallel logic on CPU (gloo, world_size 2): batch sharding inside a resolution group and the flat
gradient bucket reproduce the single-process gradient of the global batch, including complex64 parameters and
ragged shards.  The models here are plain torch stand-ins: the CUDA path is not involved (no GPU in this container)."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


class _Toy(torch.nn.Module):
    """real + complex parameters, per-sample loss averaged over the batch (as RelativeL2Loss does)"""

    def __init__(self):
        super().__init__()
        g = torch.Generator().manual_seed(3)
        self.w = torch.nn.Parameter(torch.randn(8, 8, generator=g))
        self.c = torch.nn.Parameter(torch.complex(torch.randn(5, generator=g), torch.randn(5, generator=g)))
        self.unused = torch.nn.Parameter(torch.zeros(3))

    def forward(self, x):
        f = torch.fft.rfft(x @ self.w, dim=-1)
        return torch.fft.irfft(f * self.c, n=8, dim=-1)


def _per_sample(model, x, y):
    out = model(x)
    return ((out - y).flatten(1).norm(dim=1) / (y.flatten(1).norm(dim=1) + 1e-8))


def _worker(rank, world, port, batch, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from resolution_pde_b200.distributed import GradBucket, shard_batch
    torch.manual_seed(0)
    model = _Toy()
    g = torch.Generator().manual_seed(11)
    x, y = torch.randn(batch, 8, generator=g), torch.randn(batch, 8, generator=g)
    lo, hi = shard_batch(batch, rank, world)
    bucket = GradBucket(model.parameters())
    if hi > lo:
        _per_sample(model, x[lo:hi], y[lo:hi]).mean().backward()
    bucket.allreduce_mean(weight=float(hi - lo), total_weight=float(batch))
    out = {k: (torch.view_as_real(p.grad).clone() if p.grad.is_complex() else p.grad.clone()) for k, p in model.named_parameters()}
    if rank == 0:
        q.put({k: v.numpy() for k, v in out.items()})
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("batch", [8, 5, 1])
def test_sharded_grads_equal_global_batch(batch):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, batch, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = q.get(timeout=120)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    torch.manual_seed(0)
    model = _Toy()
    g = torch.Generator().manual_seed(11)
    x, y = torch.randn(batch, 8, generator=g), torch.randn(batch, 8, generator=g)
    _per_sample(model, x, y).mean().backward()
    for k, p in model.named_parameters():
        want = torch.zeros_like(p) if p.grad is None else p.grad
        want = torch.view_as_real(want) if want.is_complex() else want
        assert torch.allclose(torch.from_numpy(got[k]), want, rtol=1e-5, atol=1e-6), k


def test_shard_batch_partitions():
    from resolution_pde_b200.distributed import shard_batch
    for world in (1, 2, 4, 8):
        for b in (1, 3, 8, 16, 17):
            spans = [shard_batch(b, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == b
            assert all(spans[i][1] == spans[i + 1][0] for i in range(world - 1))
            sizes = [hi - lo for lo, hi in spans]
            assert max(sizes) - min(sizes) <= 1


# ---------------------------------------------------------------------------------------------------------
# rollout evaluation sharded over ranks: batches r, r + world, ... per rank, one all-reduce of (sum, count)
# ---------------------------------------------------------------------------------------------------------
def _rollout_batches():
    g = torch.Generator().manual_seed(5)
    shapes = [(3, 6, 16), (3, 6, 16), (2, 1, 16), (3, 4, 16), (1, 6, 16), (2, 6, 16)]      # one batch too short to roll out
    return [torch.randn(*s, generator=g) for s in shapes]


def _rollout_batch_loss(batch):
    from oracle import rollout_port as rp                       # the CPU checker stands in for the device evaluator
    return rp.rollout_batch_loss(lambda x: 0.9 * x.roll(1, -1), batch, 4, (0.1, 1.2, 1e-8), (0.2, 0.8, 1e-8))[0]


def _rollout_worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from resolution_pde_b200.rollout import mean_over_batches
    seen = []

    def loss(b):
        seen.append(tuple(b.shape))
        return _rollout_batch_loss(b)

    got = mean_over_batches(_rollout_batches(), loss, "cpu", distributed=True)
    q.put((rank, got, len(seen)))
    dist.barrier()
    dist.destroy_process_group()


def test_rollout_evaluation_sharded_over_ranks_matches_single_process():
    from resolution_pde_b200.rollout import mean_over_batches
    want = mean_over_batches(_rollout_batches(), _rollout_batch_loss, "cpu")
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_rollout_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in range(2))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert [r[2] for r in res] == [3, 2]                         # 5 valid batches: ranks take 0,2,4 and 1,3
    for _, got, _ in res:
        assert abs(got - want) < 1e-6 * abs(want)
