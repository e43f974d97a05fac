"""Graph-captured training step for the spectral path.

The reference's loop (train/training.py:19-88) launches every op eagerly, rebuilds the coordinate grid on the
host each forward (fno.py:51-54, ffno.py:199-208) and synchronises on loss.item() every step (training.py:47).
On the small 1-D configs the HBM floor of a step is tens of microseconds, so launch latency dominates.  Here the
whole `zero_grad -> forward -> RelativeL2 -> backward` sequence of one (model, batch shape) is captured once into a
CUDA graph and replayed: the C-ABI entry points never synchronise or allocate, workspaces come from torch's
capture-aware caching allocator, DFT tables and grids are cached device tensors created during the warm-up.

One graph per input shape, as ResolutionGroupedDataLoader (train/mres_training.py:108-128) yields one
resolution per batch.  Gradients land in `p.grad` of the model's parameters (static tensors owned by the graph's
memory pool), so any torch optimizer and the flat NCCL all-reduce (distributed.GradBucket) work unchanged.
"""
from __future__ import annotations

from typing import Callable, Dict, Optional, Tuple

import torch

from . import ops


class GraphedTrainStep:
    """step(x, y) -> loss (0-dim device tensor).  x, y may live on the device or in pinned host memory."""

    def __init__(self, model: torch.nn.Module, loss_fn: Callable, warmup: int = 2, optimizer=None, bucket=None,
                 max_graphs: int = 16):
        """optimizer: an optional `optim.FusedAdamW`; its update is then captured too (a second graph replayed right
        after the first: its device table needs the gradient pointers the first capture creates) and `step(x, y)`
        becomes zero_grad + forward + loss + backward + AdamW with no host work besides the two replays.
        bucket: an optional `distributed.GradBucket`; when a process group with more than one rank is initialised its
        pack -> NCCL all-reduce -> scaled unpack is captured INTO the step's graph right after the backward pass, so the
        data-parallel step is still one replay with no eager tail (the collective is enqueued by the graph on the
        capture stream; every rank replays the same sequence).
        max_graphs: ragged tail batches of a resolution-grouped epoch (train/mres_training.py:108-128) create one extra
        graph per (shape, batch) pair; the least recently used graph is dropped beyond this bound."""
        self.model = model
        self.loss_fn = loss_fn
        self.warmup = warmup
        self.optimizer = optimizer
        self.bucket = bucket
        self.max_graphs = max_graphs
        self._graphs: Dict[Tuple, Tuple] = {}

    def _reduce(self):
        if self.bucket is not None:
            self.bucket.allreduce_mean()

    def _capture(self, x: torch.Tensor, y: torch.Tensor):
        dev = next(self.model.parameters()).device
        sx = torch.empty(x.shape, dtype=torch.float32, device=dev)
        sy = torch.empty(y.shape, dtype=torch.float32, device=dev)
        sx.copy_(x)
        sy.copy_(y)
        side = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side):          # warm-up off the capture: builds tables/grids, sizes the allocator
            for _ in range(self.warmup):
                self.model.zero_grad(set_to_none=True)
                self.loss_fn(self.model(sx), sy).backward()
                self._reduce()
        torch.cuda.current_stream(dev).wait_stream(side)
        torch.cuda.synchronize(dev)
        self.model.zero_grad(set_to_none=True)  # grads are (re)created inside the capture, in the graph's pool
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            loss = self.loss_fn(self.model(sx), sy)
            with ops.pgrad_side_stream(dev):       # parameter-gradient-only kernels on a second captured stream (ops.py)
                loss.backward()
            self._reduce()
        # the gradient tensors written by this graph (its pool owns them); re-attached to the parameters after a replay
        params = [q for q in self.model.parameters() if q.grad is not None]
        grads = [q.grad for q in params]
        g_opt = None
        if self.optimizer is not None:
            self.optimizer.prepare()            # device tables for exactly these gradient tensors (H2D, outside capture)
            self.optimizer.sync_lr()
            g_opt = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g_opt, pool=g.pool()):
                self.optimizer.launch()
        return g, sx, sy, loss, params, grads, g_opt

    def __call__(self, x: torch.Tensor, y: torch.Tensor) -> torch.Tensor:
        key = (tuple(x.shape), tuple(y.shape))
        ent = self._graphs.pop(key, None)
        if ent is None:
            while len(self._graphs) >= self.max_graphs:        # bounded cache: drop the least recently used graph
                self._graphs.pop(next(iter(self._graphs)))
            ent = self._capture(x, y)
        self._graphs[key] = ent                                 # (re)insert as most recently used
        g, sx, sy, loss, params, grads, g_opt = ent
        if x.data_ptr() != sx.data_ptr():
            sx.copy_(x, non_blocking=True)
        if y.data_ptr() != sy.data_ptr():
            sy.copy_(y, non_blocking=True)
        g.replay()
        if params and (params[0].grad is not grads[0] or params[-1].grad is not grads[-1]):
            for q, gr in zip(params, grads):   # only after a zero_grad(set_to_none=True) or a switch between graphs
                q.grad = gr
        if g_opt is not None:
            self.optimizer.sync_lr()
            g_opt.replay()
        return loss

    def release(self) -> None:
        """Drops every captured graph (and the memory pools they own).  With a `bucket` the graphs hold captured NCCL
        kernels: call this BEFORE torch.distributed.destroy_process_group() -- tearing the communicator down under a
        live graph blocks (measured on B200 / NCCL 2.28, tools/dp_check.py)."""
        self._graphs.clear()
        import gc
        gc.collect()
        if torch.cuda.is_available():
            torch.cuda.synchronize()

    def static_inputs(self, x_shape, y_shape) -> Optional[Tuple[torch.Tensor, torch.Tensor]]:
        ent = self._graphs.get((tuple(x_shape), tuple(y_shape)))
        return (ent[1], ent[2]) if ent else None


class ResolutionGroupedBatches:
    """Device-resident counterpart of the reference's ResolutionGroupedDataLoader (train/mres_training.py:75-131; SURVEY
    section 8f rank 1).  Same grouping rule (the samples' LAST dimension, :87-90), same batches (consecutive runs of
    `batch_size` inside a group, ragged tail kept), same order (groups in order of first appearance, then
    `random.shuffle` of the batch list when shuffle=True, :108-128) -- but every group is stacked ONCE into one device
    tensor, so an epoch does no per-step collation or H2D copy: a batch is a view of the group's tensor.

    rank / world: the data-parallel split of SURVEY section 8e -- every rank iterates the SAME batch sequence (same resolution
    in the same step) and takes its `shard_batch` slice of each batch; `weights()` gives the per-batch (local, global)
    sample counts for the weighted gradient all-reduce of ragged shards."""

    def __init__(self, dataset, batch_size: int, shuffle: bool = True, device=None, rank: int = 0, world: int = 1):
        from collections import OrderedDict
        from .distributed import shard_batch
        self.batch_size, self.shuffle, self.rank, self.world = batch_size, shuffle, rank, world
        groups = OrderedDict()
        for i in range(len(dataset)):
            x, y = dataset[i]
            groups.setdefault(int(x.shape[-1]), []).append((x, y))
        self.groups = OrderedDict()
        for res, items in groups.items():
            gx = torch.stack([torch.as_tensor(a) for a, _ in items]).to(device=device, dtype=torch.float32)
            gy = torch.stack([torch.as_tensor(b) for _, b in items]).to(device=device, dtype=torch.float32)
            self.groups[res] = (gx, gy)
        self._shard = shard_batch

    def _plan(self):
        import random
        plan = []
        for res, (gx, _) in self.groups.items():
            n = gx.shape[0]
            perm = torch.randperm(n).tolist() if self.shuffle else None     # DataLoader(shuffle=True) inside each group
            for i in range(0, n, self.batch_size):
                plan.append((res, i, min(n, i + self.batch_size), perm))
        if self.shuffle:
            random.shuffle(plan)
        return plan

    def __iter__(self):
        for res, lo, hi, perm in self._plan():
            gx, gy = self.groups[res]
            a, b = self._shard(hi - lo, self.rank, self.world)
            if perm is None:
                yield gx[lo + a:lo + b], gy[lo + a:lo + b]
            else:
                idx = torch.as_tensor(perm[lo + a:lo + b], device=gx.device, dtype=torch.long)
                yield gx.index_select(0, idx), gy.index_select(0, idx)

    def weights(self):
        """(local, global) sample counts of every batch of one UNSHUFFLED pass, for GradBucket.allreduce_mean(weight=...)"""
        out = []
        for res, (gx, _) in self.groups.items():
            n = gx.shape[0]
            for i in range(0, n, self.batch_size):
                g = min(n, i + self.batch_size) - i
                a, b = self._shard(g, self.rank, self.world)
                out.append((b - a, g))
        return out

    def __len__(self):
        return sum((gx.shape[0] + self.batch_size - 1) // self.batch_size for gx, _ in self.groups.values())
