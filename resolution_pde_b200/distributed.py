"""Data-parallel plumbing for the spectral path: one process per GPU, batch sharded within each
resolution group, one flat fp32 all-reduce of the gradients per step (NCCL over NVLink/NVSwitch on the GPU
box, gloo on CPU for the host-logic tests).  Replaces the reference's single-process nn.DataParallel
(main_2d.py:88-94, 146-149); parameters are replicated, nothing but the batch is sharded.
"""
from __future__ import annotations

from typing import Iterable, List, Sequence, Tuple

import torch
import torch.distributed as dist


def shard_batch(batch_size: int, rank: int, world: int) -> Tuple[int, int]:
    """[start, stop) of this rank's samples inside one resolution-group batch.  Every rank works on the same
    resolution in the same step; ragged tails (mres_training.py:87-106 keeps them) are split as evenly as
    possible, so a rank may get an empty slice when batch_size < world."""
    base, extra = divmod(batch_size, world)
    start = rank * base + min(rank, extra)
    return start, start + base + (1 if rank < extra else 0)


class GradBucket:
    """Flat fp32 gradient bucket.  complex64 parameters (FNO spectral weights) are carried as their
    (re, im) float view because NCCL has no complex dtype."""

    def __init__(self, params: Iterable[torch.nn.Parameter]):
        self.params: List[torch.nn.Parameter] = [p for p in params if p.requires_grad]
        self.sizes = [p.numel() * (2 if p.is_complex() else 1) for p in self.params]
        self.total = sum(self.sizes)
        dev = self.params[0].device if self.params else torch.device("cpu")
        self.flat = torch.zeros(self.total, dtype=torch.float32, device=dev)
        self.nbytes = self.total * 4

    def _views(self, tensors: Sequence[torch.Tensor]):
        return [torch.view_as_real(t).reshape(-1) if t.is_complex() else t.reshape(-1) for t in tensors]

    def _flat_views(self):
        views, off = [], 0
        for n in self.sizes:
            views.append(self.flat[off:off + n])
            off += n
        return views

    def pack(self) -> torch.Tensor:
        """One multi-tensor copy of every gradient into the flat buffer (missing grads are zero)."""
        flat_views = self._flat_views()
        have = [(fv, p.grad) for fv, p in zip(flat_views, self.params) if p.grad is not None]
        for fv, p in zip(flat_views, self.params):
            if p.grad is None:
                fv.zero_()
        if have:
            torch._foreach_copy_([fv for fv, _ in have], self._views([g for _, g in have]))
        return self.flat

    def unpack(self, scale: float = 1.0) -> None:
        if scale != 1.0:
            self.flat.mul_(scale)
        for p in self.params:
            if p.grad is None:
                p.grad = torch.zeros_like(p)
        torch._foreach_copy_(self._views([p.grad for p in self.params]), self._flat_views())

    def allreduce_mean(self, group=None, weight: float = 1.0, total_weight: float = None) -> None:
        """grad <- sum_r weight_r * grad_r / sum_r weight_r.  With equal per-rank batches weight = 1; with ragged
        shards pass weight = local batch size so the result equals the single-process mean over the global batch."""
        if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
            return
        self.pack()
        if weight != 1.0:
            self.flat.mul_(weight)
        dist.all_reduce(self.flat, op=dist.ReduceOp.SUM, group=group)
        tw = total_weight if total_weight is not None else float(dist.get_world_size(group)) * weight
        self.unpack(1.0 / tw)
