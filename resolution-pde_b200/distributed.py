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

    def pack(self) -> torch.Tensor:
        off = 0
        for p, n in zip(self.params, self.sizes):
            if p.grad is None:
                self.flat[off:off + n].zero_()
            else:
                self.flat[off:off + n].copy_(self._views([p.grad])[0])
            off += n
        return self.flat

    def unpack(self, scale: float = 1.0) -> None:
        off = 0
        for p, n in zip(self.params, self.sizes):
            src = self.flat[off:off + n]
            if p.grad is None:
                p.grad = torch.zeros_like(p)
            dst = self._views([p.grad])[0]
            if scale != 1.0:
                torch.mul(src, scale, out=dst)
            else:
                dst.copy_(src)
            off += n

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
