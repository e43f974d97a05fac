"""CPU restatement of the reference's resolution-grouped batching -- TEST INFRASTRUCTURE (only tests/ and bench.py's CPU arm
may import it).  Follows train/mres_training.py:75-131 (ResolutionGroupedDataLoader): samples are grouped by the size of
their last dimension (:87-90), every group is cut into consecutive batches of `batch_size` with the ragged tail kept
(DataLoader default drop_last=False, :101-106), groups are visited in order of first appearance (:110-120); with
shuffle=False the batch sequence is deterministic.  Pinned by tests/golden/training/mres_fno2d.npz, whose batch sequence
and loss histories were produced by the UNMODIFIED reference loader and train() (oracle/make_golden_mres.py)."""
from __future__ import annotations

from collections import OrderedDict
from typing import List, Sequence, Tuple

import torch


def grouped_batches(samples: Sequence[Tuple[torch.Tensor, torch.Tensor]], batch_size: int) -> List[Tuple[torch.Tensor, torch.Tensor]]:
    groups = OrderedDict()
    for x, y in samples:
        groups.setdefault(int(x.shape[-1]), []).append((x, y))
    out = []
    for items in groups.values():
        for i in range(0, len(items), batch_size):
            chunk = items[i:i + batch_size]
            out.append((torch.stack([a for a, _ in chunk]), torch.stack([b for _, b in chunk])))
    return out
