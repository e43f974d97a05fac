"""CPU: the oracle's restatement of the reference's resolution-grouped batching and training loop against the fixture the
UNMODIFIED reference produced (oracle/make_golden_mres.py): batch sequence (ragged tails 4,4,2 / 4,1 / 3) and the loss
histories of a mixed-resolution FNO2d run; plus the host logic of the device-resident ResolutionGroupedBatches (no GPU
needed: it runs on CPU tensors) including its data-parallel sharding."""
import os

import numpy as np
import torch

from helpers import GOLD

FIX = os.path.join(GOLD, "training", "mres_fno2d.npz")


def load():
    z = np.load(FIX)
    n, nv = int(z["n_samples"]), int(z["n_val"])
    samples = [(torch.from_numpy(z[f"sx/{i}"]), torch.from_numpy(z[f"sy/{i}"])) for i in range(n)]
    val = [(torch.from_numpy(z[f"vx/{i}"]), torch.from_numpy(z[f"vy/{i}"])) for i in range(nv)]
    params = {}
    for k in z.files:
        if k.startswith("param/"):
            t = torch.from_numpy(z[k])
            params[k[6:]] = torch.view_as_complex(t.contiguous()) if "weights" in k else t
    return z, samples, val, params


def test_oracle_grouping_and_training_match_the_reference():
    from oracle import mres_port, training_port
    z, samples, val, params = load()
    bs = int(z["batch_size"])
    batches = mres_port.grouped_batches(samples, bs)
    assert [(int(x.shape[-1]), int(x.shape[0])) for x, _ in batches] == [tuple(r) for r in z["batch_seq"].tolist()]
    vb = mres_port.grouped_batches(val, bs)
    hist, vhist, _ = training_port.train("FNO2d", params, batches, vb, float(z["lr"]), int(z["epochs"]), activation="gelu", n_blocks=4)
    np.testing.assert_allclose(hist, z["loss_history"], rtol=2e-5)
    np.testing.assert_allclose(vhist, z["val_history"], rtol=2e-5)


def test_grouped_batches_class_matches_the_reference_sequence_and_shards():
    from resolution_pde_b200.training import ResolutionGroupedBatches
    z, samples, _, _ = load()
    bs = int(z["batch_size"])
    full = ResolutionGroupedBatches(samples, bs, shuffle=False)
    seq = [(int(x.shape[-1]), int(x.shape[0])) for x, _ in full]
    assert seq == [tuple(r) for r in z["batch_seq"].tolist()] and len(full) == len(seq)
    # two ranks: same batch sequence, the shards tile every batch exactly, weights add up
    parts = [list(ResolutionGroupedBatches(samples, bs, shuffle=False, rank=r, world=2)) for r in range(2)]
    for (fx, fy), (ax, ay), (bx, by) in zip(full, parts[0], parts[1]):
        assert torch.equal(torch.cat([ax, bx]), fx) and torch.equal(torch.cat([ay, by]), fy)
    w0 = ResolutionGroupedBatches(samples, bs, shuffle=False, rank=0, world=2).weights()
    w1 = ResolutionGroupedBatches(samples, bs, shuffle=False, rank=1, world=2).weights()
    assert all(a[0] + b[0] == a[1] == b[1] for a, b in zip(w0, w1))
    # shuffle keeps the multiset of (resolution, batch size) pairs
    torch.manual_seed(0)
    sh = sorted((int(x.shape[-1]), int(x.shape[0])) for x, _ in ResolutionGroupedBatches(samples, bs, shuffle=True))
    assert sh == sorted(seq)
