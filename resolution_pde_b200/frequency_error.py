"""Frequency-resolved error metrics on the device.

The reference isolates one Fourier mode (1-D) or one radial band (2-D) at a time, inverse-transforms it and takes the
spatial L2 norm: `num_modes` (or `num_radial_bins`) inverse FFTs of the whole batch and two `.item()` syncs each
(`utils/frequency_error.py:37-91`, `93-161`).  By Parseval the spatial norm of a band is a weighted sum of |c|^2 over
its modes, so one analysis pass (the truncated-DFT contraction of the spectral convolutions: `rpde_dft_table` kinds 0 / 2
through `rpde_matmul`) gives the power spectrum of (y_hat - y) and of y, and every mode / band follows from it:

    1-D   || irfft(c_k e_k) ||^2 = w_k |c_k|^2 / H,   w_k = 2, or 1 with the imaginary part dropped for DC / Nyquist
    2-D   || irfft2(band) ||^2   = sum_{(ky,kx) in band} w_kx |c|^2 / (H W)   (bands are symmetric in ky)

Same function names, arguments and return values (numpy arrays) as the reference; fp32 CUDA inputs, one sync at the end.
"""
from __future__ import annotations

import numpy as np
import torch

from . import _lib, ops

_masks = {}


def _check(a: torch.Tensor, b: torch.Tensor):
    if not (a.is_cuda and b.is_cuda) or a.dtype != torch.float32 or b.dtype != torch.float32 or a.shape != b.shape:
        raise ValueError("frequency_error: y_hat and y must be fp32 CUDA tensors of one shape (no CPU fallback)")


def decompose_error_by_frequency_1d(y_hat: torch.Tensor, y: torch.Tensor, num_modes=None):
    """(error_per_mode, solution_magnitude_per_mode, frequencies) of (B, C, H) fields (frequency_error.py:37-91)."""
    _check(y_hat, y)
    H = y.shape[-1]
    n_freq = H // 2 + 1
    k = n_freq if num_modes is None else min(int(num_modes), n_freq)
    fa = ops.dft_table(0, H, k, 0, 1.0, y.device).view(2 * k, H)
    rows = torch.cat([(y_hat - y).reshape(-1, H), y.reshape(-1, H)], dim=0)        # one contraction for both
    spec = ops.matmul(rows, fa, False, True).view(2, -1, k, 2).double()            # (which, row, mode, re|im)
    edge = torch.zeros(k, dtype=torch.bool, device=y.device)
    edge[0] = True
    if H % 2 == 0 and k == n_freq:
        edge[-1] = True
    re2, im2 = spec[..., 0].pow(2).sum(1), spec[..., 1].pow(2).sum(1)              # (2, k)
    power = torch.where(edge, re2, 2.0 * (re2 + im2)) / H
    out = power.sqrt().cpu().numpy()
    freqs = torch.fft.rfftfreq(H).numpy()                                          # float32, as the reference returns
    return out[0], out[1], freqs[:k]


def _radial_bins(H: int, W: int, bins: int, device):
    """bin index of every (ky, kx) point of the half spectrum, -1 where the reference's masks select nothing"""
    key = (H, W, bins, str(device))
    m = _masks.get(key)
    if m is None:
        # the reference's own comparisons: float32 radial frequencies against the float64 bin edges (frequency_error.py:115-134)
        fy = torch.fft.fftfreq(H).view(-1, 1)
        fx = torch.fft.rfftfreq(W).view(1, -1)
        radial = torch.sqrt(fy ** 2 + fx ** 2)
        edges = np.linspace(0, 0.5, bins + 1)
        idx = torch.full(radial.shape, -1, dtype=torch.int32)
        for i in range(bins):
            idx[(radial >= edges[i]) & (radial < edges[i + 1])] = i
        m = idx.reshape(-1).contiguous().to(device)
        _masks[key] = m
    return m


def decompose_error_by_frequency_2d(y_hat: torch.Tensor, y: torch.Tensor, num_radial_bins: int = 64):
    """(error_per_bin, solution_magnitude_per_bin, radial_freqs) of (B, C, H, W) fields (frequency_error.py:93-161).
    Even H (the C2C row table keeps [0, H/2) U [-H/2, 0) = every row)."""
    _check(y_hat, y)
    H, W = y.shape[-2], y.shape[-1]
    if H & 1:
        raise NotImplementedError("decompose_error_by_frequency_2d: odd heights are not implemented")
    kx, m = W // 2 + 1, H // 2
    dev = y.device
    x3 = torch.cat([(y_hat - y).reshape(-1, H, W), y.reshape(-1, H, W)], dim=0)
    B2 = x3.shape[0]
    fa_x = ops.dft_table(0, W, kx, 0, 1.0, dev).view(2 * kx, W)
    g_an = ops.dft_table(2, H, 2 * m, m, 1.0, dev).view(4 * m, 2 * H)
    s1 = ops.matmul(x3.view(B2 * H, W), fa_x, False, True)                          # ((b, h), (k, ri))
    s1 = s1.view(B2, H, kx, 2).permute(3, 1, 0, 2).reshape(2 * H, B2 * kx)          # ((ri, h), (b, k))
    s2 = ops.matmul(g_an, s1).view(H, 2, 2, B2 // 2, kx).double()                   # (row j = ky, re|im, which, b, kx)
    power = s2.pow(2).sum(dim=(1, 3))                                               # (ky, which, kx): sum over re|im and b
    w = torch.full((kx,), 2.0, dtype=torch.float64, device=dev)
    w[0] = 1.0
    if W % 2 == 0:
        w[-1] = 1.0
    power = (power * w / (H * W)).permute(1, 0, 2).reshape(2, H * kx)               # (which, (ky, kx)); row j is ky in fft order
    bins = _radial_bins(H, W, int(num_radial_bins), dev)                             # (H*kx,) band index of every mode
    power = power.contiguous()
    sums = torch.empty((2, int(num_radial_bins)), dtype=torch.float64, device=dev)
    with ops._Ctx(power) as c:                                                       # band totals: one library kernel (fp64)
        _lib.check(c.lib.rpde_binned_sums_f64(power.data_ptr(), bins.data_ptr(), sums.data_ptr(), 2, int(num_radial_bins),
                                              power.shape[1], c.stream), "rpde_binned_sums_f64")
    out = sums.sqrt().cpu().numpy()
    edges = np.linspace(0, 0.5, int(num_radial_bins) + 1)
    return out[0], out[1], (edges[:-1] + edges[1:]) / 2
