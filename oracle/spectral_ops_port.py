"""TEST INFRASTRUCTURE ONLY - CPU restatement of the reference's spectral data-preparation operators.

Pinned by tests/golden/specops/*.npz, generated from the unmodified reference by oracle/make_golden_spectral_ops.py
(tests/test_oracle_golden.py checks this file against them).  Nothing in the product imports it.
Each function states the reference lines it follows; evaluated in float64 it is the arbiter of the GPU tests.
"""
from __future__ import annotations

import torch


def lowpass_filter_1d(data: torch.Tensor, cutoff_ratio: float = 0.25) -> torch.Tensor:
    """utils/low_pass_filter.py:3-40: rfft along the last axis, bins >= int(n_bins * ratio) set to zero, irfft(n=N)."""
    spec = torch.fft.rfft(data, dim=-1)
    keep = int(spec.shape[-1] * cutoff_ratio)
    spec = torch.cat([spec[..., :keep], torch.zeros_like(spec[..., keep:])], dim=-1)
    return torch.fft.irfft(spec, n=data.shape[-1], dim=-1)


def lowpass_filter_2d(data: torch.Tensor, cutoff_ratio: float = 0.25) -> torch.Tensor:
    """utils/low_pass_filter.py:42-100: rfft2, rectangular mask |f_y| <= ratio/2 (fftfreq) x f_x <= ratio/2 (rfftfreq),
    both grids in float32 as the reference builds them, irfft2(s=(S, S))."""
    s = data.shape[-1]
    spec = torch.fft.rfft2(data, dim=(-2, -1))
    cut = cutoff_ratio * 0.5
    keep_y = torch.abs(torch.fft.fftfreq(s)) <= cut
    keep_x = torch.abs(torch.fft.rfftfreq(s)) <= cut
    mask = (keep_y[:, None] & keep_x[None, :]).to(spec.real.dtype)
    return torch.fft.irfft2(spec * mask, s=(s, s), dim=(-2, -1))


def resize_1d(x: torch.Tensor, out_size: int) -> torch.Tensor:
    """utils/res_utils.py:93-125: rfft, the first min(n_in/2+1, out/2+1) bins copied into a zero spectrum of out/2+1
    bins, irfft(n=out), times out/in."""
    n_in = x.shape[-1]
    spec = torch.fft.rfft(x, dim=-1)
    out_bins = out_size // 2 + 1
    k = min(spec.shape[-1], out_bins)
    padded = torch.zeros(*x.shape[:-1], out_bins, dtype=spec.dtype)
    padded[..., :k] = spec[..., :k]
    return torch.fft.irfft(padded, n=out_size, dim=-1) * (out_size / n_in)


def resize(x: torch.Tensor, out_size, permute: bool = False) -> torch.Tensor:
    """utils/res_utils.py:29-50: rfft2; rows [0, top) and the last `bot` rows, columns [0, kx) copied into a zero
    spectrum of the output size; irfft2(s=out); times (H_out/H_in)(W_out/W_in)."""
    if permute:
        x = x.permute(0, 3, 1, 2)
    h_in, w_in = x.shape[-2], x.shape[-1]
    h_out, w_out = out_size
    spec = torch.fft.rfft2(x)
    top = min((h_in + 1) // 2, (h_out + 1) // 2)
    bot = min(h_in // 2, h_out // 2)
    kx = min(spec.shape[-1], w_out // 2 + 1)
    padded = torch.zeros(*x.shape[:-2], h_out, w_out // 2 + 1, dtype=spec.dtype)
    padded[..., :top, :kx] = spec[..., :top, :kx]
    if bot > 0:
        padded[..., h_out - bot:, :kx] = spec[..., h_in - bot:, :kx]
    y = torch.fft.irfft2(padded, s=(h_out, w_out)) * (h_out / h_in) * (w_out / w_in)
    if permute:
        y = y.permute(0, 2, 3, 1)
    return y


def _select(n_old: int, n: int) -> torch.Tensor:
    """utils/res_utils.py:21-22, 79-82: fftfreq(n_old, d=1/n_old) within [-n/2, n/2 - 1]"""
    freqs = torch.fft.fftfreq(n_old, d=1.0 / n_old, dtype=torch.float64)
    return (freqs >= -n / 2) & (freqs <= n / 2 - 1)


def downsample_1d(u: torch.Tensor, n: int) -> torch.Tensor:
    """utils/res_utils.py:62-91: complex fft (norm='forward') along the last axis, the modes with -n/2 <= k <= n/2 - 1
    kept in fft order, ifft (norm='forward') of that shorter spectrum, real part.  Output length = number of kept modes
    (n for even n < n_old, n - 1 for odd n, n_old when nothing is cut)."""
    sel = _select(u.shape[-1], n)
    spec = torch.fft.fft(u, dim=-1, norm="forward")[..., sel]
    return torch.fft.ifft(spec, dim=-1, norm="forward").real


def downsample(u: torch.Tensor, n: int) -> torch.Tensor:
    """utils/res_utils.py:19-27: the same on both axes of (B, C, S, S) fields: fft2 (norm='forward'), rows and
    columns with -n/2 <= k <= n/2 - 1, ifft2 (norm='forward'), real part."""
    sel = _select(u.shape[-2], n)
    spec = torch.fft.fft2(u, norm="forward")[:, :, sel, :][:, :, :, sel]
    return torch.fft.ifft2(spec, norm="forward").real


def decompose_error_by_frequency_1d(y_hat: torch.Tensor, y: torch.Tensor, num_modes=None):
    """utils/frequency_error.py:37-91: for every retained rfft bin, keep only that bin of y_hat and of y, irfft(n=H), and
    take the L2 norm over the whole batch of the difference and of the target part."""
    import numpy as np
    H = y.shape[-1]
    fh, ft = torch.fft.rfft(y_hat, dim=-1), torch.fft.rfft(y, dim=-1)
    k = ft.shape[-1] if num_modes is None else min(num_modes, ft.shape[-1])
    err, mag = np.zeros(k), np.zeros(k)
    for i in range(k):
        sel = torch.zeros(ft.shape[-1], dtype=ft.real.dtype)
        sel[i] = 1
        sh = torch.fft.irfft(fh * sel, n=H, dim=-1)
        st = torch.fft.irfft(ft * sel, n=H, dim=-1)
        err[i] = torch.norm(sh - st).item()
        mag[i] = torch.norm(st).item()
    return err, mag, torch.fft.rfftfreq(H).numpy()[:k]


def decompose_error_by_frequency_2d(y_hat: torch.Tensor, y: torch.Tensor, num_radial_bins: int = 64):
    """utils/frequency_error.py:93-161: radial bands [e_i, e_{i+1}) of sqrt(fy^2 + fx^2) (float32 grids, float64 edges from
    linspace(0, 0.5, bins + 1)); each band of rfft2 is inverse-transformed and normed; empty bands give 0."""
    import numpy as np
    H, W = y.shape[-2], y.shape[-1]
    fh, ft = torch.fft.rfft2(y_hat, dim=(-2, -1)), torch.fft.rfft2(y, dim=(-2, -1))
    radial = torch.sqrt(torch.fft.fftfreq(H).view(-1, 1) ** 2 + torch.fft.rfftfreq(W).view(1, -1) ** 2)
    edges = np.linspace(0, 0.5, num_radial_bins + 1)
    err, mag = np.zeros(num_radial_bins), np.zeros(num_radial_bins)
    for i in range(num_radial_bins):
        band = (radial >= edges[i]) & (radial < edges[i + 1])
        if band.sum() == 0:
            continue
        sh = torch.fft.irfft2(fh * band, s=(H, W), dim=(-2, -1))
        st = torch.fft.irfft2(ft * band, s=(H, W), dim=(-2, -1))
        err[i] = torch.norm(sh - st).item()
        mag[i] = torch.norm(st).item()
    return err, mag, (edges[:-1] + edges[1:]) / 2


def global_mean_std(samples):
    """dataloaders/ks_naive_markov.py:377-390 (= ns_naive_true_multires.py:513-525), transcribed: every value of every sample
    goes through a Python list into one float32 tensor; mean() and the unbiased std() of that tensor.  (Inline code of a
    dataloader factory that needs h5py files: not importable, hence transcribed rather than called.)"""
    values = []
    for t in samples:
        values.extend(t.flatten().tolist())
    v = torch.tensor(values)
    return float(v.mean()), float(v.std())


def stride_downsample(x, factor: int, spatial_dims: int = 2, channels_last: bool = False):
    """dataloaders/ns_naive_true_multires.py:363-366: data[:, :, ::f, ::f, :] (and x[..., ::f] in the 1-D loaders)"""
    sl = [slice(None)] * x.dim()
    last = x.dim() - (1 if channels_last else 0)
    for d in range(last - spatial_dims, last):
        sl[d] = slice(None, None, factor)
    return x[tuple(sl)]
