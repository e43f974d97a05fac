"""Spectral data-preparation operators on the device.

The reference low-pass filters and resamples its datasets with full FFTs on the host or with torch.fft
(`utils/low_pass_filter.py:3-100`, `utils/res_utils.py:29-50, 93-125`).  All four operators below keep only a block of
low modes, so each is the same pair of truncated-DFT contractions the spectral convolutions use: an analysis table
(`rpde_dft_table` kind 0 / 2) and a synthesis table (kind 1 / 3) applied through the contraction engine
(`rpde_matmul`), with the retained modes as the only intermediate.  Function names, argument meaning and output
shapes follow the reference; inputs must be CUDA fp32 tensors (no CPU fallback).

    lowpass_filter_1d(data, cutoff_ratio)      utils/low_pass_filter.py:3-40
    lowpass_filter_2d(data, cutoff_ratio)      utils/low_pass_filter.py:42-100
    resize_1d(x, out_size)                     utils/res_utils.py:93-125
    resize(x, out_size, permute=False)         utils/res_utils.py:29-50   (even sizes: see resize)
"""
from __future__ import annotations

import torch

from . import ops


def _axis_last(x2d: torch.Tensor, n_out: int, k: int, scale: float) -> torch.Tensor:
    """irfft(rfft(x)[..., :k] zero-padded, n=n_out) * (scale * n_out) along the last axis of a (T, n_in) matrix."""
    n_in = x2d.shape[1]
    if k <= 0:
        return torch.zeros((x2d.shape[0], n_out), dtype=torch.float32, device=x2d.device)
    fa = ops.dft_table(0, n_in, k, 0, 1.0, x2d.device).view(2 * k, n_in)          # rows (k, re|im)
    fs = ops.dft_table(1, n_out, k, 0, scale, x2d.device).view(n_out, 2 * k)      # c_k weights: irfft semantics
    spec = ops.matmul(x2d, fa, False, True)                                        # (T, 2k)
    return ops.matmul(spec, fs, False, True)                                       # (T, n_out)


def _axis_rows(x3d: torch.Tensor, k: int, scale: float) -> torch.Tensor:
    """The same real low-pass along axis 1 of a (B, H, W) tensor (H -> H), through a [H][B*W] view."""
    B, H, W = x3d.shape
    if k <= 0:
        return torch.zeros_like(x3d)
    fa = ops.dft_table(0, H, k, 0, 1.0, x3d.device).view(2 * k, H)
    fs = ops.dft_table(1, H, k, 0, scale, x3d.device).view(H, 2 * k)
    xt = x3d.permute(1, 0, 2).reshape(H, B * W)
    spec = ops.matmul(fa, xt)                                                      # (2k, B*W)
    y = ops.matmul(fs, spec)                                                       # (H, B*W)
    return y.view(H, B, W).permute(1, 0, 2).contiguous()


def _check(x: torch.Tensor):
    if not x.is_cuda or x.dtype != torch.float32:
        raise ValueError("spectral_ops: inputs must be fp32 CUDA tensors (no CPU fallback)")


def lowpass_filter_1d(data: torch.Tensor, cutoff_ratio: float = 0.25) -> torch.Tensor:
    """Zero the rfft modes >= int((N/2+1) * cutoff_ratio) along the last axis (low_pass_filter.py:3-40).
    data: (batch, time, channels, N) or (batch, time, N); same shape out."""
    _check(data)
    n = data.shape[-1]
    k = int((n // 2 + 1) * cutoff_ratio)                                           # low_pass_filter.py:27-28
    y = _axis_last(data.reshape(-1, n), n, min(k, n // 2 + 1), 1.0 / n)
    return y.view(data.shape)


def lowpass_filter_2d(data: torch.Tensor, cutoff_ratio: float = 0.25) -> torch.Tensor:
    """Rectangular mask |f_y| <= cutoff_ratio/2 and f_x <= cutoff_ratio/2 on rfft2 (low_pass_filter.py:42-100).
    data: (batch, time, channels, S, S) or (batch, time, S, S).  The mask is symmetric in f_y, so the operator is the
    product of two real 1-D low-passes (rows, then columns)."""
    _check(data)
    s_y, s_x = data.shape[-2], data.shape[-1]
    if s_y != s_x:
        raise ValueError("lowpass_filter_2d: the reference assumes height == width")
    cutoff = cutoff_ratio * 0.5
    # the reference compares float32 frequency grids against the python float: count the kept modes the same way
    kx = int((torch.fft.rfftfreq(s_x) <= cutoff).sum())
    fy = torch.fft.fftfreq(s_y)
    my = torch.abs(fy) <= cutoff
    ky = int(my[: s_y // 2 + 1].sum())        # kept k = 0 .. ky-1 (index S/2 is the Nyquist row, -0.5 in fftfreq)
    x3 = data.reshape(-1, s_y, s_x)
    y = _axis_rows(x3, ky, 1.0 / s_y)
    y = _axis_last(y.reshape(-1, s_x), s_x, kx, 1.0 / s_x)
    return y.view(data.shape)


def resize_1d(x: torch.Tensor, out_size: int) -> torch.Tensor:
    """Fourier interpolation / decimation along the last axis (res_utils.py:93-125)."""
    _check(x)
    n_in = x.shape[-1]
    k = min(n_in // 2 + 1, out_size // 2 + 1)                                      # res_utils.py:113
    y = _axis_last(x.reshape(-1, n_in), out_size, k, 1.0 / n_in)                   # irfft 1/out * (out/in)
    return y.view(*x.shape[:-1], out_size)


def resize(x: torch.Tensor, out_size, permute: bool = False) -> torch.Tensor:
    """2-D Fourier resize of the last two axes (res_utils.py:29-50).  Even input and output sizes (the reference's
    datasets: 64 ... 1024): the kept row block is then [0, m) U [-m, 0) on both sides, which is the C2C row table of the
    FNO2d transform; odd sizes keep one more positive than negative row and are not implemented."""
    _check(x)
    if permute:
        x = x.permute(0, 3, 1, 2)
    h_in, w_in = x.shape[-2], x.shape[-1]
    h_out, w_out = int(out_size[0]), int(out_size[1])
    if (h_in | h_out) & 1:
        raise NotImplementedError("resize: odd heights are not implemented")
    m = min(h_in, h_out) // 2                                                      # top_freqs1 == bot_freqs1 (res_utils.py:35-40)
    kx = min(w_in // 2 + 1, w_out // 2 + 1)
    lead = x.shape[:-2]
    x3 = x.reshape(-1, h_in, w_in).contiguous()
    B = x3.shape[0]
    dev = x.device
    fa_x = ops.dft_table(0, w_in, kx, 0, 1.0, dev).view(2 * kx, w_in)
    g_an = ops.dft_table(2, h_in, 2 * m, m, 1.0, dev).view(4 * m, 2 * h_in)        # rows (j, re|im), cols (re|im, h)
    g_sy = ops.dft_table(3, h_out, 2 * m, m, 1.0, dev).view(2 * h_out, 4 * m)      # rows (h, re|im), cols (j, re|im)
    fs_x = ops.dft_table(1, w_out, kx, 0, 1.0 / (h_in * w_in), dev).view(w_out, 2 * kx)
    s1 = ops.matmul(x3.view(B * h_in, w_in), fa_x, False, True)                    # (B*h_in, (k, ri))
    s1 = s1.view(B, h_in, kx, 2).permute(3, 1, 0, 2).reshape(2 * h_in, B * kx)     # ((ri, h), (b, k))
    s2 = ops.matmul(g_an, s1)                                                      # ((j, ri), (b, k))
    t3 = ops.matmul(g_sy, s2)                                                      # ((h_out, ri), (b, k))
    t3 = t3.view(h_out, 2, B, kx).permute(2, 0, 3, 1).reshape(B * h_out, 2 * kx)   # ((b, h_out), (k, ri))
    y = ops.matmul(t3, fs_x, False, True).view(*lead, h_out, w_out)
    if permute:
        y = y.permute(0, 2, 3, 1)
    return y
