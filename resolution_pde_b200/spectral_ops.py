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
    resize(x, out_size, permute=False)         utils/res_utils.py:29-50
    downsample_1d(u, N), downsample(u, N)      utils/res_utils.py:62-91, 19-27 (scipy on numpy there; CUDA tensors here)
    stride_downsample(x, factor, dims)         dataloaders/ns_naive_true_multires.py:363-366 (data[:, :, ::f, ::f, :])
    global_mean_std(tensors)                   dataloaders/ks_naive_markov.py:377-390, ns_naive_true_multires.py:513-525
"""
from __future__ import annotations

import torch

from . import _lib, ops


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


def _axis_rows_to(x3d: torch.Tensor, h_out: int, k: int, scale: float) -> torch.Tensor:
    """_axis_last along axis 1 of a (B, H, W) tensor: H -> h_out."""
    B, H, W = x3d.shape
    fa = ops.dft_table(0, H, k, 0, 1.0, x3d.device).view(2 * k, H)
    fs = ops.dft_table(1, h_out, k, 0, scale, x3d.device).view(h_out, 2 * k)
    xt = x3d.permute(1, 0, 2).reshape(H, B * W)
    y = ops.matmul(fs, ops.matmul(fa, xt))                                         # (h_out, B*W)
    return y.view(h_out, B, W).permute(1, 0, 2).contiguous()


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


_row_tables = {}


def _row_table(kind: int, h: int, top: int, bot: int, device) -> torch.Tensor:
    """C2C row table (rpde_dft_table kind 2 analysis / 3 synthesis) over the rows [0, top) U [h - bot, h), top - bot in {0, 1}.
    The library table keeps a symmetric block [0, top) U [h - top, h); for odd sizes the reference keeps one more positive
    than negative row (res_utils.py:35-40), so the first negative row is removed once and the result cached."""
    if bot == top:
        t = ops.dft_table(kind, h, 2 * top, top, 1.0, device)
        return t.view(4 * top, 2 * h) if kind == 2 else t.view(2 * h, 4 * top)
    key = (kind, h, top, bot, str(device))
    t = _row_tables.get(key)
    if t is None:
        full = ops.dft_table(kind, h, 2 * top, top, 1.0, device)
        # h == 2*top - 1: rows top-1 and top alias one frequency and the synthesis table already zeroes the first of them
        drop = top - 1 if (kind == 3 and h == 2 * top - 1) else top
        keep = torch.tensor([j for j in range(2 * top) if j != drop], device=device)
        if kind == 2:
            t = full.view(2 * top, 2, 2 * h).index_select(0, keep).reshape(2 * (top + bot), 2 * h).contiguous()
        else:
            t = full.view(2 * h, 2 * top, 2).index_select(1, keep).reshape(2 * h, 2 * (top + bot)).contiguous()
        _row_tables[key] = t
    return t


def resize(x: torch.Tensor, out_size, permute: bool = False) -> torch.Tensor:
    """2-D Fourier resize of the last two axes (res_utils.py:29-50): the kept row block [0, top) U [-bot, 0) on both sides is
    the C2C row table of the FNO2d transform (one more positive than negative row when the smaller height is odd)."""
    _check(x)
    if permute:
        x = x.permute(0, 3, 1, 2)
    h_in, w_in = x.shape[-2], x.shape[-1]
    h_out, w_out = int(out_size[0]), int(out_size[1])
    top = min((h_in + 1) // 2, (h_out + 1) // 2)                                   # res_utils.py:35-40
    bot = min(h_in // 2, h_out // 2)
    kx = min(w_in // 2 + 1, w_out // 2 + 1)
    lead = x.shape[:-2]
    x3 = x.reshape(-1, h_in, w_in).contiguous()
    B = x3.shape[0]
    dev = x.device
    fa_x = ops.dft_table(0, w_in, kx, 0, 1.0, dev).view(2 * kx, w_in)
    g_an = _row_table(2, h_in, top, bot, dev)                                      # rows (j, re|im), cols (re|im, h)
    g_sy = _row_table(3, h_out, top, bot, dev)                                     # rows (h, re|im), cols (j, re|im)
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


def _downsample_len(n_old: int, n: int):
    """number of integer modes of fftfreq(n_old, 1/n_old) inside [-n/2, n/2 - 1] (res_utils.py:21-22, 79-82): 2*floor(n/2)
    when that cuts anything, else None (every mode kept: the operator is the identity)"""
    ne = 2 * (int(n) // 2)
    return None if ne >= n_old + (n_old & 1) else ne


def downsample_1d(u: torch.Tensor, N: int) -> torch.Tensor:
    """Spectral decimation along the last axis (res_utils.py:62-91): complex fft, modes -N/2 <= k <= N/2 - 1, ifft, real part.
    For real input the kept block is k = 0 .. Ne/2 of the half spectrum with the lone k = -Ne/2 mode entering through its
    real part only: exactly the truncated analysis / irfft-weighted synthesis pair of resize_1d at Ne = 2*floor(N/2) points."""
    _check(u)
    n_old = u.shape[-1]
    ne = _downsample_len(n_old, N)
    if ne is None:
        return u.clone()
    if ne <= 0:
        return u.new_zeros(*u.shape[:-1], 0)
    y = _axis_last(u.reshape(-1, n_old), ne, ne // 2 + 1, 1.0 / n_old)
    return y.view(*u.shape[:-1], ne)


def downsample(u: torch.Tensor, N: int) -> torch.Tensor:
    """2-D spectral decimation of (B, C, S, S) fields (res_utils.py:19-27): fft2, rows and columns -N/2 <= k <= N/2 - 1,
    ifft2, real part.  Writing the kept block per axis as (real 1-D decimation) + i (lone -Ne/2 mode's imaginary part),
    the real part of the 2-D result is the product of the two 1-D decimations minus the product of the two imaginary
    parts, a rank-one corner term:  D_y D_x u  -  (-1)^(x+y) / S^2  *  sum u sin(a y') sin(a x'),  a = pi Ne / S."""
    _check(u)
    if u.dim() != 4:
        raise ValueError("downsample expects (B, C, S, S)")
    s_y, s_x = u.shape[-2], u.shape[-1]
    ne = _downsample_len(s_y, N)
    if ne is None:
        return u.clone()
    if s_x != s_y:
        raise ValueError("downsample: the reference selects both axes with the row frequencies (height == width)")
    if ne <= 0:
        return u.new_zeros(*u.shape[:-2], 0, 0)
    dev = u.device
    x3 = u.reshape(-1, s_y, s_x).contiguous()
    B = x3.shape[0]
    k = ne // 2 + 1
    v = _axis_last(x3.view(B * s_y, s_x), ne, k, 1.0 / s_x).view(B, s_y, ne)
    w = _axis_rows_to(v, ne, k, 1.0 / s_y)                                          # (B, ne, ne) = D_y D_x u
    # corner term: c[b] = s^T U_b s / S^2 through two thin contractions, then one rank-one update
    a = torch.arange(s_y, dtype=torch.float64, device=dev) * (torch.pi * ne / s_y)
    sv = torch.sin(a).to(torch.float32).view(s_y, 1)
    c = ops.matmul(ops.matmul(x3.view(B * s_y, s_x), sv).view(B, s_y), sv) / float(s_y * s_x)   # (B, 1)
    sign = 1.0 - 2.0 * (torch.arange(ne, device=dev) & 1).to(torch.float32)
    w -= c.view(B, 1, 1) * (sign.view(ne, 1) * sign.view(1, ne))
    return w.view(*u.shape[:-2], ne, ne)


def stride_downsample(x: torch.Tensor, factor: int, spatial_dims: int = 2, channels_last: bool = False) -> torch.Tensor:
    """Naive down-sampling: every `factor`-th grid point along the last `spatial_dims` axes (before a trailing channel axis
    when channels_last), i.e. x[..., ::f, ::f] / x[..., ::f, ::f, :] as dataloaders/ns_naive_true_multires.py:363-366 and the 1-D
    loaders' x[..., ::f] do on the host.  One strided-gather kernel on the device."""
    _check(x)
    if factor < 1 or spatial_dims not in (1, 2):
        raise ValueError("stride_downsample: factor >= 1, one or two spatial axes")
    x = x.contiguous()
    C = x.shape[-1] if channels_last else 1
    sp = x.shape[-spatial_dims - (1 if channels_last else 0):x.dim() - (1 if channels_last else 0)]
    H, W = (1, sp[0]) if spatial_dims == 1 else (sp[0], sp[1])
    lead = x.shape[:x.dim() - spatial_dims - (1 if channels_last else 0)]
    outer = 1
    for d in lead:
        outer *= d
    fh = 1 if spatial_dims == 1 else factor
    Ho, Wo = (H + fh - 1) // fh, (W + factor - 1) // factor
    out_sp = (Wo,) if spatial_dims == 1 else (Ho, Wo)
    y = torch.empty(tuple(lead) + out_sp + ((C,) if channels_last else ()), dtype=torch.float32, device=x.device)
    with ops._Ctx(x) as c:
        _lib.check(c.lib.rpde_stride_downsample(x.data_ptr(), y.data_ptr(), outer, H, W, C, fh, factor, c.stream), "rpde_stride_downsample")
    return y


def global_mean_std(tensors):
    """(mean, std) over every element of every tensor of an iterable -- the reference's 'simple global normalization'
    statistics (it extends a Python list with x.flatten().tolist() for every sample and calls torch.tensor(...).mean() and
    .std(), unbiased: dataloaders/ks_naive_markov.py:377-390, ns_naive_true_multires.py:513-525).  One streaming pass per tensor,
    fp64 accumulators on the device, one host read at the end.  Tensors may differ in shape (mixed resolutions)."""
    acc = None
    out = None
    for t in ([tensors] if isinstance(tensors, torch.Tensor) else tensors):
        _check(t)
        t = t.contiguous()
        if acc is None:
            acc = torch.zeros(3, dtype=torch.float64, device=t.device)
            out = torch.empty(2, dtype=torch.float32, device=t.device)
        with ops._Ctx(t) as c:
            _lib.check(c.lib.rpde_moments_accumulate(t.data_ptr(), t.numel(), acc.data_ptr(), c.stream), "rpde_moments_accumulate")
    if acc is None:
        raise ValueError("global_mean_std: no tensors")
    with ops._Ctx(out) as c:
        _lib.check(c.lib.rpde_moments_finalize(acc.data_ptr(), out.data_ptr(), c.stream), "rpde_moments_finalize")
    m, s = out.tolist()
    return m, s
