"""torch.autograd.Function wrappers over the C ABI (include/rpde_b200.h).

PyTorch is plumbing here: it owns device memory (caching allocator), the current stream and autograd's
tape.  Every arithmetic step is a call into librpde_b200.so on the tensor's device and torch's current
stream.  No CPU path exists: a non-CUDA tensor raises.
"""
from __future__ import annotations

import math
import os
import threading
from typing import Optional, Sequence, Tuple

import numpy as np
import torch

from . import _lib

ACT_NONE, ACT_GELU, ACT_RELU = 0, 1, 2
ACT_SAVE_GRAD, DACT_SAVED = 0x100, 3   # forward stores act'(z); backward multiplies by the stored factor
MIX_FULL, MIX_LOWPASS = 0, 1
PREC = {"fp32": 0, "3xtf32": 1, "tf32": 2, "bf16": 3}
_ACT_BY_NAME = {"identity": ACT_NONE, "none": ACT_NONE, "gelu": ACT_GELU, "relu": ACT_RELU}

_state = threading.local()
_default_precision = 0
_fno_fused = os.environ.get("RPDE_FNO_FUSED", "1") != "0"   # fused FNOBlock1d forward (rpde_fno_block1d_fwd); 0 = composed path
# fused backward prologue (rpde_fno_block1d_bwd: act' multiply + synthesis adjoint in one launch).  FNO1d Burgers fwd+bwd:
# 0.723 ms composed -> 0.679 ms fused forward -> 0.638 ms with this prologue (profiles/README.md r1i); 0 = composed prologue
_fno_fused_bwd = os.environ.get("RPDE_FNO_FUSED_BWD", "1") != "0"
# rpde_fno_block1d_bwd can also produce dx (analysis adjoint + bypass dgrad) in one kernel when it is handed the bypass
# weight.  With the tensor-path synthesis kernel (csrc/fno2d_mma.cuh: N % 128 == 0, at most 64 channels) that is the faster
# choice on launch-bound batches (FNO1d Burgers, batch 16: 0.383 against 0.399 ms/step) and the slower one on large ones
# (batch 256: 3.34 against 2.80 ms), so the default is automatic; RPDE_FNO_FUSED_DX=1 / 0 forces it on / off.
_fno_fused_dx = {"1": True, "0": False}.get(os.environ.get("RPDE_FNO_FUSED_DX", ""), None)


def _use_fused_dx(B: int, I: int, O: int, n: int) -> bool:
    if _fno_fused_dx is not None:
        return _fno_fused_dx
    return B <= 64 and n % 128 == 0 and I <= 64 and O <= 64


def set_precision(name: str) -> None:
    """'fp32' (CUDA-core FMA, default), '3xtf32' (tcgen05, fp32-accurate), 'tf32' (tcgen05, ~5e-4) or 'bf16' (fused BF16
    FeedForward kernels where the shape allows, the TF32 engine elsewhere; model-level tolerance 1e-2 / 4e-2)."""
    global _default_precision
    _default_precision = PREC[name]


def get_precision() -> str:
    return {v: k for k, v in PREC.items()}[_default_precision]


def _sync_precision(lib) -> None:
    # the C side keeps the engine choice per thread (DataParallel runs one thread per device)
    if getattr(_state, "prec", None) != _default_precision:
        _lib.check(lib.rpde_set_precision(_default_precision), "rpde_set_precision")
        _state.prec = _default_precision


def act_id(act) -> Optional[int]:
    """Map the reference's activation spellings (callables or registry names) to an epilogue id."""
    import torch.nn.functional as F
    if act is None:
        return ACT_NONE
    if isinstance(act, str):
        return _ACT_BY_NAME[act]
    if act is F.relu or act is torch.relu or isinstance(act, torch.nn.ReLU):
        return ACT_RELU
    if act is F.gelu or (isinstance(act, torch.nn.GELU) and getattr(act, "approximate", "none") == "none"):
        return ACT_GELU
    if isinstance(act, torch.nn.Identity):
        return ACT_NONE
    return None  # unknown callable: caller applies it outside the fused epilogue


def _p(t: Optional[torch.Tensor]):
    return None if t is None else t.data_ptr()


def _need_cuda(*ts: torch.Tensor) -> None:
    for t in ts:
        if t is not None and not t.is_cuda:
            raise RuntimeError("resolution_pde_b200 runs on CUDA tensors only (B200, sm_100a); there is no CPU fallback")
        if t is not None and t.dtype != torch.float32:
            raise TypeError(f"resolution_pde_b200 expects float32 tensors, got {t.dtype}")


class _Ctx:
    """device guard + library + current stream for one op call"""

    def __init__(self, t: torch.Tensor):
        self.dev = t.device
        self.guard = torch.cuda.device(self.dev)

    def __enter__(self):
        self.guard.__enter__()
        self.lib = _lib.load()
        _sync_precision(self.lib)
        self.stream = torch.cuda.current_stream(self.dev).cuda_stream
        return self

    def __exit__(self, *a):
        return self.guard.__exit__(*a)


# ---------------------------------------------------------------------------------------------
# cached device tables: DFT matrices and coordinate grids
# ---------------------------------------------------------------------------------------------
_tables = {}
_tables_lock = threading.Lock()


def dft_table(kind: int, n: int, k: int, m1: int, scale: float, device) -> torch.Tensor:
    key = (kind, n, k, m1, float(scale), str(device))
    t = _tables.get(key)
    if t is None:
        with _tables_lock:
            t = _tables.get(key)
            if t is None:
                t = torch.empty((2 if kind in (0, 1, 4) else 4) * k * n, dtype=torch.float32, device=device)
                with _Ctx(t) as c:
                    _lib.check(c.lib.rpde_dft_table(_p(t), kind, n, k, m1, float(scale), c.stream), "rpde_dft_table")
                _tables[key] = t
    return t


def norm_scales(norm: Optional[str], n: int) -> Tuple[float, float]:
    """(forward, inverse) scale factors of torch.fft for a transform of length n."""
    if norm is None or norm == "backward":
        return 1.0, 1.0 / n
    if norm == "ortho":
        return 1.0 / math.sqrt(n), 1.0 / math.sqrt(n)
    if norm == "forward":
        return 1.0 / n, 1.0
    raise RuntimeError(f"Invalid normalization mode: \"{norm}\"")


def coord_grid(values, n: int, device, lo: float = 0.0, hi: float = 1.0) -> torch.Tensor:
    """np.linspace(lo, hi, n) in fp64 rounded to fp32 (fno.py:51,121-125; ffno.py:92,201-205), cached per device,
    or a user grid converted the way the reference does (torch.tensor(..., dtype=torch.float))."""
    if values is not None:
        t = values if isinstance(values, torch.Tensor) else torch.tensor(values, dtype=torch.float)
        return t.to(device=device, dtype=torch.float32).reshape(-1).contiguous()
    key = ("grid", n, lo, hi, str(device))
    t = _tables.get(key)
    if t is None:
        t = torch.tensor(np.linspace(lo, hi, n), dtype=torch.float).to(device)
        _tables[key] = t
    return t


def _ws(nbytes: int, device) -> torch.Tensor:
    return torch.empty(max(int(nbytes), 256), dtype=torch.uint8, device=device)


# ---------------------------------------------------------------------------------------------
# FNO spectral convolutions (+ optional fused bypass/activation of the FNO block)
# ---------------------------------------------------------------------------------------------
def _spectral_tables(ndim, shape, modes, device):
    if ndim == 1:
        n, k = shape[0], modes[0]
        return (dft_table(0, n, k, 0, 1.0, device), dft_table(1, n, k, 0, 1.0 / n, device))
    h, w = shape
    m1, m2 = modes
    return (dft_table(0, w, m2, 0, 1.0, device), dft_table(4, w, m2, 0, 1.0 / w, device),
            dft_table(2, h, 2 * m1, m1, 1.0, device), dft_table(3, h, 2 * m1, m1, 1.0 / h, device))


def _g2t_table(h, m1, device):
    """the H-axis analysis table (kind 2, [(j,ro)][(ri,h)]) transposed to [(h,ri)][(j,ro)]: both H-axis kernels of the fused
    FNOBlock2d read their table as [(h, r)][a], the kind-3 table has that layout as generated"""
    key = ("fno2d_g2t", h, m1, str(device))
    t = _tables.get(key)
    if t is None:
        j = 2 * m1
        g2 = dft_table(2, h, j, m1, 1.0, device)
        t = g2.view(j, 2, 2, h).permute(3, 2, 0, 1).reshape(2 * h, 2 * j).contiguous()
        _tables[key] = t
    return t


def _check_modes(ndim, spatial, modes):
    if ndim == 1:
        if modes[0] > spatial[0] // 2 + 1:
            raise RuntimeError(f"SpectralConv1d: modes1={modes[0]} exceeds N//2+1={spatial[0] // 2 + 1}")
    else:
        if modes[1] > spatial[1] // 2 + 1 or modes[0] > spatial[0]:
            raise RuntimeError(f"SpectralConv2d: modes {tuple(modes)} exceed the spectrum of a {tuple(spatial)} grid")


# ---------------------------------------------------------------------------------------------
# Parameter-gradient side stream.  On the launch-bound FNO steps (FNO1d at batch 16: ~60 kernels of 4-16 us) the kernels
# that only produce a parameter gradient are off the critical path: nothing in the backward pass reads their result.
# training.GraphedTrainStep opens `pgrad_side_stream` while it captures; those kernels are then enqueued on a second
# stream (forked after everything the main stream has enqueued so far, joined once at the end of the backward pass), so the
# captured graph runs them next to the input-gradient chain.  Outside that context everything stays on one stream.
# ---------------------------------------------------------------------------------------------
_pgrad = {}     # device index -> [stream, kept tensors, forked]; autograd runs CUDA nodes on its own per-device threads, so
                # this is process-wide state keyed by device, not thread-local


class pgrad_side_stream:
    def __init__(self, device):
        self.dev = torch.device(device)
        self.idx = self.dev.index if self.dev.index is not None else torch.cuda.current_device()
        self.stream = torch.cuda.Stream(device=self.dev)

    def __enter__(self):
        _pgrad[self.idx] = [self.stream, [], False]
        return self

    def __exit__(self, *a):
        st = _pgrad.pop(self.idx, None)
        if st is not None and st[2]:   # a forked stream must rejoin before the capture ends; its inputs were kept alive until here
            torch.cuda.current_stream(self.dev).wait_stream(self.stream)
        return False


def _pgrad_stream(c: "_Ctx", *reads):
    """stream handle for a kernel that only writes parameter gradients; `reads`: the tensors it reads (kept alive until the join,
    so that the allocator cannot hand their memory to later main-stream work)"""
    st = _pgrad.get(c.dev.index if c.dev.index is not None else torch.cuda.current_device())
    if st is None:
        return c.stream
    st[0].wait_stream(torch.cuda.current_stream(c.dev))
    st[2] = True
    st[1].extend(t for t in reads if t is not None)
    return st[0].cuda_stream


def _dw_db(w: torch.Tensor, O: int, has_bias: bool):
    """gradient buffers of a 1x1 conv laid out back to back (dw | db): rpde_conv1x1_bwd_params then zero-fills both with
    one memset (its streaming kernel accumulates with atomics)"""
    if not has_bias:
        return torch.empty_like(w), None
    flat = torch.empty(w.numel() + O, dtype=torch.float32, device=w.device)
    return flat[:w.numel()].view_as(w), flat[w.numel():]


class _FNOSpectralFn(torch.autograd.Function):
    """y = act(spectral(x; weights) [+ conv1x1(x; wb, bb)])   channels-first, 1-D or 2-D."""

    @staticmethod
    def forward(ctx, x, w1, w2, wb, bb, act, ndim, grad_on=True):
        _need_cuda(x, wb, bb)
        x = x.contiguous()
        B, I = x.shape[0], x.shape[1]
        spatial = tuple(x.shape[2:])
        O = w1.shape[1]
        modes = tuple(w1.shape[2:])
        _check_modes(ndim, spatial, modes)
        w1r = torch.view_as_real(w1.detach().contiguous())
        w2r = torch.view_as_real(w2.detach().contiguous()) if w2 is not None else None
        S = int(np.prod(spatial))
        y = torch.empty((B, O) + spatial, dtype=torch.float32, device=x.device)
        need_grad = grad_on and any(ctx.needs_input_grad)
        z = torch.empty_like(y) if (act != ACT_NONE and need_grad) else None
        tabs = _spectral_tables(ndim, spatial, modes, x.device)
        nmodes = int(np.prod(modes)) * (2 if ndim == 2 else 1)
        spec = torch.empty(nmodes * B * 2 * I, dtype=torch.float32, device=x.device)
        if bb is not None and wb is None:
            raise RuntimeError("fno_spectral: a bias needs the bypass weight")
        fused = ndim == 1 and _fno_fused and spatial[0] <= 2048 and 8 * modes[0] * (I + 1) <= 200 * 1024 and B <= 65535
        if fused:
            # FNOBlock1d in three launches (analysis, mixing, synthesis + bypass + bias + activation), fp32 FMA
            n, k = spatial[0], modes[0]
            fs4 = dft_table(4, n, k, 0, 1.0 / n, x.device)
            wbc = wb.detach().reshape(O, I).contiguous() if wb is not None else None
            up = lambda v: (v + 255) // 256 * 256
            nb = up(k * B * 2 * O * 4) + (0 if need_grad else up(k * B * 2 * I * 4))
            ws = _ws(nb, x.device)
            with _Ctx(x) as c:
                _lib.check(c.lib.rpde_fno_block1d_fwd(_p(x), _p(w1r), _p(tabs[0]), _p(fs4), _p(wbc), _p(bb),
                                                      _p(spec) if need_grad else None, _p(y), _p(z), B, I, O, n, k, act,
                                                      _p(ws), nb, c.stream), "rpde_fno_block1d_fwd")
            ctx.save_for_backward(x, w1, w2, wb, spec, z)
            ctx.cfg = (act, ndim, B, I, O, spatial, modes, S, bb is not None, True)
            return y
        if ndim == 2 and _fno_fused and w2 is not None:
            with _Ctx(x) as c:
                if c.lib.rpde_fno_block2d_supported(B, I, O, spatial[0], spatial[1], modes[0], modes[1]):
                    # FNOBlock2d in five fp32-FMA launches (csrc/fno2d_fused.cuh), engine-independent
                    wbc = wb.detach().reshape(O, I).contiguous() if wb is not None else None
                    nb = c.lib.rpde_fno_block2d_workspace_bytes(B, I, O, spatial[0], spatial[1], modes[0], modes[1])
                    ws = _ws(nb, x.device)
                    g2t = _g2t_table(spatial[0], modes[0], x.device)
                    _lib.check(c.lib.rpde_fno_block2d_fwd(_p(x), _p(w1r), _p(w2r), _p(tabs[0]), _p(tabs[1]), _p(g2t), _p(tabs[3]),
                                                          _p(wbc), _p(bb), _p(spec) if need_grad else None, _p(y), _p(z), B, I, O,
                                                          spatial[0], spatial[1], modes[0], modes[1], act, _p(ws), nb, c.stream),
                               "rpde_fno_block2d_fwd")
                    ctx.save_for_backward(x, w1, w2, wb, spec, z)
                    ctx.cfg = (act, ndim, B, I, O, spatial, modes, S, bb is not None, True)
                    return y
        with _Ctx(x) as c:
            beta = 0
            if wb is not None:
                wbc = wb.detach().reshape(O, I).contiguous()
                _lib.check(c.lib.rpde_conv1x1_fwd(_p(x), _p(wbc), _p(bb), _p(y), None, B, I, O, S, ACT_NONE, 0, c.stream),
                           "rpde_conv1x1_fwd")
                beta = 1
            if ndim == 1:
                nb = c.lib.rpde_spectral1d_workspace_bytes(B, I, O, spatial[0], modes[0])
                ws = _ws(nb, x.device)
                _lib.check(c.lib.rpde_spectral1d_fwd(_p(x), _p(w1r), _p(tabs[0]), _p(tabs[1]), _p(spec), _p(y), _p(z), B, I, O,
                                                     spatial[0], modes[0], beta, act, _p(ws), nb, c.stream),
                           "rpde_spectral1d_fwd")
            else:
                nb = c.lib.rpde_spectral2d_workspace_bytes(B, I, O, spatial[0], spatial[1], modes[0], modes[1])
                ws = _ws(nb, x.device)
                _lib.check(c.lib.rpde_spectral2d_fwd(_p(x), _p(w1r), _p(w2r), _p(tabs[0]), _p(tabs[1]), _p(tabs[2]), _p(tabs[3]),
                                                     _p(spec), _p(y), _p(z), B, I, O, spatial[0], spatial[1], modes[0],
                                                     modes[1], beta, act, _p(ws), nb, c.stream), "rpde_spectral2d_fwd")
        ctx.save_for_backward(x, w1, w2, wb, spec, z)
        ctx.cfg = (act, ndim, B, I, O, spatial, modes, S, bb is not None, False)
        return y

    @staticmethod
    def backward(ctx, dy):
        x, w1, w2, wb, spec, z = ctx.saved_tensors
        act, ndim, B, I, O, spatial, modes, S, has_bias, fwd_fused = ctx.cfg
        dy = dy.contiguous()
        need_x = ctx.needs_input_grad[0]
        w1r = torch.view_as_real(w1.detach().contiguous())
        w2r = torch.view_as_real(w2.detach().contiguous()) if w2 is not None else None
        tabs = _spectral_tables(ndim, spatial, modes, x.device)
        dx = torch.empty_like(x) if need_x else None
        dw1 = torch.empty_like(w1r)
        dw2 = torch.empty_like(w2r) if w2 is not None else None
        dwb = dbb = None
        fused = fwd_fused and _fno_fused_bwd          # one decision, taken in forward (same shape limits)
        dx_done = False
        if fwd_fused and ndim == 2:
            g2t = _g2t_table(spatial[0], modes[0], x.device)
            dpre = torch.empty_like(dy) if act != ACT_NONE else None
            wbc = wb.detach().reshape(O, I).contiguous() if wb is not None else None
            with _Ctx(x) as c:
                nb = c.lib.rpde_fno_block2d_workspace_bytes(B, I, O, spatial[0], spatial[1], modes[0], modes[1])
                ws = _ws(nb, x.device)
                _lib.check(c.lib.rpde_fno_block2d_bwd(_p(dy), _p(z) if dpre is not None else None, act, _p(w1r), _p(w2r), _p(tabs[0]),
                                                      _p(tabs[1]), _p(tabs[3]), _p(g2t), _p(spec), _p(wbc), _p(dpre), _p(dx), _p(dw1),
                                                      _p(dw2), B, I, O, spatial[0], spatial[1], modes[0], modes[1], _p(ws), nb,
                                                      c.stream), "rpde_fno_block2d_bwd")
                if dpre is None:
                    dpre = dy
                if wb is not None:
                    dwb, dbb = _dw_db(wb, O, has_bias)
                    _lib.check(c.lib.rpde_conv1x1_bwd_params(_p(dpre), _p(x), _p(dwb), _p(dbb), B, I, O, S,
                                                             _pgrad_stream(c, dpre, x) if S % 32 == 0 and max(I, O) <= 64 else c.stream),
                               "rpde_conv1x1_bwd_params")
            return (dx, torch.view_as_complex(dw1), torch.view_as_complex(dw2), dwb, dbb, None, None, None)
        with _Ctx(x) as c:
            if fused:
                # act' multiply + projection on the synthesis table in one launch, then the spectral backward proper
                n, k = spatial[0], modes[0]
                dpre = torch.empty_like(dy) if act != ACT_NONE else None
                fs4 = dft_table(4, n, k, 0, 1.0 / n, x.device)
                nb = c.lib.rpde_spectral1d_workspace_bytes(B, I, O, n, k)
                ws = _ws(nb, x.device)
                wbc = wb.detach().reshape(O, I).contiguous() if wb is not None else None
                dx_done = bool(_use_fused_dx(B, I, O, n) and wbc is not None and need_x)   # dx then includes the bypass conv's input gradient
                _lib.check(c.lib.rpde_fno_block1d_bwd(_p(dy), _p(z) if dpre is not None else None, act, _p(w1r), _p(tabs[0]), _p(fs4),
                                                      _p(spec), _p(wbc) if dx_done else None, _p(dpre), _p(dx), 0, _p(dw1), B, I, O,
                                                      n, k, _p(ws), nb, c.stream), "rpde_fno_block1d_bwd")
                if dpre is None:
                    dpre = dy
            elif act != ACT_NONE:
                dpre = torch.empty_like(dy)
                _lib.check(c.lib.rpde_act_bwd(_p(dy), _p(z), _p(dpre), dy.numel(), act, c.stream), "rpde_act_bwd")
            else:
                dpre = dy
            if fused:
                pass
            elif ndim == 1:
                nb = c.lib.rpde_spectral1d_workspace_bytes(B, I, O, spatial[0], modes[0])
                ws = _ws(nb, x.device)
                _lib.check(c.lib.rpde_spectral1d_bwd(_p(dpre), _p(w1r), _p(tabs[0]), _p(tabs[1]), _p(spec), _p(dx), 0, _p(dw1),
                                                     B, I, O, spatial[0], modes[0], _p(ws), nb, c.stream),
                           "rpde_spectral1d_bwd")
            else:
                nb = c.lib.rpde_spectral2d_workspace_bytes(B, I, O, spatial[0], spatial[1], modes[0], modes[1])
                ws = _ws(nb, x.device)
                _lib.check(c.lib.rpde_spectral2d_bwd(_p(dpre), _p(w1r), _p(w2r), _p(tabs[0]), _p(tabs[1]), _p(tabs[2]), _p(tabs[3]),
                                                     _p(spec), _p(dx), 0, _p(dw1), _p(dw2), B, I, O, spatial[0], spatial[1],
                                                     modes[0], modes[1], _p(ws), nb, c.stream), "rpde_spectral2d_bwd")
            if wb is not None:
                wbc = wb.detach().reshape(O, I).contiguous()
                if need_x and not (fused and dx_done):
                    _lib.check(c.lib.rpde_conv1x1_bwd_input(_p(dpre), _p(wbc), None, 0, _p(dx), 1, B, I, O, S, c.stream),
                               "rpde_conv1x1_bwd_input")
                dwb, dbb = _dw_db(wb, O, has_bias)
                _lib.check(c.lib.rpde_conv1x1_bwd_params(_p(dpre), _p(x), _p(dwb), _p(dbb), B, I, O, S,
                                                         _pgrad_stream(c, dpre, x) if S % 32 == 0 and max(I, O) <= 64 else c.stream),
                           "rpde_conv1x1_bwd_params")
        return (dx, torch.view_as_complex(dw1), torch.view_as_complex(dw2) if dw2 is not None else None, dwb, dbb,
                None, None, None)


def fno_spectral(x, w1, w2=None, bypass_weight=None, bypass_bias=None, act: int = ACT_NONE):
    # autograd.Function.forward always runs with grad mode off: sample it here so torch.no_grad() inference
    # (rollouts, evaluation) saves nothing and keeps the spectrum in the workspace
    return _FNOSpectralFn.apply(x, w1, w2, bypass_weight, bypass_bias, act, x.dim() - 2, torch.is_grad_enabled())


# ---------------------------------------------------------------------------------------------
# F-FNO factorised Fourier pass (forward_fourier), channels-last
# ---------------------------------------------------------------------------------------------
class _FFNOFourierFn(torch.autograd.Function):
    """1-D: x (B,N,C), weights (w,).  2-D: x (B,M,N,C), weights (w_y, w_x); result = x-pass + y-pass."""

    @staticmethod
    def forward(ctx, x, n_modes, mix, norm, grad_on, with_skip, *weights):
        # with_skip: x is returned as a second output (the layer's skip connection, F-FNO `x + layer(x)`).  Its gradient then
        # arrives HERE together with the gradient of the spectral result, and the first backward pass accumulates onto it
        # in place: no separate add kernel for the two uses of x (autograd would insert one).
        _need_cuda(x, *weights)
        x_in = x
        x = x.contiguous()
        C = x.shape[-1]
        B = x.shape[0]
        ndim = x.dim() - 2
        y = torch.empty_like(x)
        # (outer, n, rest, weight index) per pass; the 2-D layer transforms the last axis with weight 0 first
        if ndim == 1:
            passes = [(B, x.shape[1], 1, 0)]
        else:
            passes = [(B * x.shape[1], x.shape[2], 1, 0), (B, x.shape[1], x.shape[2], 1)]
        specs = []
        with _Ctx(x) as c:
            for pi, (outer, n, rest, wi) in enumerate(passes):
                k = min(n_modes, n // 2 + 1)
                sf, si = norm_scales(norm, n)
                fa = dft_table(0, n, k, 0, sf, x.device)
                fs = dft_table(1, n, k, 0, si, x.device)
                spec = torch.empty(outer * 2 * k * rest * C, dtype=torch.float32, device=x.device)
                nb = c.lib.rpde_ffno_axis_workspace_bytes(outer, n, rest, C, k)
                ws = _ws(nb, x.device)
                w = weights[wi].detach().contiguous()
                _lib.check(c.lib.rpde_ffno_axis_fwd(_p(x), _p(w), _p(fa), _p(fs), _p(spec), _p(y), 1 if pi else 0, outer, n,
                                                    rest, C, k, w.shape[2], mix, _p(ws), nb, c.stream),
                           "rpde_ffno_axis_fwd")
                specs.append(spec)
        keep_spec = grad_on and mix == MIX_FULL and any(ctx.needs_input_grad[6:])
        ctx.save_for_backward(*weights, *(specs if keep_spec else []))
        ctx.cfg = (n_modes, mix, norm, passes, C, len(weights), keep_spec, tuple(x.shape))
        if with_skip:
            return y, x_in
        return y

    @staticmethod
    def backward(ctx, dy, dskip=None):
        n_modes, mix, norm, passes, C, nw, keep_spec, xshape = ctx.cfg
        saved = ctx.saved_tensors
        weights, specs = saved[:nw], saved[nw:]
        dy = dy.contiguous()
        need_x = ctx.needs_input_grad[0]
        acc0 = 0
        dx = None
        if need_x:
            if dskip is not None and dskip.is_contiguous() and dskip.dtype == torch.float32 and tuple(dskip.shape) == xshape:
                dx, acc0 = dskip, 1        # the skip connection's gradient: the passes add onto it in place
            else:
                dx = torch.empty(xshape, dtype=torch.float32, device=dy.device)
                if dskip is not None:
                    dx.copy_(dskip)
                    acc0 = 1
        dws = [None] * nw
        with _Ctx(dy) as c:
            for pi, (outer, n, rest, wi) in enumerate(passes):
                k = min(n_modes, n // 2 + 1)
                sf, si = norm_scales(norm, n)
                fa = dft_table(0, n, k, 0, sf, dy.device)
                fs = dft_table(1, n, k, 0, si, dy.device)
                w = weights[wi].detach().contiguous()
                want_w = keep_spec and ctx.needs_input_grad[6 + wi]
                dw = None
                if want_w:
                    dw = torch.empty_like(w)
                    dws[wi] = dw
                spec = specs[pi] if keep_spec else None
                nb = c.lib.rpde_ffno_axis_workspace_bytes(outer, n, rest, C, k)
                ws = _ws(nb, dy.device)
                _lib.check(c.lib.rpde_ffno_axis_bwd(_p(dy), _p(w), _p(fa), _p(fs), _p(spec), _p(dx), 1 if pi else acc0, _p(dw), 0,
                                                    outer, n, rest, C, k, w.shape[2], mix, _p(ws), nb, c.stream),
                           "rpde_ffno_axis_bwd")
        if mix != MIX_FULL:
            dws = [torch.zeros_like(w) if ctx.needs_input_grad[6 + i] else None for i, w in enumerate(weights)]
        return (dx, None, None, None, None, None, *dws)


def ffno_fourier(x, weights: Sequence[torch.Tensor], n_modes: int, mix: int = MIX_FULL, norm: str = "ortho", with_skip: bool = False):
    """with_skip=True returns (fourier(x), x): route the layer's skip connection through the second output (see forward)."""
    return _FFNOFourierFn.apply(x, n_modes, mix, norm, torch.is_grad_enabled(), bool(with_skip), *weights)


# ---------------------------------------------------------------------------------------------
# Linear / FeedForward (channels-last)
# ---------------------------------------------------------------------------------------------
class _LinearFn(torch.autograd.Function):
    """y = x W^T + b over the last dim; cf_spatial>0 writes y channels-first (B, out, *spatial)."""

    @staticmethod
    def forward(ctx, x, w, b, out_shape_cf):
        _need_cuda(x, w, b)
        x = x.contiguous()
        w = w.contiguous()
        out_f, in_f = w.shape
        T = x.numel() // in_f
        if out_shape_cf is not None:
            y = torch.empty(out_shape_cf, dtype=torch.float32, device=x.device)
            S = int(np.prod(out_shape_cf[2:]))
        else:
            y = torch.empty(x.shape[:-1] + (out_f,), dtype=torch.float32, device=x.device)
            S = 0
        with _Ctx(x) as c:
            _lib.check(c.lib.rpde_linear_fwd(_p(x), _p(w), _p(b), _p(y), None, T, in_f, out_f, ACT_NONE, S, c.stream),
                       "rpde_linear_fwd")
        ctx.save_for_backward(x, w)
        ctx.cfg = (T, in_f, out_f, S, b is not None)
        return y

    @staticmethod
    def backward(ctx, dy):
        x, w = ctx.saved_tensors
        T, in_f, out_f, S, has_b = ctx.cfg
        dy = dy.contiguous()
        dx = dw = db = None
        with _Ctx(x) as c:
            if ctx.needs_input_grad[0]:
                dx = torch.empty_like(x)
                _lib.check(c.lib.rpde_linear_bwd_input(_p(dy), _p(w), None, 0, _p(dx), 0, None, T, in_f, out_f, S, c.stream),
                           "rpde_linear_bwd_input")
            if ctx.needs_input_grad[1] or (has_b and ctx.needs_input_grad[2]):
                dw = torch.empty_like(w)
                db = torch.empty(out_f, dtype=torch.float32, device=x.device) if has_b else None
                _lib.check(c.lib.rpde_linear_bwd_params(_p(dy), _p(x), _p(dw), _p(db), T, in_f, out_f, S, c.stream),
                           "rpde_linear_bwd_params")
        return dx, dw, db, None


def linear(x, w, b=None, channels_first_out: Optional[Tuple[int, ...]] = None):
    return _LinearFn.apply(x, w, b, channels_first_out)


_ff3_images = {}


def _ff3_image(c, Ws, cache: bool) -> torch.Tensor:
    """BF16 shared-memory image of the three FeedForward weights (rpde_ff3_pack_weights).  Under no_grad (constant
    weights: rollouts, evaluation) the image is cached on the parameters' storage and version counters."""
    key = tuple((w.data_ptr(), w._version) for w in Ws) if cache else None
    if key is not None:
        hit = _ff3_images.get(key[0][0])
        if hit is not None and hit[0] == key:
            return hit[1]
    img = torch.empty(int(c.lib.rpde_ff3_weight_image_bytes()), dtype=torch.uint8, device=Ws[0].device)
    _lib.check(c.lib.rpde_ff3_pack_weights(_p(Ws[0]), _p(Ws[1]), _p(Ws[2]), _p(img), c.stream), "rpde_ff3_pack_weights")
    if key is not None:
        _ff3_images[key[0][0]] = (key, img)
    return img


class _FusedFF3Fn(torch.autograd.Function):
    """y = residual + LayerNorm(FeedForward_3(x)) in one BF16 tcgen05 kernel (rpde_ff3_fwd); the backward pass recomputes
    the hidden activations on chip from x (rpde_ff3_bwd) - nothing but x is kept between the two."""

    @staticmethod
    def forward(ctx, x, residual, use_ln, eps, grad_on, w1, b1, w2, b2, w3, b3, gamma, beta):
        _need_cuda(x, residual, w1, b1, w2, b2, w3, b3, gamma, beta)
        x = x.contiguous()
        T = x.numel() // x.shape[-1]
        res = residual.contiguous() if residual is not None else None
        Ws = [w1.contiguous(), w2.contiguous(), w3.contiguous()]
        bs = [b1.contiguous(), b2.contiguous(), b3.contiguous()]
        g = gamma.contiguous() if use_ln else None
        bt = beta.contiguous() if use_ln else None
        y = torch.empty_like(x)
        need_grad = grad_on and any(ctx.needs_input_grad)
        with _Ctx(x) as c:
            img = _ff3_image(c, Ws, cache=not need_grad)
            _lib.check(c.lib.rpde_ff3_fwd(_p(x), _p(res), _p(img), _p(bs[0]), _p(bs[1]), _p(bs[2]), _p(g), _p(bt), _p(y), T,
                                          float(eps), c.stream), "rpde_ff3_fwd")
        if need_grad:
            ctx.save_for_backward(x, img, *bs, g)
            ctx.cfg = (T, use_ln, float(eps), residual is not None)
        return y

    @staticmethod
    def backward(ctx, dy):
        x, img, b1, b2, b3, g = ctx.saved_tensors
        T, use_ln, eps, has_res = ctx.cfg
        dy = dy.contiguous()
        dev = dy.device
        f = lambda *s: torch.empty(s, dtype=torch.float32, device=dev)
        ds = torch.empty_like(x)
        dw1, dw2, dw3 = f(256, 64), f(256, 256), f(64, 256)
        db1, db2, db3 = f(256), f(256), f(64)
        dg, dbt = (f(64), f(64)) if use_ln else (None, None)
        with _Ctx(dy) as c:
            nb = int(c.lib.rpde_ff3_bwd_workspace_bytes(T))
            ws = _ws(nb, dev)
            _lib.check(c.lib.rpde_ff3_bwd(_p(x), _p(dy), _p(img), _p(b1), _p(b2), _p(b3), _p(g), _p(ds), _p(dw1), _p(db1),
                                          _p(dw2), _p(db2), _p(dw3), _p(db3), _p(dg), _p(dbt), T, eps, _p(ws), nb, c.stream),
                       "rpde_ff3_bwd")
        return (ds if ctx.needs_input_grad[0] else None, dy if has_res else None, None, None, None,
                dw1, db1, dw2, db2, dw3, db3, dg, dbt)


def _ff3_applies(x, weights, biases, final_act) -> bool:
    if _default_precision != PREC["bf16"] or len(weights) != 3 or final_act != ACT_NONE or any(b is None for b in biases):
        return False
    c = x.shape[-1]
    h = weights[0].shape[0]
    return (tuple(weights[0].shape) == (h, c) and tuple(weights[1].shape) == (h, h) and tuple(weights[2].shape) == (c, h)
            and bool(_lib.load().rpde_ff3_supported(c, h)))


class _FeedForwardFn(torch.autograd.Function):
    """FeedForward (custom_layer.py:49-68, dropout 0) fused with what follows it in the F-FNO layer:
       u = L_n(gelu(... gelu(L_1(x))));  v = LayerNorm(u) if ln;  y = residual + act(v)."""

    @staticmethod
    def forward(ctx, x, residual, n_layers, use_ln, final_act, eps, grad_on, *params):
        _need_cuda(x, residual, *params)
        x = x.contiguous()
        C_in = x.shape[-1]
        T = x.numel() // C_in
        Ws = [params[2 * i].contiguous() for i in range(n_layers)]
        bs = [params[2 * i + 1] for i in range(n_layers)]
        gamma = params[2 * n_layers] if use_ln else None
        beta = params[2 * n_layers + 1] if use_ln else None
        need_grad = grad_on and any(ctx.needs_input_grad)
        acts, zs = [x], []
        dev = x.device
        res = residual.contiguous() if residual is not None else None
        with _Ctx(x) as c:
            a = x
            for i in range(n_layers - 1):
                out_f, in_f = Ws[i].shape
                h = torch.empty((T, out_f), dtype=torch.float32, device=dev)
                z = torch.empty_like(h) if need_grad else None
                # z holds gelu'(pre-activation): the backward pass multiplies, no erf/exp there
                _lib.check(c.lib.rpde_linear_fwd(_p(a), _p(Ws[i]), _p(bs[i]), _p(h), _p(z), T, in_f, out_f,
                                                 ACT_GELU | (ACT_SAVE_GRAD if z is not None else 0), 0, c.stream),
                           "rpde_linear_fwd")
                acts.append(h)
                zs.append(z)
                a = h
            if n_layers:
                out_f, in_f = Ws[-1].shape
                u = torch.empty(x.shape[:-1] + (out_f,), dtype=torch.float32, device=dev)
                _lib.check(c.lib.rpde_linear_fwd(_p(a), _p(Ws[-1]), _p(bs[-1]), _p(u), None, T, in_f, out_f, ACT_NONE, 0,
                                                 c.stream), "rpde_linear_fwd")
            else:  # activation / LayerNorm / residual only (used by the dropout path)
                u, out_f = x, C_in
            mean = rstd = v = None
            y = u
            if use_ln:
                mean = torch.empty(T, dtype=torch.float32, device=dev) if need_grad else None
                rstd = torch.empty(T, dtype=torch.float32, device=dev) if need_grad else None
                v = torch.empty_like(u)
                fuse_res = res if final_act == ACT_NONE else None
                _lib.check(c.lib.rpde_layernorm_fwd(_p(u), _p(fuse_res), _p(gamma), _p(beta), _p(v), _p(mean), _p(rstd), T, out_f,
                                                    eps, c.stream), "rpde_layernorm_fwd")
                y = v
                if final_act == ACT_NONE:
                    v = None  # residual already added; nothing else to keep
            if final_act != ACT_NONE or (res is not None and not use_ln):
                src = v if use_ln else u
                y = torch.empty_like(u)
                _lib.check(c.lib.rpde_act_fwd(_p(src), _p(res), _p(y), y.numel(), final_act, c.stream), "rpde_act_fwd")
        ctx.n_layers, ctx.use_ln, ctx.final_act, ctx.T = n_layers, use_ln, final_act, T
        ctx.has_res = residual is not None
        ctx.nz = len(zs)
        pre = (v if use_ln else u) if final_act != ACT_NONE else None
        ctx.save_for_backward(*Ws, gamma, u if use_ln else None, mean, rstd, pre, *(acts if n_layers else []), *zs)
        ctx.has_bias = [b is not None for b in bs]
        return y

    @staticmethod
    def backward(ctx, dy):
        n = ctx.n_layers
        saved = ctx.saved_tensors
        Ws = saved[:n]
        gamma, u, mean, rstd, pre = saved[n:n + 5]
        acts = saved[n + 5:n + 5 + n]
        zs = saved[n + 5 + n:]
        T = ctx.T
        dy = dy.contiguous()
        dev = dy.device
        grads = [None] * (2 * n + (2 if ctx.use_ln else 0))
        dres = dy if ctx.has_res else None
        db_next = None  # bias gradient of layer i, when an earlier kernel already reduced the columns of g
        with _Ctx(dy) as c:
            g = dy
            if ctx.final_act != ACT_NONE:
                t = torch.empty_like(dy)
                _lib.check(c.lib.rpde_act_bwd(_p(g), _p(pre), _p(t), t.numel(), ctx.final_act, c.stream), "rpde_act_bwd")
                g = t
            if ctx.use_ln:
                Cn = gamma.numel()
                du = torch.empty_like(dy)
                dgamma = torch.empty_like(gamma)
                dbeta = torch.empty_like(gamma)
                # the column sums of du are the bias gradient of the last Linear: produced in the same pass
                if n > 0 and ctx.has_bias[n - 1]:
                    db_next = torch.empty(Cn, dtype=torch.float32, device=dev)
                _lib.check(c.lib.rpde_layernorm_bwd(_p(g), _p(u), _p(gamma), _p(mean), _p(rstd), _p(du), _p(dgamma), _p(dbeta),
                                                    _p(db_next), T, Cn, c.stream), "rpde_layernorm_bwd")
                grads[2 * n], grads[2 * n + 1] = dgamma, dbeta
                g = du
            dx = g if (n == 0 and ctx.needs_input_grad[0]) else None
            for i in range(n - 1, -1, -1):
                out_f, in_f = Ws[i].shape
                dw = torch.empty_like(Ws[i])
                if db_next is not None:
                    db, fused = db_next, True
                else:
                    db, fused = (torch.empty(out_f, dtype=torch.float32, device=dev) if ctx.has_bias[i] else None), False
                db_next = None
                _lib.check(c.lib.rpde_linear_bwd_params(_p(g), _p(acts[i]), _p(dw), None if fused else _p(db), T, in_f, out_f, 0,
                                                        c.stream), "rpde_linear_bwd_params")
                grads[2 * i], grads[2 * i + 1] = dw, db
                if i > 0 or ctx.needs_input_grad[0]:
                    gin = torch.empty((T, in_f), dtype=torch.float32, device=dev)
                    zprev = zs[i - 1] if i > 0 else None
                    if i > 0 and ctx.has_bias[i - 1]:
                        db_next = torch.empty(in_f, dtype=torch.float32, device=dev)
                    _lib.check(c.lib.rpde_linear_bwd_input(_p(g), _p(Ws[i]), _p(zprev), DACT_SAVED if i > 0 else 0, _p(gin), 0,
                                                           _p(db_next), T, in_f, out_f, 0, c.stream), "rpde_linear_bwd_input")
                    g = gin
                    if i == 0:
                        dx = gin.reshape(acts[0].shape)
        return (dx, dres, None, None, None, None, None, *grads)


def feed_forward(x, residual, weights, biases, ln_weight=None, ln_bias=None, final_act: int = ACT_NONE, eps: float = 1e-5):
    n = len(weights)
    if _ff3_applies(x, weights, biases, final_act):
        return _FusedFF3Fn.apply(x, residual, ln_weight is not None, eps, torch.is_grad_enabled(), weights[0], biases[0],
                                 weights[1], biases[1], weights[2], biases[2], ln_weight, ln_bias)
    params = []
    for w, b in zip(weights, biases):
        params += [w, b]
    use_ln = ln_weight is not None
    if use_ln:
        params += [ln_weight, ln_bias]
    return _FeedForwardFn.apply(x, residual, n, use_ln, final_act, eps, torch.is_grad_enabled(), *params)


# ---------------------------------------------------------------------------------------------
# pointwise channels-first conv (+ fused GELU MLP), weight norm, grid concat, loss
# ---------------------------------------------------------------------------------------------
class _Conv1x1Fn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, w, b):
        _need_cuda(x, w, b)
        x = x.contiguous()
        O, I = w.shape[0], w.shape[1]
        B = x.shape[0]
        spatial = tuple(x.shape[2:])
        S = int(np.prod(spatial))
        w2 = w.detach().reshape(O, I).contiguous()
        y = torch.empty((B, O) + spatial, dtype=torch.float32, device=x.device)
        with _Ctx(x) as c:
            _lib.check(c.lib.rpde_conv1x1_fwd(_p(x), _p(w2), _p(b), _p(y), None, B, I, O, S, ACT_NONE, 0, c.stream),
                       "rpde_conv1x1_fwd")
        ctx.save_for_backward(x, w)
        ctx.cfg = (B, I, O, S, b is not None)
        return y

    @staticmethod
    def backward(ctx, dy):
        x, w = ctx.saved_tensors
        B, I, O, S, has_b = ctx.cfg
        dy = dy.contiguous()
        w2 = w.detach().reshape(O, I).contiguous()
        dx = dw = db = None
        with _Ctx(x) as c:
            if ctx.needs_input_grad[0]:
                dx = torch.empty_like(x)
                _lib.check(c.lib.rpde_conv1x1_bwd_input(_p(dy), _p(w2), None, 0, _p(dx), 0, B, I, O, S, c.stream),
                           "rpde_conv1x1_bwd_input")
            dw, db = _dw_db(w, O, has_b)
            _lib.check(c.lib.rpde_conv1x1_bwd_params(_p(dy), _p(x), _p(dw), _p(db), B, I, O, S, c.stream),
                       "rpde_conv1x1_bwd_params")
        return dx, dw, db


def conv1x1(x, w, b=None):
    return _Conv1x1Fn.apply(x, w, b)


_mlp_fused = True   # tests flip it to run the composed projection MLP (two 1x1 convs through the contraction engine)


class _MLPFn(torch.autograd.Function):
    """projection MLP (fno_blocks.py:35-45, 73-83): conv1x1 -> GELU -> conv1x1, GELU fused in both directions."""

    @staticmethod
    def forward(ctx, x, w1, b1, w2, b2, grad_on=True):
        _need_cuda(x, w1, b1, w2, b2)
        x = x.contiguous()
        B, I = x.shape[0], x.shape[1]
        spatial = tuple(x.shape[2:])
        S = int(np.prod(spatial))
        H, O = w1.shape[0], w2.shape[0]
        w1c = w1.detach().reshape(H, I).contiguous()
        w2c = w2.detach().reshape(O, H).contiguous()
        need_grad = grad_on and any(ctx.needs_input_grad)
        y = torch.empty((B, O) + spatial, dtype=torch.float32, device=x.device)
        with _Ctx(x) as c:
            if _mlp_fused and c.lib.rpde_mlp_supported(I, H, O, S):
                # one kernel, hidden layer in registers (csrc/fno2d_mma.cuh); backward recomputes it from x
                _lib.check(c.lib.rpde_mlp_fwd(_p(x), _p(w1c), _p(b1), _p(w2c), _p(b2), _p(y), B, I, H, O, S, c.stream),
                           "rpde_mlp_fwd")
                ctx.save_for_backward(x, w1, w2, b1)
                ctx.cfg = (B, I, H, O, S, b1 is not None, b2 is not None, True)
                return y
        h = torch.empty((B, H) + spatial, dtype=torch.float32, device=x.device)
        z = torch.empty_like(h) if need_grad else None
        with _Ctx(x) as c:
            _lib.check(c.lib.rpde_conv1x1_fwd(_p(x), _p(w1c), _p(b1), _p(h), _p(z), B, I, H, S, ACT_GELU, 0, c.stream),
                       "rpde_conv1x1_fwd")
            _lib.check(c.lib.rpde_conv1x1_fwd(_p(h), _p(w2c), _p(b2), _p(y), None, B, H, O, S, ACT_NONE, 0, c.stream),
                       "rpde_conv1x1_fwd")
        ctx.save_for_backward(x, w1, w2, h, z)
        ctx.cfg = (B, I, H, O, S, b1 is not None, b2 is not None, False)
        return y

    @staticmethod
    def backward(ctx, dy):
        B, I, H, O, S, hb1, hb2, fused = ctx.cfg
        dy = dy.contiguous()
        dev = dy.device
        if fused:
            x, w1, w2, b1 = ctx.saved_tensors
            w1c = w1.detach().reshape(H, I).contiguous()
            w2c = w2.detach().reshape(O, H).contiguous()
            dx = torch.empty_like(x) if ctx.needs_input_grad[0] else None
            # dw1 | db1 | dw2 | db2 back to back: the kernel accumulates them with atomics and zero-fills them with one memset
            flat = torch.empty(H * I + H + O * H + O, dtype=torch.float32, device=dev)
            dw1 = flat[:H * I].view_as(w1)
            db1f = flat[H * I:H * I + H]
            dw2 = flat[H * I + H:H * I + H + O * H].view_as(w2)
            db2f = flat[H * I + H + O * H:]
            db1, db2 = (db1f if hb1 else None), (db2f if hb2 else None)
            if not (hb1 and hb2):       # a missing bias breaks the back-to-back layout the C side looks for: separate tensors
                dw1, dw2 = torch.empty_like(w1), torch.empty_like(w2)
                db1 = torch.empty(H, dtype=torch.float32, device=dev) if hb1 else None
                db2 = torch.empty(O, dtype=torch.float32, device=dev) if hb2 else None
            with _Ctx(x) as c:
                _lib.check(c.lib.rpde_mlp_bwd(_p(x), _p(dy), _p(w1c), _p(b1), _p(w2c), _p(dx), _p(dw1), _p(db1), _p(dw2), _p(db2),
                                              B, I, H, O, S, c.stream), "rpde_mlp_bwd")
            return dx, dw1, db1, dw2, db2, None
        x, w1, w2, h, z = ctx.saved_tensors
        w1c = w1.detach().reshape(H, I).contiguous()
        w2c = w2.detach().reshape(O, H).contiguous()
        dx = None
        with _Ctx(x) as c:
            dw2 = torch.empty_like(w2)
            db2 = torch.empty(O, dtype=torch.float32, device=dev) if hb2 else None
            _lib.check(c.lib.rpde_conv1x1_bwd_params(_p(dy), _p(h), _p(dw2), _p(db2), B, H, O, S, c.stream),
                       "rpde_conv1x1_bwd_params")
            dz = torch.empty_like(h)
            _lib.check(c.lib.rpde_conv1x1_bwd_input(_p(dy), _p(w2c), _p(z), ACT_GELU, _p(dz), 0, B, H, O, S, c.stream),
                       "rpde_conv1x1_bwd_input")
            dw1 = torch.empty_like(w1)
            db1 = torch.empty(H, dtype=torch.float32, device=dev) if hb1 else None
            _lib.check(c.lib.rpde_conv1x1_bwd_params(_p(dz), _p(x), _p(dw1), _p(db1), B, I, H, S, c.stream),
                       "rpde_conv1x1_bwd_params")
            if ctx.needs_input_grad[0]:
                dx = torch.empty_like(x)
                _lib.check(c.lib.rpde_conv1x1_bwd_input(_p(dz), _p(w1c), None, 0, _p(dx), 0, B, I, H, S, c.stream),
                           "rpde_conv1x1_bwd_input")
        return dx, dw1, db1, dw2, db2, None


def mlp_gelu(x, w1, b1, w2, b2):
    return _MLPFn.apply(x, w1, b1, w2, b2, torch.is_grad_enabled())


class _WeightNormFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, v, g):
        _need_cuda(v, g)
        v = v.contiguous()
        g = g.contiguous()
        O, I = v.shape
        w = torch.empty_like(v)
        norms = torch.empty(O, dtype=torch.float32, device=v.device)
        with _Ctx(v) as c:
            _lib.check(c.lib.rpde_weight_norm_fwd(_p(v), _p(g), _p(w), _p(norms), O, I, c.stream), "rpde_weight_norm_fwd")
        ctx.save_for_backward(v, g, norms)
        return w

    @staticmethod
    def backward(ctx, dw):
        v, g, norms = ctx.saved_tensors
        O, I = v.shape
        dw = dw.contiguous()
        dv = torch.empty_like(v)
        dg = torch.empty_like(g)
        with _Ctx(v) as c:
            _lib.check(c.lib.rpde_weight_norm_bwd(_p(dw), _p(v), _p(g), _p(norms), _p(dv), _p(dg), O, I, c.stream),
                       "rpde_weight_norm_bwd")
        return dv, dg


def weight_norm(v, g):
    return _WeightNormFn.apply(v, g)


class _ConcatGridFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, g0, g1, channels_last):
        _need_cuda(x, g0, g1)
        x = x.contiguous()
        B, Cin = x.shape[0], x.shape[1]
        S0 = x.shape[2]
        S1 = x.shape[3] if x.dim() == 4 else 1
        G = (g0 is not None) + (g1 is not None)
        if g0 is not None and g0.numel() != S0 or g1 is not None and g1.numel() != S1:
            raise RuntimeError("coordinate grid length does not match the input resolution")
        sp = tuple(x.shape[2:])
        shape = (B,) + sp + (Cin + G,) if channels_last else (B, Cin + G) + sp
        out = torch.empty(shape, dtype=torch.float32, device=x.device)
        with _Ctx(x) as c:
            _lib.check(c.lib.rpde_concat_grid_fwd(_p(x), _p(g0), _p(g1), _p(out), B, Cin, S0, S1, int(channels_last), c.stream),
                       "rpde_concat_grid_fwd")
        ctx.cfg = (B, Cin, S0, S1, G, int(channels_last), tuple(x.shape))
        return out

    @staticmethod
    def backward(ctx, dout):
        B, Cin, S0, S1, G, cl, xshape = ctx.cfg
        if not ctx.needs_input_grad[0]:
            return None, None, None, None
        dout = dout.contiguous()
        dx = torch.empty(xshape, dtype=torch.float32, device=dout.device)
        with _Ctx(dout) as c:
            _lib.check(c.lib.rpde_concat_grid_bwd(_p(dout), _p(dx), B, Cin, S0, S1, G, cl, c.stream), "rpde_concat_grid_bwd")
        return dx, None, None, None


def concat_grid(x, g0=None, g1=None, channels_last=False):
    return _ConcatGridFn.apply(x, g0, g1, channels_last)


class _RelL2Fn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, y, mode):
        _need_cuda(x, y)
        if x.shape != y.shape:
            raise RuntimeError(f"RelativeL2Loss: shape mismatch {tuple(x.shape)} vs {tuple(y.shape)}")
        x = x.contiguous()
        y = y.contiguous()
        B = x.shape[0]
        S = x.numel() // B
        acc = torch.empty(2 * B, dtype=torch.float32, device=x.device)
        per = torch.empty(B, dtype=torch.float32, device=x.device)
        loss = torch.empty((), dtype=torch.float32, device=x.device)
        with _Ctx(x) as c:
            _lib.check(c.lib.rpde_rel_l2_fwd(_p(x), _p(y), _p(acc), _p(per), _p(loss), B, S, mode, c.stream), "rpde_rel_l2_fwd")
        ctx.save_for_backward(x, y, acc)
        ctx.cfg = (B, S, mode)
        return loss if mode else per

    @staticmethod
    def backward(ctx, g):
        x, y, acc = ctx.saved_tensors
        B, S, mode = ctx.cfg
        if ctx.needs_input_grad[1]:
            raise NotImplementedError("RelativeL2Loss: gradient w.r.t. the target is not part of the training path")
        g = g.contiguous().to(torch.float32)
        dx = torch.empty_like(x)
        with _Ctx(x) as c:
            _lib.check(c.lib.rpde_rel_l2_bwd(_p(x), _p(y), _p(acc), _p(g), _p(dx), B, S, mode, c.stream), "rpde_rel_l2_bwd")
        return dx, None, None


def relative_l2(x, y, size_average=True, reduction=True):
    mode = 0 if not reduction else (1 if size_average else 2)
    return _RelL2Fn.apply(x, y, mode)


def matmul(a: torch.Tensor, b: torch.Tensor, trans_a=False, trans_b=False) -> torch.Tensor:
    """test hook for the contraction engines: row-major C = op(A) op(B)"""
    _need_cuda(a, b)
    a, b = a.contiguous(), b.contiguous()
    M = a.shape[1] if trans_a else a.shape[0]
    K = a.shape[0] if trans_a else a.shape[1]
    N = b.shape[0] if trans_b else b.shape[1]
    out = torch.empty((M, N), dtype=torch.float32, device=a.device)
    with _Ctx(a) as c:
        _lib.check(c.lib.rpde_matmul(_p(a), _p(b), _p(out), M, N, K, int(trans_a), int(trans_b), c.stream), "rpde_matmul")
    return out
