"""CPU-side checks: the C-ABI library exports every symbol include/rpde_b200.h declares, the drop-in modules
keep the reference's constructor signatures / state-dict layout (checked against the reference-generated
fixtures), and the product path refuses to run without CUDA (no CPU fallback)."""
import glob
import inspect
import os
import re

import pytest
import torch

from helpers import GOLD, build_candidate, load_fixture

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CASES = sorted(os.path.basename(p) for p in glob.glob(os.path.join(GOLD, "*.npz")) if "relative_l2" not in p)


def test_library_exports_every_declared_symbol():
    import resolution_pde_b200 as R
    hdr = open(os.path.join(ROOT, "include", "rpde_b200.h")).read()
    declared = set(re.findall(r"\b(rpde_[a-z0-9_]+)\s*\(", hdr))
    declared.discard("rpde_stream_t")
    assert len(declared) >= 30
    lib = R._lib.load()
    for name in sorted(declared):
        assert hasattr(lib, name), name
    assert declared == set(R._lib.SIGNATURES), declared ^ set(R._lib.SIGNATURES)
    assert lib.rpde_version() >= 100  # host-only call, no GPU needed


@pytest.mark.parametrize("name", CASES)
def test_state_dict_layout_matches_reference(name):
    _, params, _, hyper = load_fixture(name)
    m, _ = build_candidate(name, params, hyper)
    sd = m.state_dict()
    assert list(sd.keys()) == list(params.keys())
    for k, v in params.items():
        assert sd[k].shape == v.shape and sd[k].dtype == v.dtype, k
    m.load_state_dict(params, strict=True)
    for k, v in m.state_dict().items():
        assert torch.equal(v, params[k])


def test_constructor_signatures_match_reference():
    import resolution_pde_b200 as R
    want = {
        "SpectralConv1d": ["in_channels", "out_channels", "modes1"],
        "SpectralConv2d": ["in_channels", "out_channels", "modes1", "modes2"],
        "FSpectralConv1d": ["d_model", "modes", "forecast_ff", "backcast_ff", "fourier_weight", "factor", "ff_weight_norm",
                            "n_ff_layers", "layer_norm", "use_fork", "dropout", "mode", "activation", "fft_norm", "kwargs"],
        "FSpectralConv2d": ["d_model", "modes", "forecast_ff", "backcast_ff", "fourier_weight", "factor", "ff_weight_norm",
                            "n_ff_layers", "layer_norm", "use_fork", "dropout", "mode"],
        "FNO1d": ["in_channels", "out_channels", "modes", "width", "grid", "activation", "n_blocks"],
        "FNO2d": ["in_channels", "out_channels", "modes1", "modes2", "width", "grid", "activation", "n_blocks"],
        "FFNO1D": ["in_channels", "out_channels", "width", "n_layers", "n_modes", "factor", "ff_weight_norm", "n_ff_layers",
                   "layer_norm", "dropout", "grid", "mode", "fft_norm", "activation", "use_grid"],
        "FFNO2D": ["in_channels", "out_channels", "width", "n_layers", "n_modes", "factor", "ff_weight_norm", "n_ff_layers",
                   "layer_norm", "grid", "dropout", "mode", "use_grid"],
        "FeedForward": ["dim", "factor", "n_layers", "ff_weight_norm", "layer_norm", "dropout"],
        "RelativeL2Loss": ["size_average", "reduction"],
    }
    for cls, args in want.items():
        got = [p for p in inspect.signature(getattr(R, cls).__init__).parameters if p != "self"]
        assert got == args, (cls, got)


def test_ffno1d_use_grid_quirk():
    import resolution_pde_b200 as R
    # ffno.py:51: use_grid is overwritten by `grid`, so grid=None never adds the coordinate channel
    m = R.FFNO1D(1, 1, width=8, n_layers=1, n_modes=4, use_grid=True)
    assert m.in_proj.in_features == 1
    m = R.FFNO1D(1, 1, width=8, n_layers=1, n_modes=4, grid=[0.1 * i for i in range(16)])
    assert m.in_proj.in_features == 2


def test_no_cpu_fallback():
    import resolution_pde_b200 as R
    m = R.SpectralConv1d(2, 2, 3)
    with pytest.raises(RuntimeError, match="CUDA"):
        m(torch.randn(1, 2, 16))
    with pytest.raises(RuntimeError, match="CUDA"):
        R.RelativeL2Loss()(torch.randn(2, 4), torch.randn(2, 4))


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "resolution_pde_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in src and "/root/reference" not in src, f


def test_fused_adamw_host_contract():
    """Constructor mirrors torch.optim.AdamW; unsupported switches raise; CPU tensors raise at step (no fallback)."""
    import pytest
    import torch
    import resolution_pde_b200 as R
    p = torch.nn.Parameter(torch.randn(5))
    with pytest.raises(ValueError):
        R.FusedAdamW([p], amsgrad=True)
    with pytest.raises(ValueError):
        R.FusedAdamW([p], betas=(1.0, 0.999))
    opt = R.FusedAdamW([p], lr=2e-3, weight_decay=0.1)
    ref = torch.optim.AdamW([torch.nn.Parameter(torch.randn(5))], lr=2e-3, weight_decay=0.1)
    for k in ("lr", "betas", "eps", "weight_decay"):
        assert opt.param_groups[0][k] == ref.param_groups[0][k]
    assert set(ref.param_groups[0]) <= set(opt.param_groups[0]) | {"params"}   # state dicts load both ways
    opt.step()                                                                  # no gradients yet: nothing to do
    p.grad = torch.randn(5)
    with pytest.raises(RuntimeError):
        opt.step()
