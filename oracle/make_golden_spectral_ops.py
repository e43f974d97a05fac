"""Generate tests/golden/specops/specops_*.npz from the UNMODIFIED reference utilities (build container only).

Run:  python oracle/make_golden_spectral_ops.py        (needs /root/reference; CPU only)
utils/low_pass_filter.py and utils/res_utils.py are imported as they are; seeded fp32 inputs, reference outputs.
"""
from __future__ import annotations

import importlib.util
import os

import numpy as np
import torch

REF = os.environ.get("RPDE_REFERENCE", "/root/reference")
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden", "specops")


def _stub_plotting():
    """frequency_error.py imports matplotlib at module top for its plotting helpers; the metric functions never use it"""
    import sys
    import types
    for name in ("matplotlib", "matplotlib.pyplot", "matplotlib.patches", "matplotlib.gridspec", "matplotlib.colors", "seaborn", "wandb"):
        if name not in sys.modules:
            try:
                __import__(name)
            except Exception:
                sys.modules[name] = types.ModuleType(name)


def _load(name, rel):
    spec = importlib.util.spec_from_file_location(name, os.path.join(REF, rel))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def main():
    lp = _load("ref_low_pass_filter", "utils/low_pass_filter.py")
    ru = _load("ref_res_utils", "utils/res_utils.py")
    g = torch.Generator().manual_seed(2024)
    cases = {}
    for shape, ratio in [((2, 3, 2, 64), 0.25), ((2, 3, 48), 0.5), ((1, 2, 1, 33), 0.3), ((2, 2, 32), 1.0), ((1, 1, 16), 0.05)]:
        x = torch.randn(*shape, generator=g)
        cases[f"lowpass1d_{'x'.join(map(str, shape))}_r{ratio}"] = dict(x=x, ratio=ratio, y=lp.lowpass_filter_1d(x.clone(), ratio))
    for shape, ratio in [((2, 2, 1, 32, 32), 0.25), ((1, 3, 16, 16), 0.5), ((1, 1, 2, 24, 24), 0.125), ((1, 2, 15, 15), 0.4),
                         ((1, 1, 16, 16), 1.0)]:
        x = torch.randn(*shape, generator=g)
        cases[f"lowpass2d_{'x'.join(map(str, shape))}_r{ratio}"] = dict(x=x, ratio=ratio, y=lp.lowpass_filter_2d(x.clone(), ratio))
    for shape, out in [((2, 3, 64), 32), ((2, 2, 32), 64), ((1, 4, 48), 48), ((3, 33), 20), ((2, 20), 33), ((2, 64), 16)]:
        x = torch.randn(*shape, generator=g)
        cases[f"resize1d_{'x'.join(map(str, shape))}_to{out}"] = dict(x=x, out=out, y=ru.resize_1d(x.clone(), out))
    for shape, out, perm in [((2, 2, 32, 32), (16, 16), False), ((1, 3, 16, 16), (32, 32), False), ((2, 1, 24, 16), (12, 32), False),
                             ((1, 16, 16, 2), (8, 8), True), ((1, 2, 16, 20), (16, 20), False), ((1, 1, 32, 30), (16, 15), False)]:
        x = torch.randn(*shape, generator=g)
        cases[f"resize2d_{'x'.join(map(str, shape))}_to{out[0]}x{out[1]}_p{int(perm)}"] = dict(
            x=x, out=np.array(out), permute=perm, y=ru.resize(x.clone(), out, permute=perm))
    for shape, n in [((2, 3, 32), 16), ((2, 2, 33), 16), ((3, 32), 15), ((2, 20), 8), ((2, 16), 16), ((1, 2, 16), 32), ((2, 9), 8)]:
        x = torch.randn(*shape, generator=g)
        cases[f"downsample1d_{'x'.join(map(str, shape))}_to{n}"] = dict(x=x, out=n, y=ru.downsample_1d(x.numpy().copy(), n))
    for shape, n in [((2, 3, 32, 32), 16), ((1, 2, 33, 33), 16), ((2, 1, 32, 32), 15), ((1, 1, 20, 20), 8), ((1, 2, 16, 16), 16),
                     ((1, 1, 9, 9), 4)]:
        x = torch.randn(*shape, generator=g)
        cases[f"downsample2d_{'x'.join(map(str, shape))}_to{n}"] = dict(x=x, out=n, y=ru.downsample(x.numpy().copy(), n))
    _stub_plotting()
    fe = _load("ref_frequency_error", "utils/frequency_error.py")
    for shape, nm in [((3, 2, 64), None), ((2, 1, 48), 10), ((2, 2, 33), None), ((1, 1, 32), 100)]:
        yh, yt = torch.randn(*shape, generator=g), torch.randn(*shape, generator=g)
        e, m, f = fe.decompose_error_by_frequency_1d(yh.clone(), yt.clone(), nm)
        cases[f"freqerr1d_{'x'.join(map(str, shape))}_m{nm}"] = dict(y_hat=yh, x=yt, num=-1 if nm is None else nm, err=e, mag=m, freqs=f)
    for shape, nb in [((2, 2, 32, 32), 16), ((1, 1, 16, 24), 64), ((2, 1, 24, 16), 8), ((1, 2, 64, 64), 64)]:
        yh, yt = torch.randn(*shape, generator=g), torch.randn(*shape, generator=g)
        e, m, f = fe.decompose_error_by_frequency_2d(yh.clone(), yt.clone(), nb)
        cases[f"freqerr2d_{'x'.join(map(str, shape))}_b{nb}"] = dict(y_hat=yh, x=yt, num=nb, err=e, mag=m, freqs=f)
    # odd heights (one more positive than negative row kept, res_utils.py:35-40); drawn last so earlier cases keep their inputs
    for shape, out in [((1, 2, 15, 16), (8, 8)), ((1, 1, 16, 16), (15, 12)), ((2, 1, 15, 15), (9, 9)), ((1, 1, 9, 12), (15, 16))]:
        x = torch.randn(*shape, generator=g)
        cases[f"resize2d_{'x'.join(map(str, shape))}_to{out[0]}x{out[1]}_p0"] = dict(
            x=x, out=np.array(out), permute=False, y=ru.resize(x.clone(), out, permute=False))
    os.makedirs(OUT, exist_ok=True)
    for name, c in cases.items():
        np.savez_compressed(os.path.join(OUT, f"specops_{name}.npz"),
                            **{k: (v.numpy() if torch.is_tensor(v) else np.asarray(v)) for k, v in c.items()})
    print(f"wrote {len(cases)} fixtures")


if __name__ == "__main__":
    main()
