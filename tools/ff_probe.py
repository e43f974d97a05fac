"""Bring-up probe of the fused BF16 FeedForward kernels (rpde_ff3_*): calls the C ABI directly and compares the output, every
gradient AND every intermediate the backward kernel leaves in its workspace (BF16 copies of x, H1, H2, dZ1, dZ2, dZ3)
against a torch fp32 evaluation on the same device.  `python tools/ff_probe.py [T] [--no-ln] [--time]`"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import resolution_pde_b200 as R  # noqa: E402
from resolution_pde_b200 import _lib  # noqa: E402


def rel(a, b):
    a, b = a.double(), b.double()
    return ((a - b).norm() / b.norm().clamp_min(1e-30)).item()


def main():
    T = int(sys.argv[1]) if len(sys.argv) > 1 and sys.argv[1].isdigit() else 128 * 5 + 37
    use_ln = "--no-ln" not in sys.argv
    dev = torch.device("cuda:0")
    lib = _lib.load()
    torch.manual_seed(0)
    C, H = 64, 256
    x = torch.randn(T, C, device=dev)
    res = torch.randn(T, C, device=dev)
    dy = torch.randn(T, C, device=dev)
    w1 = torch.randn(H, C, device=dev) / 8
    w2 = torch.randn(H, H, device=dev) / 16
    w3 = torch.randn(C, H, device=dev) / 16
    b1, b2, b3 = torch.randn(H, device=dev) * 0.1, torch.randn(H, device=dev) * 0.1, torch.randn(C, device=dev) * 0.1
    gamma, beta = 1 + 0.1 * torch.randn(C, device=dev), 0.1 * torch.randn(C, device=dev)
    st = torch.cuda.current_stream().cuda_stream
    p = lambda t: None if t is None else t.data_ptr()

    # reference (fp32 torch, weights/inputs as given)
    params = [t.clone().requires_grad_(True) for t in (x, w1, b1, w2, b2, w3, b3, gamma, beta)]
    xr, w1r, b1r, w2r, b2r, w3r, b3r, gr, br = params
    z1 = xr @ w1r.t() + b1r
    h1 = torch.nn.functional.gelu(z1)
    z2 = h1 @ w2r.t() + b2r
    h2 = torch.nn.functional.gelu(z2)
    z3 = h2 @ w3r.t() + b3r
    for t in (z1, z2, z3, h1, h2):
        t.retain_grad()
    u = torch.nn.functional.layer_norm(z3, (C,), gr, br, 1e-5) if use_ln else z3
    yr = res + u
    yr.backward(dy)

    img = torch.empty(lib.rpde_ff3_weight_image_bytes(), dtype=torch.uint8, device=dev)
    _lib.check(lib.rpde_ff3_pack_weights(p(w1), p(w2), p(w3), p(img), st), "pack")
    y = torch.empty_like(x)
    _lib.check(lib.rpde_ff3_fwd(p(x), p(res), p(img), p(b1), p(b2), p(b3), p(gamma) if use_ln else None, p(beta) if use_ln else None,
                                p(y), T, 1e-5, st), "fwd")
    torch.cuda.synchronize()
    print(f"T={T} ln={use_ln}")
    print("fwd  y      rel err", rel(y, yr.detach()), " (y - res) rel err", rel(y - res, (yr - res).detach()))

    f = lambda *s: torch.full(s, float("nan"), device=dev)
    ds, dw1, dw2, dw3, db1, db2, db3, dg, dbt = f(T, C), f(H, C), f(H, H), f(C, H), f(H), f(H), f(C), f(C), f(C)
    nb = lib.rpde_ff3_bwd_workspace_bytes(T)
    ws = torch.zeros(nb, dtype=torch.uint8, device=dev)
    _lib.check(lib.rpde_ff3_bwd(p(x), p(dy), p(img), p(b1), p(b2), p(b3), p(gamma) if use_ln else None, p(ds), p(dw1), p(db1), p(dw2),
                                p(db2), p(dw3), p(db3), p(dg) if use_ln else None, p(dbt) if use_ln else None, T, 1e-5, p(ws), nb, st), "bwd")
    torch.cuda.synchronize()
    up = lambda v: (v + 255) // 256 * 256
    off = 0

    def take(cols):
        nonlocal off
        n = T * cols * 2
        t = ws[off:off + n].view(torch.bfloat16).reshape(T, cols).float()
        off += up(n)
        return t

    xb, dz3s, h1s, h2s, dz1s, dz2s = take(C), take(C), take(H), take(H), take(H), take(H)
    print("bwd  xb     rel err", rel(xb, x))
    print("bwd  H1     rel err", rel(h1s, h1.detach()))
    print("bwd  H2     rel err", rel(h2s, h2.detach()))
    print("bwd  dZ3    rel err", rel(dz3s, z3.grad))
    print("bwd  dZ2    rel err", rel(dz2s, z2.grad))
    print("bwd  dZ1    rel err", rel(dz1s, z1.grad))
    print("bwd  ds     rel err", rel(ds, xr.grad))
    for name, got, want in (("dw1", dw1, w1r.grad), ("dw2", dw2, w2r.grad), ("dw3", dw3, w3r.grad), ("db1", db1, b1r.grad),
                            ("db2", db2, b2r.grad), ("db3", db3, b3r.grad)) + ((("dgamma", dg, gr.grad), ("dbeta", dbt, br.grad)) if use_ln else ()):
        print(f"bwd  {name:6s} rel err", rel(got, want))

    if "--time" in sys.argv:
        for name, fn in (("fwd", lambda: lib.rpde_ff3_fwd(p(x), p(res), p(img), p(b1), p(b2), p(b3), p(gamma), p(beta), p(y), T, 1e-5, st)),
                         ("bwd", lambda: lib.rpde_ff3_bwd(p(x), p(dy), p(img), p(b1), p(b2), p(b3), p(gamma), p(ds), p(dw1), p(db1), p(dw2),
                                                          p(db2), p(dw3), p(db3), p(dg), p(dbt), T, 1e-5, p(ws), nb, st))):
            for _ in range(3):
                fn()
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record()
            for _ in range(10):
                fn()
            e.record()
            torch.cuda.synchronize()
            print(f"time {name}: {s.elapsed_time(e) / 10:.3f} ms for T={T}")


if __name__ == "__main__":
    main()
