"""GPU test of rpde_rel_l2_fwd_static, the allocation-free variant of rpde_rel_l2_fwd (accumulator in a file-scope
__device__ array of the library instead of a caller buffer; include/rpde_b200.h).  Called through the C ABI with
ctypes, compared with the existing entry point on two calls in a row (the second with fewer, different samples, so
stale sums of the first call would show).  Tolerance 1e-5 relative L2, the bound of tests/test_gpu_cabi_ops.py: the
per-sample sums are fp32 atomics whose order differs between launches.

This file is collected last on purpose: a failure here hides no other test."""
import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
TOL = 1e-5
MAX_B = 4096  # RPDE_REL_L2_STATIC_MAX_B


def rel(a, b):
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    return ((a - b).norm() / b.norm().clamp_min(1e-30)).item()


def _st():
    return torch.cuda.current_stream().cuda_stream


def _both(lib, x, y, mode):
    """(per_sample, loss) from the existing entry point and from the static-scratch one; every buffer is allocated
    here, before the calls."""
    B, S = x.shape[0], x[0].numel()
    acc = torch.empty(2 * B, device=DEV)
    ps_a, loss_a = torch.full((B,), -1.0, device=DEV), torch.full((1,), -1.0, device=DEV)
    ps_b, loss_b = torch.full((B,), -2.0, device=DEV), torch.full((1,), -2.0, device=DEV)
    torch.cuda.synchronize()
    rc = lib.rpde_rel_l2_fwd(x.data_ptr(), y.data_ptr(), acc.data_ptr(), ps_a.data_ptr(), loss_a.data_ptr(), B, S, mode,
                             _st())
    assert rc == 0, lib.rpde_last_error()
    free0 = torch.cuda.mem_get_info()[0]
    rc = lib.rpde_rel_l2_fwd_static(x.data_ptr(), y.data_ptr(), ps_b.data_ptr(), loss_b.data_ptr(), B, S, mode, _st())
    assert rc == 0, lib.rpde_last_error()
    torch.cuda.synchronize()
    free1 = torch.cuda.mem_get_info()[0]
    return (ps_a, loss_a), (ps_b, loss_b), free0 - free1


def test_rel_l2_static_matches_caller_scratch_twice():
    import resolution_pde_b200 as R
    lib = R._lib.load()
    torch.cuda.set_device(0)
    torch.manual_seed(7)
    # call 1: 16 fields of 256 x 256, mean reduction; call 2: 5 ragged-size fields, sum reduction, other values
    x1, y1 = torch.randn(16, 1, 256, 256, device=DEV), torch.randn(16, 1, 256, 256, device=DEV)
    x2, y2 = 3.0 * torch.randn(5, 1000, device=DEV) + 1.0, 0.1 * torch.randn(5, 1000, device=DEV)
    for k, (x, y, mode) in enumerate([(x1, y1, 1), (x2, y2, 2), (x1, y1, 0)]):
        (ps_a, loss_a), (ps_b, loss_b), grew = _both(lib, x, y, mode)
        assert rel(ps_b, ps_a) < TOL, (k, rel(ps_b, ps_a))
        d = (x - y).double().flatten(1).norm(dim=1) / (y.double().flatten(1).norm(dim=1) + 1e-8)
        assert rel(ps_b, d) < TOL
        if mode:
            assert rel(loss_b, loss_a) < TOL, (k, loss_a.item(), loss_b.item())
            assert rel(loss_b, d.mean() if mode == 1 else d.sum()) < TOL
        else:
            assert loss_b.item() == -2.0  # reduce_mode 0 writes only per_sample
        if k:  # after the first call has loaded the kernels: the call takes no device memory
            assert grew <= 0, grew


def test_rel_l2_static_rejects_batches_beyond_the_array():
    import resolution_pde_b200 as R
    lib = R._lib.load()
    B = MAX_B + 1
    x, y = torch.randn(B, 8, device=DEV), torch.randn(B, 8, device=DEV)
    ps, loss = torch.zeros(B, device=DEV), torch.full((1,), 5.0, device=DEV)
    rc = lib.rpde_rel_l2_fwd_static(x.data_ptr(), y.data_ptr(), ps.data_ptr(), loss.data_ptr(), B, 8, 1, _st())
    assert rc == -2, rc  # RPDE_E_WORKSPACE
    assert b"RPDE_REL_L2_STATIC_MAX_B" in lib.rpde_last_error()
    torch.cuda.synchronize()
    assert loss.item() == 5.0 and ps.abs().max().item() == 0.0  # nothing was launched
    # the largest accepted batch still works
    (ps_a, loss_a), (ps_b, loss_b), _ = _both(lib, x[:MAX_B], y[:MAX_B], 1)
    assert rel(ps_b, ps_a) < TOL and rel(loss_b, loss_a) < TOL
