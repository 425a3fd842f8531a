"""pccb_cc_energy_static: the allocation-free variant of pccb_cc_energy (static __device__ scratch, result in a
caller-owned device double; include/pyccb200.h).  It has no counterpart in the reference - it exists for callers
that cannot call cudaMalloc - so its check is against the existing entry point, through the C ABI, on two calls
in a row (the scratch array is shared by all calls and keeps the previous call's contents).

This file is named so that pytest collects it LAST: a failure here must
not hide any other test.  It passes alone in a fresh process and after the rest of the suite."""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _case(be, no, nv, seed, dtype):
    g = torch.Generator(device="cpu").manual_seed(seed)
    n = no + nv
    F = torch.randn(n, n, generator=g, dtype=torch.float64).to(dtype).to(be.device)
    t1 = (0.1 * torch.randn(no, nv, generator=g, dtype=torch.float64)).to(dtype).to(be.device)
    t2 = (0.1 * torch.randn(no, no, nv, nv, generator=g, dtype=torch.float64)).to(dtype).to(be.device)
    oovv = torch.randn(no, no, nv, nv, generator=g, dtype=torch.float64).to(dtype).to(be.device)
    return F, t1, t2, oovv


def _static(be, F, t1, t2, oovv, no, nv, out):
    from pycc_b200.backend import dtype_code
    fov = F[:no, no:]
    return be.lib.pccb_cc_energy_static(be.h, dtype_code(t2), fov.data_ptr(), F.stride(0), t1.data_ptr(), t2.data_ptr(),
                                        oovv.data_ptr(), no, nv, out.data_ptr())


@pytest.mark.parametrize("dtype", [torch.float64, torch.float32])
def test_cc_energy_static_matches_cc_energy_twice(dtype):
    from pycc_b200.backend import CudaBackend
    be = CudaBackend()
    be.bind_stream()
    out = torch.full((1,), float("nan"), dtype=torch.float64, device=be.device)
    # two calls in a row on DIFFERENT problems: a bigger one first, so that the second call runs over scratch
    # slots that still hold the first call's partial sums
    for (no, nv, seed) in ((7, 70, 1), (5, 33, 2)):
        F, t1, t2, oovv = _case(be, no, nv, seed, dtype)
        want = be.cc_energy(F[:no, no:], F.stride(0), t1, t2, oovv, no, nv)        # existing entry point
        out.fill_(float("nan"))
        assert _static(be, F, t1, t2, oovv, no, nv, out) == 0, be.lib.pccb_last_error(be.h).decode()
        be.synchronize()
        got = out.item()
        # same kernels, same fixed summation order: identical up to nothing; the suite's energy bar is 1e-12
        assert abs(got - want) <= 1e-12 * max(1.0, abs(want)), (got, want)
        # and against an independent FP64 formula (pycc/ccwfn.py:717-718) on the same stored values
        t1d, t2d, gd = t1.double(), t2.double(), oovv.double()
        tau = t2d + t1d[:, None, :, None] * t1d[None, :, None, :]
        ref = float(2.0 * (F[:no, no:].double() * t1d).sum() + (tau * (2.0 * gd - gd.transpose(2, 3))).sum())
        assert abs(got - ref) <= 1e-9 * max(1.0, abs(ref))
    # the same problem twice gives the same bits (the array is reused, nothing stale is summed)
    assert _static(be, F, t1, t2, oovv, no, nv, out) == 0
    be.synchronize()
    assert out.item() == got


def test_cc_energy_static_rejects_oversize():
    """ceil(nv/32)^2 * no^2 beyond the static array is an argument error with a message, not a fault; no device
    memory is touched (the pointers below are never dereferenced)."""
    from pycc_b200.backend import CudaBackend
    from pycc_b200 import _lib
    be = CudaBackend()
    cap = be.lib.pccb_cc_energy_static_max_partials()
    assert cap == 1 << 17
    no, nv = 40, 400                     # 13^2 * 1600 = 270400 > 131072
    assert ((nv + 31) // 32) ** 2 * no * no > cap
    dummy = torch.zeros(8, dtype=torch.float64, device=be.device)
    p = dummy.data_ptr()
    rc = be.lib.pccb_cc_energy_static(be.h, _lib.F64, p, no + nv, p, p, p, no, nv, p)
    assert rc != 0
    msg = be.lib.pccb_last_error(be.h).decode()
    assert "static scratch" in msg and "270400" in msg
