"""Complex-amplitude CCSD and Lambda residuals on the B200 (two real planes through the FP64
tensor-core kernels) against vectors from the reference's own code (tests/golden/complex/)."""
import contextlib
import io

import numpy as np
import pytest
import torch

from tests import complex_checks as cc
import pycc_b200

pytestmark = pytest.mark.gpu


def make_wfn(g, model):
    return pycc_b200.ccwfn.from_arrays(g.F, g.blocks, g.no, model=model)


@pytest.mark.parametrize("model", cc.MODELS)
@pytest.mark.parametrize("name", cc.CASES)
def test_complex_residuals(name, model):
    cc.check_complex_residuals(make_wfn, name, model)


def test_complex_residual_is_holomorphic_mid_size():
    """r(t) is a polynomial in the amplitudes, so r(conj t) = conj r(t) for real integrals and a real
    Fock matrix; and a purely real complex call equals the real path.  (o, v) = (10, 72)."""
    from pycc_b200.synthetic import SyntheticSystem
    s = SyntheticSystem(10, 72, seed=5)
    with contextlib.redirect_stdout(io.StringIO()):
        w = pycc_b200.ccwfn.from_arrays(s.F, s.blocks, s.no)
    gen = torch.Generator(device="cpu").manual_seed(11)
    t1 = (0.05 * torch.randn(10, 72, 2, generator=gen, dtype=torch.float64)).cuda()
    t2 = (0.02 * torch.randn(10, 10, 72, 72, 2, generator=gen, dtype=torch.float64)).cuda()
    t1c, t2c = torch.view_as_complex(t1), torch.view_as_complex(t2)
    r1, r2 = w.residuals(s.F, t1c, t2c, real_time=True)
    q1, q2 = w.residuals(s.F, t1c.conj().resolve_conj(), t2c.conj().resolve_conj(), real_time=True)
    assert (q1 - r1.conj()).abs().max().item() < 1e-12
    assert (q2 - r2.conj()).abs().max().item() < 1e-12
    x1, x2 = w.residuals(s.F, t1c.real.contiguous(), t2c.real.contiguous())
    y1, y2 = w.residuals(s.F, torch.complex(t1c.real, torch.zeros_like(t1c.real)),
                         torch.complex(t2c.real, torch.zeros_like(t2c.real)))
    assert (y2.real - x2).abs().max().item() < 1e-12 and y2.imag.abs().max().item() == 0.0
    assert (y1.real - x1).abs().max().item() < 1e-12
