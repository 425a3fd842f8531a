"""HBAR and Lambda on the B200 through the C ABI, against reference-generated vectors
(tests/golden/lambda/, made by tests/golden/make_golden_lambda.py), plus a size-independent
property at a shape no CPU reference covers."""
import contextlib
import io

import numpy as np
import pytest
import torch

from tests import lambda_checks as lc
import pycc_b200

pytestmark = pytest.mark.gpu


def make_wfn(g, model):
    return pycc_b200.ccwfn.from_arrays(g.F, g.blocks, g.no, model=model)


@pytest.mark.parametrize("model", lc.MODELS)
@pytest.mark.parametrize("name", lc.CASES)
def test_hbar_blocks_and_lambda_residuals(name, model):
    lc.check_blocks_and_residuals(make_wfn, name, model)


@pytest.mark.parametrize("model", lc.MODELS)
@pytest.mark.parametrize("name", lc.CASES)
def test_solve_lambda(name, model):
    lc.check_solve(make_wfn, name, model, chain=False, e_tol=1e-10)


@pytest.mark.parametrize("name", ["o4v8", "o5v19_offdiag"])
def test_solve_chain(name):
    lc.check_solve(make_wfn, name, "CCSD", chain=True, e_tol=1e-10)


def test_lambda_linearity_mid_size():
    """The Lambda residual is affine in (l1, l2): r(a x + (1-a) y) = a r(x) + (1-a) r(y), and the
    tensor-core path agrees with the library's scalar contraction path, at (o, v) = (10, 72)."""
    from pycc_b200.synthetic import SyntheticSystem
    s = SyntheticSystem(10, 72, seed=5)
    with contextlib.redirect_stdout(io.StringIO()):
        w = pycc_b200.ccwfn.from_arrays(s.F, s.blocks, s.no)
        for it in range(3):
            w._iteration(None, it + 1, 1, extrapolate=False)
        hb = pycc_b200.cchbar(w)
        lam = pycc_b200.cclambda(w, hb)
    gen = torch.Generator(device="cpu").manual_seed(3)
    x1, y1 = [torch.randn(lam.l1.shape, generator=gen, dtype=torch.float64).cuda() * 0.1 for _ in range(2)]
    x2, y2 = [torch.randn(lam.l2.shape, generator=gen, dtype=torch.float64).cuda() * 0.1 for _ in range(2)]
    t1, t2 = w.t1.clone(), w.t2.clone()
    a = 0.3
    rx = lam.residuals(w.H.F, t1, t2, x1, x2)
    ry = lam.residuals(w.H.F, t1, t2, y1, y2)
    rz = lam.residuals(w.H.F, t1, t2, a * x1 + (1 - a) * y1, a * x2 + (1 - a) * y2)
    for k in range(2):
        ref = a * rx[k] + (1 - a) * ry[k]
        assert (rz[k] - ref).abs().max().item() < 1e-11
    w.backend.use_naive_contract(True)
    try:
        hb.engine.build()
        rn = lam.residuals(w.H.F, t1, t2, x1, x2)
    finally:
        w.backend.use_naive_contract(False)
    for k in range(2):
        assert (rn[k] - rx[k]).abs().max().item() < 1e-11


@pytest.mark.parametrize("name", ["o3v5_offdiag", "o4v8", "o6v13"])
def test_t3_density_and_ccsd_t_lambda(name):
    """ccwfn.t3_density (S1, S2, Dvv, Doo, Dov, Goovv, Gooov, Gvvvo, E(T)) and the CCSD(T) Lambda equations on the
    device (pycc/ccwfn.py:721-855, pycc/cclambda.py:474-475, 519-520) against the reference's own vectors."""
    from pycc_b200.backend import CudaBackend
    from tests import t3density_checks as tc
    be = CudaBackend()
    tc.check_t3_density(name, be)
    tc.check_lambda_t(name, be)


@pytest.mark.parametrize("name", ["o3v5_offdiag", "o4v8", "o6v13"])
def test_cc3_lambda(name):
    """The CC3 Lambda equations on the device (pycc/cclambda.py:118-139, 160-235; cctriples.l3_ijk :219-277): one
    iteration at perturbed unsymmetric amplitudes and the solve_cc -> cchbar -> solve_lambda chain, against the
    reference's own vectors."""
    from pycc_b200.backend import CudaBackend
    from tests import lambda_cc3_checks as l3
    be = CudaBackend()
    l3.check_one_iteration(name, be)
    l3.check_solve(name, be)
    if name != "o6v13":
        l3.check_residuals_entry(name, be)


def test_sp_mode_round2_engines():
    """precision='SP' through the round-2 engines (t3_density, CC3 Lambda): FP32 storage and contractions, FP64 scalars;
    against the reference's FP64 vectors at single-precision tolerance."""
    import contextlib, io
    import numpy as np
    from tests import lambda_cc3_checks as l3, t3density_checks as tc
    g, z = l3.load("o4v8")
    with contextlib.redirect_stdout(io.StringIO()):
        w = pycc_b200.ccwfn.from_arrays(g.F, g.blocks, g.no, model="CC3", precision="SP")
        w._load(None, z["t1"], z["t2"])
        lam = pycc_b200.cclambda(w, pycc_b200.cchbar(w))
    lam._load(z["l1"], z["l2"])
    r1, r2 = lam.engine.residuals(symmetrize=True)
    assert str(r2.dtype) == "torch.float32"
    assert np.abs(r1.double().cpu().numpy() - z["r1"]).max() < 3e-5 and np.abs(r2.double().cpu().numpy() - z["r2"]).max() < 3e-5    # |r| ~ 1.5: FP32 rounding of the stored blocks
    g, z = tc.load("o4v8")
    with contextlib.redirect_stdout(io.StringIO()):
        w = pycc_b200.ccwfn.from_arrays(g.F, g.blocks, g.no, model="CCSD(T)", precision="SP")
        w._load(None, g["sym_t1"], g["sym_t2"])
        et = w.t3_density()
    assert abs(et - float(z["sym_ET"])) < 1e-7
    assert np.abs(w.S2.double().cpu().numpy() - z["sym_S2"]).max() < 1e-6
