"""Shared checks of the CC3 Lambda equations (pycc/cclambda.py:118-139, 160-235; cctriples.l3_ijk :219-277) against
tests/golden/lambda_cc3/*.npz (unmodified reference, tests/golden/make_golden_lambda_cc3.py), for the CPU
(interpreter back end) and GPU (C ABI) tests."""
import contextlib
import io
import os

import numpy as np

CASES = ("o3v5_offdiag", "o4v8")


def load(name):
    from tests.conftest import GOLDEN_DIR, Golden
    return Golden(name), np.load(os.path.join(GOLDEN_DIR, "lambda_cc3", name + ".npz"))


def _wfn(g, backend):
    import pycc_b200
    with contextlib.redirect_stdout(io.StringIO()):
        return pycc_b200.ccwfn.from_arrays(g.F, g.blocks, g.no, model="CC3", _backend=backend)


def check_one_iteration(name, backend, tol=1e-11):
    """Residuals (CCSD part + the CC3 triples terms Y1, Y2 + Y2^T) at perturbed, unsymmetric amplitudes."""
    import pycc_b200
    g, z = load(name)
    w = _wfn(g, backend)
    w._load(None, z["t1"], z["t2"])
    with contextlib.redirect_stdout(io.StringIO()):
        hb = pycc_b200.cchbar(w)
        lam = pycc_b200.cclambda(w, hb)
    lam._load(z["l1"], z["l2"])
    r1, r2 = lam.engine.residuals(symmetrize=True)
    r1, r2 = r1.cpu().numpy().copy(), r2.cpu().numpy().copy()
    # the CC3 terms are a visible part of the residual (the check is not dominated by the CCSD terms)
    lam.engine.cc3, keep = None, lam.engine.cc3
    s1, s2 = lam.engine.residuals(symmetrize=True)
    lam.engine.cc3 = keep
    assert np.abs(r1 - s1.cpu().numpy()).max() > 1e-3 and np.abs(r2 - s2.cpu().numpy()).max() > 1e-3
    assert np.abs(r1 - z["r1"]).max() < tol, np.abs(r1 - z["r1"]).max()
    assert np.abs(r2 - z["r2"]).max() < tol, np.abs(r2 - z["r2"]).max()


def check_solve(name, backend):
    """solve_cc -> cchbar -> solve_lambda from the MP2 guess: energies, iteration count, amplitudes."""
    import pycc_b200
    g, z = load(name)
    w = _wfn(g, backend)
    with contextlib.redirect_stdout(io.StringIO()):
        ecc = w.solve_cc(e_conv=1e-12, r_conv=1e-12, maxiter=100)
        hb = pycc_b200.cchbar(w)
        lam = pycc_b200.cclambda(w, hb)
        lecc = lam.solve_lambda(e_conv=1e-12, r_conv=1e-12, maxiter=100)
    assert abs(ecc - float(z["solve_ecc"])) < 1e-10
    assert len(lam.iteration_log) == int(z["solve_niter"])
    assert abs(lecc - float(z["solve_lecc"])) < 1e-10
    assert np.abs(lam.l1.cpu().numpy() - z["solve_l1"]).max() < 1e-9
    assert np.abs(lam.l2.cpu().numpy() - z["solve_l2"]).max() < 1e-9


def check_residuals_entry(name, backend, tol=1e-11):
    """cclambda.residuals(F, t1, t2, l1, l2) for CC3 (the real-time entry, pycc/cclambda.py:276-447) with a
    field-dressed Fock matrix, for a plain and a real_time wave function (t3_pert_ijk inside the t3 loops); the call is
    pure: a solve_lambda afterwards reproduces the reference's chain."""
    import pycc_b200
    from tests.conftest import GOLDEN_DIR
    g, zs = load(name)
    z = np.load(os.path.join(GOLDEN_DIR, "rtcc3", name + "_lambda.npz"))
    assert np.abs(z["rt_r1"] - z["plain_r1"]).max() > 1e-7
    for tag, rt in (("plain", False), ("rt", True)):
        with contextlib.redirect_stdout(io.StringIO()):
            w = pycc_b200.ccwfn.from_arrays(g.F, g.blocks, g.no, model="CC3", real_time=rt, _backend=backend)
            lam = pycc_b200.cclambda(w, pycc_b200.cchbar(w))
        r1, r2 = lam.residuals(z["F_field"], z["t1"], z["t2"], z["l1"], z["l2"])
        assert np.abs(r1.cpu().numpy() - z[tag + "_r1"]).max() < tol
        assert np.abs(r2.cpu().numpy() - z[tag + "_r2"]).max() < tol
