"""Shared parity checks for HBAR / Lambda against the vectors the
reference's own code produced (tests/golden/make_golden_lambda.py).  ``make_wfn(g, model)`` builds
the ccwfn under test: on the NumPy interpreter back end for the host-logic tests, on the CUDA
library for the ``gpu`` tests."""
import contextlib
import io
import os

import numpy as np

from tests.conftest import GOLDEN_DIR, Golden
import pycc_b200

BLOCKS = ("Hov", "Hvv", "Hoo", "Hoooo", "Hvvvv", "Hvovv", "Hooov", "Hovvo", "Hovov", "Hvvvo", "Hovoo")
HVVVV_SUB = (slice(None, None, 4), slice(1, None, 3))
CASES = ["o3v5_offdiag", "o4v8", "o5v19_offdiag", "o6v13"]
MODELS = ["CCSD", "CCD", "CC2"]


def load(name):
    return Golden(name), np.load(os.path.join(GOLDEN_DIR, "lambda", name + ".npz"))


def host(x):
    return x.detach().cpu().numpy()


def check_blocks_and_residuals(make_wfn, name, model, tol=1e-12):
    g, z = load(name)
    tag = model.lower()
    t1 = np.zeros_like(g["gen_t1"]) if model == "CCD" else g["gen_t1"]
    t2 = g["gen_t2"]
    with contextlib.redirect_stdout(io.StringIO()):
        w = make_wfn(g, model)
        w._load(None, t1, t2)
        hb = pycc_b200.cchbar(w)
    for b in BLOCKS:
        x = host(getattr(hb, b))
        if b == "Hvvvv" and g.nv > 8:
            x = x[HVVVV_SUB]
        assert x.shape == z["%s_%s" % (tag, b)].shape, b
        assert np.abs(x - z["%s_%s" % (tag, b)]).max() < tol, b
    lam = pycc_b200.cclambda(w, hb)
    # the guess (cclambda.py:57-58)
    assert np.abs(host(lam.l1) - 2.0 * t1).max() < 1e-14
    assert np.abs(host(lam.l2) - 2.0 * (2.0 * t2 - t2.swapaxes(2, 3))).max() < 1e-14
    l1, l2 = z[tag + "_l1"], z[tag + "_l2"]
    assert np.abs(host(lam.build_Goo(t2, l2)) - z[tag + "_Goo"]).max() < tol
    assert np.abs(host(lam.build_Gvv(t2, l2)) - z[tag + "_Gvv"]).max() < tol
    r1, r2 = lam.residuals(g.F, t1, t2, l1, l2)
    assert np.abs(host(r1) - z[tag + "_rl1"]).max() < tol
    assert np.abs(host(r2) - z[tag + "_rl2"]).max() < tol
    assert abs(lam.pseudoenergy(w.o, w.v, None, l2) - float(z[tag + "_pseudo"])) < tol
    # single builders with the reference's signatures, after the amplitudes were changed
    w._load(None, 0.5 * t1, 0.5 * t2)
    x = hb.build_Hovoo(w.o, w.v, None, None, None, None, t1, t2)
    assert np.abs(host(x) - z[tag + "_Hovoo"]).max() < tol
    x = hb.build_Hvv(w.o, w.v, g.F, None, t1, t2)
    assert np.abs(host(x) - z[tag + "_Hvv"]).max() < tol
    return w


def check_solve(make_wfn, name, model, chain, e_tol=1e-11, a_tol=1e-9):
    """``chain``: solve_cc first (the reference's workflow); otherwise the reference's converged
    t1 / t2 are loaded so that the Lambda solve is checked on its own."""
    g, z = load(name)
    tag = model.lower()
    with contextlib.redirect_stdout(io.StringIO()):
        w = make_wfn(g, model)
        if chain:
            w.solve_cc(e_conv=1e-12, r_conv=1e-12, maxiter=100)
        else:
            w._load(None, z[tag + "_solve_t1"], z[tag + "_solve_t2"])
        hb = pycc_b200.cchbar(w)
        lam = pycc_b200.cclambda(w, hb)
        buf = io.StringIO()
        with contextlib.redirect_stdout(buf):
            lecc = lam.solve_lambda(e_conv=1e-12, r_conv=1e-12, maxiter=100)
    lines = [l for l in buf.getvalue().split("\n") if l.startswith("LCC Iter")]
    assert lines[0].startswith("LCC Iter   0: LCC PseudoE = ")
    assert len(lines) - 1 == int(z[tag + "_solve_niter"]) == len(lam.iteration_log)
    assert abs(lecc - float(z[tag + "_solve_lecc"])) < e_tol
    assert np.abs(host(lam.l1) - z[tag + "_solve_l1"]).max() < a_tol
    assert np.abs(host(lam.l2) - z[tag + "_solve_l2"]).max() < a_tol
