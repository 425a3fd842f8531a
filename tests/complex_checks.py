"""Shared parity checks for the complex-amplitude residuals
against vectors from the reference's own code (tests/golden/make_golden_complex.py)."""
import contextlib
import io
import os

import numpy as np

from tests.conftest import GOLDEN_DIR, Golden
import pycc_b200

CASES = ["o3v5_offdiag", "o4v8", "o6v13"]
MODELS = ["CCSD", "CCD", "CC2"]


def host(x):
    return x.detach().cpu().numpy()


def check_complex_residuals(make_wfn, name, model, tol=1e-12):
    g = Golden(name)
    z = np.load(os.path.join(GOLDEN_DIR, "complex", name + ".npz"))
    tag = model.lower()
    t1 = np.zeros_like(z["t1"]) if model == "CCD" else z["t1"]
    l1 = np.zeros_like(z["l1"]) if model == "CCD" else z["l1"]
    t2, l2 = z["t2"], z["l2"]
    with contextlib.redirect_stdout(io.StringIO()):
        w = make_wfn(g, model)
        w._load(None, 0 * g["gen_t1"] if model == "CCD" else g["gen_t1"], g["gen_t2"])     # the object's own amplitudes
        hb = pycc_b200.cchbar(w)
        lam = pycc_b200.cclambda(w, hb)
    for ftag, F in (("field", z["F_field"]), ("complex", z["F_complex"])):
        r1, r2 = w.residuals(F, t1, t2, real_time=True)
        assert r1.is_complex() and r2.is_complex()
        assert np.abs(host(r1) - z["%s_%s_r1" % (tag, ftag)]).max() < tol
        assert np.abs(host(r2) - z["%s_%s_r2" % (tag, ftag)]).max() < tol
        q1, q2 = lam.residuals(F, t1, t2, l1, l2)
        assert np.abs(host(q1) - z["%s_%s_rl1" % (tag, ftag)]).max() < tol
        assert np.abs(host(q2) - z["%s_%s_rl2" % (tag, ftag)]).max() < tol
    # the objects' own state is untouched by residuals() calls
    assert np.abs(host(w.t2) - g["gen_t2"]).max() == 0.0 and np.abs(host(w.H.F) - g.F).max() == 0.0
    # real inputs take the real path
    r1, r2 = w.residuals(g.F, g["gen_t1"] if model != "CCD" else 0 * g["gen_t1"], g["gen_t2"])
    assert not r2.is_complex()
    # a complex call with zero imaginary parts reproduces the real residuals
    c1, c2 = w.residuals(g.F.astype(complex), (g["gen_t1"] if model != "CCD" else 0 * g["gen_t1"]).astype(complex),
                         g["gen_t2"].astype(complex))
    assert np.abs(host(c2).imag).max() == 0.0
    assert np.abs(host(c2).real - host(r2)).max() < tol and np.abs(host(c1).real - host(r1)).max() < tol


CC3_CASES = ["o3v5_offdiag", "o4v8"]


def check_complex_cc3_residuals(make_wfn, name, tol=1e-12):
    """Model CC3 with complex amplitudes (the real-time propagator's call) against the reference's own vectors
    (tests/golden/make_golden_rtcc3_complex.py): a real field-dressed and a complex Fock matrix with a real
    diagonal, with and without the t3_pert_ijk term of real_time=True."""
    g = Golden(name)
    z = np.load(os.path.join(GOLDEN_DIR, "rtcc3_complex", name + ".npz"))
    with contextlib.redirect_stdout(io.StringIO()):
        w = make_wfn(g, "CC3")
    for tag in ("field", "mixed"):
        for rt, suf in ((True, ""), (False, "_plain")):
            r1, r2 = w.residuals(z["F_" + tag], z["t1"], z["t2"], real_time=rt)
            assert r1.is_complex() and r2.is_complex()
            assert np.abs(host(r1) - z[tag + "_r1" + suf]).max() < tol, (tag, rt)
            assert np.abs(host(r2) - z[tag + "_r2" + suf]).max() < tol, (tag, rt)
    # the perturbation term is there: the two calls differ
    assert np.abs(z["field_r2"] - z["field_r2_plain"]).max() > 1e-7
    # a complex diagonal of F has no real triples denominator: refused, not approximated
    Fc = z["F_mixed"].copy()
    Fc[0, 0] += 0.25j
    try:
        w.residuals(Fc, z["t1"], z["t2"])
    except ValueError as err:
        assert "diagonal of F must be real" in str(err)
    else:
        raise AssertionError("a complex Fock diagonal was accepted")
    # the object's own state is untouched, and real inputs still take the real CC3 path
    assert np.abs(host(w.H.F) - g.F).max() == 0.0
    q1, q2 = w.residuals(g.F, g["gen_t1"], g["gen_t2"])
    assert not q2.is_complex()
    # complex inputs with zero imaginary parts reproduce the real path
    c1, c2 = w.residuals(g.F.astype(complex), g["gen_t1"].astype(complex), g["gen_t2"].astype(complex))
    assert np.abs(host(c2).imag).max() == 0.0 and np.abs(host(c2).real - host(q2)).max() < tol
    assert np.abs(host(c1).real - host(q1)).max() < tol


def check_complex_cc3_lambda_residuals(make_wfn, name, tol=1e-12):
    """cclambda.residuals for model CC3 with complex t and l (the propagator's second call, pycc/rt/rtcc.py:132)
    against the reference's own vectors (tests/golden/make_golden_rtcc3_complex_lambda.py), for a plain and a
    real_time wave function (the latter subtracts t3_pert_ijk inside the t3 loops), two Fock matrices each."""
    g = Golden(name)
    z = np.load(os.path.join(GOLDEN_DIR, "rtcc3_complex", name + "_lambda.npz"))
    for tag, rt in (("plain", False), ("rt", True)):
        with contextlib.redirect_stdout(io.StringIO()):
            w = make_wfn(g, "CC3")
            w.real_time = rt
            hb = pycc_b200.cchbar(w)
            lam = pycc_b200.cclambda(w, hb)
        for ftag in ("field", "mixed"):
            r1, r2 = lam.residuals(z["F_" + ftag], z["t1"], z["t2"], z["l1"], z["l2"])
            assert r1.is_complex() and r2.is_complex()
            assert np.abs(host(r1) - z["%s_%s_r1" % (ftag, tag)]).max() < tol, (tag, ftag)
            assert np.abs(host(r2) - z["%s_%s_r2" % (ftag, tag)]).max() < tol, (tag, ftag)
    assert np.abs(z["field_rt_r1"] - z["field_plain_r1"]).max() > 1e-7          # the field term is there
