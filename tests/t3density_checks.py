"""Shared checks of ccwfn.t3_density and the CCSD(T) Lambda equations (pycc/ccwfn.py:721-855,
pycc/cclambda.py:474-475, 519-520) against tests/golden/t3density/*.npz (unmodified reference,
tests/golden/make_golden_t3density.py), for the CPU (interpreter back end) and GPU (C ABI) tests."""
import contextlib
import io
import os

import numpy as np

FIELDS = ("S1", "S2", "Dvv", "Doo", "Dov", "Goovv", "Gooov", "Gvvvo")
CASES = ("o3v5_offdiag", "o4v8", "o6v13")


def load(name):
    from tests.conftest import GOLDEN_DIR, Golden
    return Golden(name), np.load(os.path.join(GOLDEN_DIR, "t3density", name + ".npz"))


def _wfn(g, backend, **kw):
    import pycc_b200
    with contextlib.redirect_stdout(io.StringIO()):
        return pycc_b200.ccwfn.from_arrays(g.F, g.blocks, g.no, model="CCSD(T)", _backend=backend, **kw)


def check_t3_density(name, backend, tol=1e-12):
    g, z = load(name)
    offdiag = np.abs(g.F - np.diag(np.diag(g.F))).max() > 0
    for tag, amp in (("sym_", "sym"), ("", "gen")):
        w = _wfn(g, backend)
        w._load(None, g[amp + "_t1"], g[amp + "_t2"])
        et = w.t3_density()
        assert abs(et - float(z[tag + "ET"])) < tol
        for f in FIELDS:
            if f == "Doo" and tag == "" and offdiag:
                continue        # unsymmetric t2 AND non-diagonal F: outside the symmetry Doo relies on (engine docstring)
            ref = z[tag + f]
            assert np.abs(ref).max() > 1e-5, f
            assert np.abs(getattr(w, f).cpu().numpy() - ref).max() < tol, (name, tag, f)
        # E(T) of t3_density is the "Vikings" form: the reference's own t_vikings value on the same amplitudes ...
        if amp + "_et_vikings" in g.z.files:
            assert abs(et - float(g[amp + "_et_vikings"])) < tol
        # ... which equals Lee-Rendell's t_tjl for pair-symmetric amplitudes (reference test_005:38-40)
        if amp == "sym":
            import pycc_b200
            assert abs(pycc_b200.t_tjl(w) - et) < 1e-11


def check_lambda_t(name, backend):
    import pycc_b200
    g, z = load(name)
    w = _wfn(g, backend)
    w._load(None, g["gen_t1"], g["gen_t2"])
    with contextlib.redirect_stdout(io.StringIO()):
        w.t3_density()
        hb = pycc_b200.cchbar(w)
        lam = pycc_b200.cclambda(w, hb)
    o, v, l1, l2 = w.o, w.v, z["l1"], z["l2"]
    t2 = g["gen_t2"]
    Gvv, Goo = lam.build_Gvv(t2, l2), lam.build_Goo(t2, l2)
    r1 = lam.r_L1(o, v, l1, l2, hb.Hov, hb.Hvv, hb.Hoo, hb.Hovvo, hb.Hovov, hb.Hvvvo, hb.Hovoo, hb.Hvovv, hb.Hooov, Gvv, Goo)
    r2 = lam.r_L2(o, v, l1, l2, w.H.L, hb.Hov, hb.Hvv, hb.Hoo, hb.Hoooo, hb.Hvvvv, hb.Hovvo, hb.Hovov, hb.Hvvvo, hb.Hovoo,
                  hb.Hvovv, hb.Hooov, Gvv, Goo)
    assert np.abs(r1.cpu().numpy() - z["rl1"]).max() < 1e-12
    assert np.abs(r2.cpu().numpy() - z["rl2"]).max() < 1e-12
    # the chain solve_cc(make_t3_density) -> cchbar -> solve_lambda from the MP2 guess
    w = _wfn(g, backend, make_t3_density=True)
    with contextlib.redirect_stdout(io.StringIO()):
        ecc = w.solve_cc(e_conv=1e-12, r_conv=1e-12, maxiter=100)
        hb = pycc_b200.cchbar(w)
        lam = pycc_b200.cclambda(w, hb)
        lecc = lam.solve_lambda(e_conv=1e-12, r_conv=1e-12, maxiter=100)
    assert abs(ecc - float(z["solve_ecc_total"])) < 1e-10
    assert np.abs(w.S2.cpu().numpy() - z["solve_S2"]).max() < 1e-10
    assert len(lam.iteration_log) == int(z["solve_niter"])
    assert abs(lecc - float(z["solve_lecc"])) < 1e-10
    assert np.abs(lam.l1.cpu().numpy() - z["solve_l1"]).max() < 1e-9
    assert np.abs(lam.l2.cpu().numpy() - z["solve_l2"]).max() < 1e-9
