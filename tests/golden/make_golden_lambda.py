"""
Golden vectors for HBAR (pycc/cchbar.py) and the Lambda equations (pycc/cclambda.py), produced by
the reference's own, unmodified code on the same seeded systems as make_golden.py:

    python tests/golden/make_golden_lambda.py     # in the build container only

-> tests/golden/lambda/<case>.npz (inputs are taken from tests/golden/<case>.npz).  Per model
(CCSD, CCD, CC2): every HBAR block at the perturbed test amplitudes (the v^4 block subsampled for
the two larger cases), Goo / Gvv / r_L1 / r_L2 / pseudoenergy at perturbed, unsymmetric Lambda
amplitudes, and a full solve_cc -> cchbar -> solve_lambda chain (pseudoenergy, iteration count,
converged l1 / l2 and the converged t1 / t2 they belong to).
"""
import contextlib
import io
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from refharness import reference_ccwfn_from_arrays, import_reference      # noqa: E402
from pycc_b200.synthetic import SyntheticSystem                             # noqa: E402
from make_golden import CASES, test_amplitudes                              # noqa: E402

BLOCKS = ("Hov", "Hvv", "Hoo", "Hoooo", "Hvvvv", "Hvovv", "Hooov", "Hovvo", "Hovov", "Hvvvo", "Hovoo")
HVVVV_SUB = (slice(None, None, 4), slice(1, None, 3))      # rows kept of Hvvvv when nv > 8


def test_lambda_amplitudes(t1, t2, seed):
    rng = np.random.default_rng(2000 + seed)
    l1 = 2.0 * t1 + 0.05 * rng.standard_normal(t1.shape)
    l2 = 2.0 * (2.0 * t2 - t2.swapaxes(2, 3)) + 0.05 * rng.standard_normal(t2.shape)
    return l1, l2


def run_case(name, no, nv, seed, offdiag):
    import_reference()
    from pycc.cchbar import cchbar
    from pycc.cclambda import cclambda
    s = SyntheticSystem(no, nv, seed=seed, fock_offdiag=offdiag)
    ERI = s.full_eri()
    out = dict(no=no, nv=nv, seed=seed)
    t1, t2 = test_amplitudes(s, seed, False)
    for model in ("CCSD", "CCD", "CC2"):
        tag = model.lower()
        w = reference_ccwfn_from_arrays(s.F, ERI, no, model=model)
        x1 = np.zeros_like(t1) if model == "CCD" else t1
        w.t1, w.t2 = x1.copy(), t2.copy()
        with contextlib.redirect_stdout(io.StringIO()):
            hb = cchbar(w)
        for b in BLOCKS:
            x = getattr(hb, b)
            if b == "Hvvvv" and nv > 8:
                x = x[HVVVV_SUB]
            out["%s_%s" % (tag, b)] = x
        lam = cclambda(w, hb)
        l1, l2 = test_lambda_amplitudes(x1, t2, seed)
        if model == "CCD":
            l1 = np.zeros_like(l1)
        out[tag + "_l1"], out[tag + "_l2"] = l1, l2
        out[tag + "_Goo"], out[tag + "_Gvv"] = lam.build_Goo(t2, l2), lam.build_Gvv(t2, l2)
        r1, r2 = lam.residuals(w.H.F, x1, t2, l1, l2)
        out[tag + "_rl1"], out[tag + "_rl2"] = r1, r2
        out[tag + "_pseudo"] = lam.pseudoenergy(w.o, w.v, w.H.ERI, l2)
        # full chain from the MP2 guess
        w = reference_ccwfn_from_arrays(s.F, ERI, no, model=model)
        buf = io.StringIO()
        with contextlib.redirect_stdout(buf):
            w.solve_cc(e_conv=1e-12, r_conv=1e-12, maxiter=100)
            hb = cchbar(w)
            lam = cclambda(w, hb)
            lecc = lam.solve_lambda(e_conv=1e-12, r_conv=1e-12, maxiter=100)
        log = [l for l in buf.getvalue().split("\n") if l.startswith("LCC Iter")]
        out[tag + "_solve_t1"], out[tag + "_solve_t2"] = w.t1, w.t2
        out[tag + "_solve_lecc"] = lecc
        out[tag + "_solve_niter"] = len(log) - 1
        out[tag + "_solve_l1"], out[tag + "_solve_l2"] = lam.l1, lam.l2
        print(name, model, lecc, len(log) - 1)
    os.makedirs(os.path.join(HERE, "lambda"), exist_ok=True)
    np.savez_compressed(os.path.join(HERE, "lambda", name + ".npz"), **out)


if __name__ == "__main__":
    for c in CASES:
        run_case(*c)
