"""
Golden vectors for complex-amplitude residuals (the real-time propagator's use of the hot path,
pycc/rt/rtcc.py:101-143), produced by the reference's own, unmodified code:

    python tests/golden/make_golden_complex.py     # in the build container only

-> tests/golden/complex/<case>.npz.  Per model (CCSD, CCD, CC2): complex Fock matrix, t1, t2, l1, l2
and the reference's ``ccwfn.residuals(F, t1, t2, real_time=True)`` and
``cclambda.residuals(F, t1, t2, l1, l2)``, on an object whose own amplitudes are the real test
amplitudes of make_golden.py.  A second Fock matrix is real (field-dressed, as rtcc
builds it) to pin the mixed real/complex input path.
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
from make_golden_lambda import test_lambda_amplitudes                       # noqa: E402


def complexify(x, rng, scale):
    return x + 1j * scale * rng.standard_normal(x.shape)


def run_case(name, no, nv, seed, offdiag):
    import_reference()
    from pycc.cchbar import cchbar
    from pycc.cclambda import cclambda
    s = SyntheticSystem(no, nv, seed=seed, fock_offdiag=offdiag)
    ERI = s.full_eri()
    rng = np.random.default_rng(3000 + seed)
    t1r, t2r = test_amplitudes(s, seed, False)
    l1r, l2r = test_lambda_amplitudes(t1r, t2r, seed)
    n = no + nv
    V = rng.standard_normal((n, n))
    Ffield = s.F + 0.01 * (V + V.T)                       # real, field-dressed (rtcc.py:119-121)
    W = rng.standard_normal((n, n))
    Fc = Ffield + 1j * 0.005 * (W + W.T)
    out = dict(no=no, nv=nv, seed=seed, F_field=Ffield, F_complex=Fc)
    out["t1"], out["t2"] = complexify(t1r, rng, 0.03), complexify(t2r, rng, 0.02)
    out["l1"], out["l2"] = complexify(l1r, rng, 0.05), complexify(l2r, rng, 0.05)
    for model in ("CCSD", "CCD", "CC2"):
        tag = model.lower()
        w = reference_ccwfn_from_arrays(s.F, ERI, no, model=model)
        t1 = np.zeros_like(out["t1"]) if model == "CCD" else out["t1"]
        l1 = np.zeros_like(out["l1"]) if model == "CCD" else out["l1"]
        t2, l2 = out["t2"], out["l2"]
        # the object's own amplitudes stay real (the ground state in a propagation); the reference's CC2
        # Lambda code reads them, and H.F, besides the arguments (cclambda.py:481-486, 522-524)
        w.t1, w.t2 = (np.zeros_like(t1r) if model == "CCD" else t1r.copy()), t2r.copy()
        with contextlib.redirect_stdout(io.StringIO()):
            hb = cchbar(w)
        lam = cclambda(w, hb)
        for ftag, F in (("field", Ffield), ("complex", Fc)):
            r1, r2 = w.residuals(F, t1, t2, real_time=True)
            out["%s_%s_r1" % (tag, ftag)], out["%s_%s_r2" % (tag, ftag)] = r1, r2
            q1, q2 = lam.residuals(F, t1, t2, l1, l2)
            out["%s_%s_rl1" % (tag, ftag)], out["%s_%s_rl2" % (tag, ftag)] = q1, q2
        print(name, model, abs(r2).max(), abs(q2).max())
    os.makedirs(os.path.join(HERE, "complex"), exist_ok=True)
    np.savez_compressed(os.path.join(HERE, "complex", name + ".npz"), **out)


if __name__ == "__main__":
    for c in CASES:
        if c[0] != "o5v19_offdiag":      # three cases keep the fixture small
            run_case(*c)
