"""
Golden vectors for the CC3 Lambda equations (pycc/cclambda.py:118-139, 160-235 with cctriples.l3_ijk, :219-277),
produced by the reference's own, unmodified code:

    python tests/golden/make_golden_lambda_cc3.py      # in the build container only

-> tests/golden/lambda_cc3/<case>.npz (inputs are taken from tests/golden/<case>.npz).  Per case:
  * ONE Lambda iteration at perturbed (unconverged, unsymmetric) amplitudes: the reference's solve_lambda has no
    separate CC3 residual function, so it is run with maxiter=1 and DIIS off from (t1, t2, l1, l2); the residuals
    follow from the update, r = (l_new - l_old) * D  (stored as r1 / r2, symmetrised r2 as the loop adds it);
  * the chain solve_cc -> cchbar -> solve_lambda from the MP2 guess (energies, iteration count, l1 / l2).
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


def run_case(name, no, nv, seed, offdiag):
    import_reference()
    from pycc.cchbar import cchbar
    from pycc.cclambda import cclambda
    s = SyntheticSystem(no, nv, seed=seed, fock_offdiag=offdiag)
    ERI = s.full_eri()
    out = dict(no=no, nv=nv, seed=seed)
    # one iteration at perturbed amplitudes
    t1, t2 = test_amplitudes(s, seed, False)
    l1, l2 = test_lambda_amplitudes(t1, t2, seed)
    w = reference_ccwfn_from_arrays(s.F, ERI, no, model="CC3")
    w.t1, w.t2 = t1.copy(), t2.copy()
    with contextlib.redirect_stdout(io.StringIO()):
        hb = cchbar(w)
        lam = cclambda(w, hb)
        lam.l1, lam.l2 = l1.copy(), l2.copy()
        lam.solve_lambda(e_conv=1e-30, r_conv=1e-30, maxiter=1, max_diis=8, start_diis=100)
    out["t1"], out["t2"], out["l1"], out["l2"] = t1, t2, l1, l2
    out["r1"] = (lam.l1 - l1) * w.Dia
    out["r2"] = (lam.l2 - l2) * w.Dijab
    # full chain from the MP2 guess
    w = reference_ccwfn_from_arrays(s.F, ERI, no, model="CC3")
    buf = io.StringIO()
    with contextlib.redirect_stdout(buf):
        ecc = w.solve_cc(e_conv=1e-12, r_conv=1e-12, maxiter=100)
        hb = cchbar(w)
        lam = cclambda(w, hb)
        lecc = lam.solve_lambda(e_conv=1e-12, r_conv=1e-12, maxiter=100)
    log = [l for l in buf.getvalue().split("\n") if l.startswith("LCC Iter")]
    out["solve_ecc"] = ecc
    out["solve_t1"], out["solve_t2"] = w.t1, w.t2
    out["solve_lecc"] = lecc
    out["solve_niter"] = len(log) - 1
    out["solve_l1"], out["solve_l2"] = lam.l1, lam.l2
    print(name, np.abs(out["r1"]).max(), np.abs(out["r2"]).max(), ecc, lecc, len(log) - 1)
    os.makedirs(os.path.join(HERE, "lambda_cc3"), exist_ok=True)
    np.savez_compressed(os.path.join(HERE, "lambda_cc3", name + ".npz"), **out)


if __name__ == "__main__":
    for c in CASES:
        if c[2] <= 13:               # o^3 interpreted iterations of ~70 einsums, every Lambda iteration
            run_case(*c)
