"""
Golden vectors for the (T) contributions to the Lambda equations and densities (pycc/ccwfn.py:721-855,
``t3_density``) and for the CCSD(T) Lambda equations (pycc/cclambda.py:474-475, 519-520), produced by the
reference's own, unmodified code:

    python tests/golden/make_golden_t3density.py      # in the build container only

-> tests/golden/t3density/<case>.npz (inputs are taken from tests/golden/<case>.npz).  Per case: ET, S1, S2, Dvv,
Doo, Dov, Goovv, Gooov, Gvvvo at the perturbed test amplitudes (general, and `sym_`: pair-symmetric); r_L1 / r_L2 of model CCSD(T) at perturbed Lambda
amplitudes; and the chain solve_cc(make_t3_density) -> cchbar -> solve_lambda (energies, iteration count, l1 / l2).
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

FIELDS = ("S1", "S2", "Dvv", "Doo", "Dov", "Goovv", "Gooov", "Gvvvo")


def run_case(name, no, nv, seed, offdiag):
    import_reference()
    from pycc.cchbar import cchbar
    from pycc.cclambda import cclambda
    s = SyntheticSystem(no, nv, seed=seed, fock_offdiag=offdiag)
    ERI = s.full_eri()
    out = dict(no=no, nv=nv, seed=seed)
    # pair-symmetric amplitudes (what every iterate of solve_cc is) and general ones
    for tag, sym in (("sym_", True), ("", False)):
        t1, t2 = test_amplitudes(s, seed, sym)
        w = reference_ccwfn_from_arrays(s.F, ERI, no, model="CCSD(T)")
        w.t1, w.t2 = t1.copy(), t2.copy()
        out[tag + "ET"] = w.t3_density()
        for f in FIELDS:
            out[tag + f] = getattr(w, f)
    with contextlib.redirect_stdout(io.StringIO()):
        hb = cchbar(w)
    lam = cclambda(w, hb)
    l1, l2 = test_lambda_amplitudes(t1, t2, seed)
    out["l1"], out["l2"] = l1, l2
    o, v = w.o, w.v
    out["rl1"] = lam.r_L1(o, v, l1, l2, hb.Hov, hb.Hvv, hb.Hoo, hb.Hovvo, hb.Hovov, hb.Hvvvo, hb.Hovoo, hb.Hvovv, hb.Hooov,
                          lam.build_Gvv(t2, l2), lam.build_Goo(t2, l2))
    out["rl2"] = lam.r_L2(o, v, l1, l2, w.H.L, hb.Hov, hb.Hvv, hb.Hoo, hb.Hoooo, hb.Hvvvv, hb.Hovvo, hb.Hovov, hb.Hvvvo, hb.Hovoo,
                          hb.Hvovv, hb.Hooov, lam.build_Gvv(t2, l2), lam.build_Goo(t2, l2))
    # full chain from the MP2 guess
    w = reference_ccwfn_from_arrays(s.F, ERI, no, model="CCSD(T)")
    w.make_t3_density = True
    buf = io.StringIO()
    with contextlib.redirect_stdout(buf):
        ecc = w.solve_cc(e_conv=1e-12, r_conv=1e-12, maxiter=100)
        hb = cchbar(w)
        lam = cclambda(w, hb)
        lecc = lam.solve_lambda(e_conv=1e-12, r_conv=1e-12, maxiter=100)
    log = [l for l in buf.getvalue().split("\n") if l.startswith("LCC Iter")]
    out["solve_ecc_total"] = ecc
    out["solve_t1"], out["solve_t2"] = w.t1, w.t2
    out["solve_S1"], out["solve_S2"] = w.S1, w.S2
    out["solve_lecc"] = lecc
    out["solve_niter"] = len(log) - 1
    out["solve_l1"], out["solve_l2"] = lam.l1, lam.l2
    print(name, out["ET"], ecc, lecc, len(log) - 1)
    os.makedirs(os.path.join(HERE, "t3density"), exist_ok=True)
    np.savez_compressed(os.path.join(HERE, "t3density", name + ".npz"), **out)


if __name__ == "__main__":
    for c in CASES:
        if c[2] <= 13:               # the reference's second loop is v^3 interpreted iterations of 18 einsums
            run_case(*c)
