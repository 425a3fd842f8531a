"""
Golden vectors for the CCD and CC2 branches of the reference's builders (pycc/ccwfn.py:413-622),
produced by the reference's own, unmodified code on the same seeded systems as make_golden.py:

    python tests/golden/make_golden_models.py     # in the build container only

-> tests/golden/models/<case>.npz (inputs are taken from tests/golden/<case>.npz).
"""
import contextlib
import io
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from refharness import reference_ccwfn_from_arrays          # noqa: E402
from pycc_b200.synthetic import SyntheticSystem               # noqa: E402
from make_golden import CASES, test_amplitudes                # noqa: E402


def run_case(name, no, nv, seed, offdiag):
    s = SyntheticSystem(no, nv, seed=seed, fock_offdiag=offdiag)
    ERI = s.full_eri()
    out = dict(no=no, nv=nv, seed=seed)
    t1, t2 = test_amplitudes(s, seed, False)
    for model in ("CCD", "CC2", "CC3"):
        tag = model.lower()
        w = reference_ccwfn_from_arrays(s.F, ERI, no, model=model)
        o, v, F, L = w.o, w.v, w.H.F, w.H.L
        x1 = np.zeros_like(t1) if model == "CCD" else t1
        r1, r2 = w.residuals(F, x1, t2)
        out[tag + "_r1"], out[tag + "_r2"] = r1, r2
        out[tag + "_energy"] = w.cc_energy(o, v, F, L, x1, t2)
        out[tag + "_Wmnij"] = w.build_Wmnij(o, v, w.H.ERI, x1, t2)
        if model == "CC3":
            W3 = w.build_cc3_Wmnij(o, v, w.H.ERI, x1)
            out["cc3_Wmnij3"] = W3
            out["cc3_Wmbij"] = w.build_cc3_Wmbij(o, v, w.H.ERI, x1, W3)
            out["cc3_Wmnie"] = w.build_cc3_Wmnie(o, v, w.H.ERI, x1)
            out["cc3_Wamef"] = w.build_cc3_Wamef(o, v, w.H.ERI, x1)
            out["cc3_Wabei"] = w.build_cc3_Wabei(o, v, w.H.ERI, x1)
        if model == "CC2":
            out[tag + "_Zmbij"] = w.build_Zmbij(o, v, w.H.ERI, x1, t2)
        w = reference_ccwfn_from_arrays(s.F, ERI, no, model=model)
        buf = io.StringIO()
        with contextlib.redirect_stdout(buf):
            e = w.solve_cc(e_conv=1e-12, r_conv=1e-12, maxiter=100)
        log = [l for l in buf.getvalue().split("\n") if l.startswith("CC Iter")]
        out[tag + "_solve_ecc"] = e
        out[tag + "_solve_niter"] = len(log) - 1
        out[tag + "_solve_t1"], out[tag + "_solve_t2"] = w.t1, w.t2
        print(name, model, e, len(log) - 1)
    os.makedirs(os.path.join(HERE, "models"), exist_ok=True)
    np.savez_compressed(os.path.join(HERE, "models", name + ".npz"), **out)


if __name__ == "__main__":
    for c in CASES:
        run_case(*c)
