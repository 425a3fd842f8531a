"""
Golden vectors for the CC3 residuals with COMPLEX amplitudes - the call the real-time propagator makes for model CC3
(pycc/rt/rtcc.py:125 -> ccwfn.residuals(F, t1, t2, real_time=True), pycc/ccwfn.py:371-401 with
cctriples.t3c_ijk / t3_pert_ijk) - produced by the reference's own, unmodified code:

    python tests/golden/make_golden_rtcc3_complex.py      # in the build container only

-> tests/golden/rtcc3_complex/<case>.npz.  Inputs: the complex test amplitudes (t1, t2) and the real field-dressed
F_field of tests/golden/complex/<case>.npz, plus F_mixed = the complex Fock matrix of that file with the imaginary
part of its DIAGONAL removed (the engine divides by real triples denominators; a real field leaves the diagonal
real).  Outputs per Fock matrix: r1 / r2 with real_time=True and with real_time=False.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from refharness import reference_ccwfn_from_arrays, import_reference      # noqa: E402
from pycc_b200.synthetic import SyntheticSystem                             # noqa: E402
from make_golden import CASES                                               # noqa: E402


def run_case(name, no, nv, seed, offdiag):
    import_reference()
    s = SyntheticSystem(no, nv, seed=seed, fock_offdiag=offdiag)
    ERI = s.full_eri()
    z = np.load(os.path.join(HERE, "complex", name + ".npz"))
    t1, t2 = z["t1"], z["t2"]
    assert np.iscomplexobj(t1) and np.iscomplexobj(t2)
    F_field = z["F_field"]
    F_mixed = z["F_complex"].copy()
    F_mixed[np.diag_indices_from(F_mixed)] = F_mixed.diagonal().real
    out = dict(t1=t1, t2=t2, F_field=F_field, F_mixed=F_mixed)
    w = reference_ccwfn_from_arrays(s.F, ERI, no, model="CC3")
    for tag, F in (("field", F_field), ("mixed", F_mixed)):
        r1, r2 = w.residuals(F, t1, t2, real_time=True)
        p1, p2 = w.residuals(F, t1, t2, real_time=False)
        out[tag + "_r1"], out[tag + "_r2"], out[tag + "_r1_plain"], out[tag + "_r2_plain"] = r1, r2, p1, p2
        print(name, tag, np.abs(r1).max(), np.abs(r2).max(), np.abs(r2.imag).max(), np.abs(r2 - p2).max())
    os.makedirs(os.path.join(HERE, "rtcc3_complex"), exist_ok=True)
    np.savez_compressed(os.path.join(HERE, "rtcc3_complex", name + ".npz"), **out)


if __name__ == "__main__":
    for c in CASES:
        if c[2] <= 8:
            run_case(*c)
