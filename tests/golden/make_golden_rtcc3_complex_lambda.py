"""
Golden vectors for ``cclambda.residuals`` with model CC3 and COMPLEX amplitudes (pycc/cclambda.py:276-447: the call the
real-time propagator makes for the Lambda amplitudes, pycc/rt/rtcc.py:132), evaluated by the reference's own, unmodified
code for a plain and for a ``real_time=True`` wave function:

    python tests/golden/make_golden_rtcc3_complex_lambda.py      # in the build container only

-> tests/golden/rtcc3_complex/<case>_lambda.npz.  Inputs: the complex test amplitudes (t1, t2, l1, l2) of
tests/golden/complex/<case>.npz and the two Fock matrices of tests/golden/rtcc3_complex/<case>.npz (F_field real,
F_mixed complex with a real diagonal).
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
from make_golden import CASES                                               # noqa: E402


def run_case(name, no, nv, seed, offdiag):
    import_reference()
    from pycc.cchbar import cchbar
    from pycc.cclambda import cclambda
    s = SyntheticSystem(no, nv, seed=seed, fock_offdiag=offdiag)
    ERI = s.full_eri()
    z = np.load(os.path.join(HERE, "complex", name + ".npz"))
    zf = np.load(os.path.join(HERE, "rtcc3_complex", name + ".npz"))
    t1, t2, l1, l2 = z["t1"], z["t2"], z["l1"], z["l2"]
    out = dict(t1=t1, t2=t2, l1=l1, l2=l2, F_field=zf["F_field"], F_mixed=zf["F_mixed"])
    for ftag in ("field", "mixed"):
        for tag, rt in (("plain", False), ("rt", True)):
            w = reference_ccwfn_from_arrays(s.F, ERI, no, model="CC3")
            w.real_time = rt
            with contextlib.redirect_stdout(io.StringIO()):
                hb = cchbar(w)
                lam = cclambda(w, hb)
            r1, r2 = lam.residuals(zf["F_" + ftag], t1, t2, l1, l2)
            out["%s_%s_r1" % (ftag, tag)], out["%s_%s_r2" % (ftag, tag)] = r1, r2
        print(name, ftag, np.abs(out[ftag + "_rt_r1"] - out[ftag + "_plain_r1"]).max(),
              np.abs(out[ftag + "_rt_r2"] - out[ftag + "_plain_r2"]).max(), np.abs(out[ftag + "_rt_r2"].imag).max())
    np.savez_compressed(os.path.join(HERE, "rtcc3_complex", name + "_lambda.npz"), **out)


if __name__ == "__main__":
    for c in CASES:
        if c[2] <= 8:
            run_case(*c)
