"""
Golden vectors for ``cclambda.residuals`` with model CC3 (pycc/cclambda.py:276-447): the real-time entry, evaluated by the
reference's own, unmodified code at perturbed real amplitudes with a field-dressed (real) Fock matrix, for a plain and
for a ``real_time=True`` wave function (the latter adds t3_pert_ijk inside the t3 loops).

    python tests/golden/make_golden_rtcc3_lambda.py      # in the build container only

-> tests/golden/rtcc3/<case>_lambda.npz
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
    t1, t2 = test_amplitudes(s, seed, False)
    l1, l2 = test_lambda_amplitudes(t1, t2, seed)
    F_field = np.load(os.path.join(HERE, "complex", name + ".npz"))["F_field"]
    out = dict(F_field=F_field, t1=t1, t2=t2, l1=l1, l2=l2)
    for tag, rt in (("plain", False), ("rt", True)):
        w = reference_ccwfn_from_arrays(s.F, ERI, no, model="CC3")
        w.real_time = rt
        with contextlib.redirect_stdout(io.StringIO()):
            hb = cchbar(w)
            lam = cclambda(w, hb)
        r1, r2 = lam.residuals(F_field, t1, t2, l1, l2)
        out[tag + "_r1"], out[tag + "_r2"] = r1, r2
    print(name, np.abs(out["rt_r1"] - out["plain_r1"]).max(), np.abs(out["rt_r2"] - out["plain_r2"]).max())
    np.savez_compressed(os.path.join(HERE, "rtcc3", name + "_lambda.npz"), **out)


if __name__ == "__main__":
    for c in CASES:
        if c[2] <= 8:
            run_case(*c)
