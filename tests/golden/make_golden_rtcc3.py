"""
Golden vectors for the CC3 residuals with a field-dressed (real) Fock matrix and ``real_time=True``: the reference's
own, unmodified ``ccwfn.residuals`` (pycc/ccwfn.py:371-401) adds the ``t3_pert_ijk`` term (cctriples.py:547-562)
built from V = F - H.F.

    python tests/golden/make_golden_rtcc3.py      # in the build container only

-> tests/golden/rtcc3/<case>.npz : r1, r2 at the perturbed test amplitudes of tests/golden/<case>.npz with the
F_field of tests/golden/complex/<case>.npz; r1_plain / r2_plain are the same call with real_time=False.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from refharness import reference_ccwfn_from_arrays, import_reference      # noqa: E402
from pycc_b200.synthetic import SyntheticSystem                             # noqa: E402
from make_golden import CASES, test_amplitudes                              # noqa: E402


def run_case(name, no, nv, seed, offdiag):
    import_reference()
    s = SyntheticSystem(no, nv, seed=seed, fock_offdiag=offdiag)
    ERI = s.full_eri()
    t1, t2 = test_amplitudes(s, seed, False)
    F_field = np.load(os.path.join(HERE, "complex", name + ".npz"))["F_field"]
    assert not np.iscomplexobj(F_field)
    w = reference_ccwfn_from_arrays(s.F, ERI, no, model="CC3")
    r1, r2 = w.residuals(F_field, t1, t2, real_time=True)
    p1, p2 = w.residuals(F_field, t1, t2, real_time=False)
    print(name, np.abs(r1 - p1).max(), np.abs(r2 - p2).max())
    os.makedirs(os.path.join(HERE, "rtcc3"), exist_ok=True)
    np.savez_compressed(os.path.join(HERE, "rtcc3", name + ".npz"), F_field=F_field, t1=t1, t2=t2, r1=r1, r2=r2,
                        r1_plain=p1, r2_plain=p2)


if __name__ == "__main__":
    for c in CASES:
        if c[2] <= 13:
            run_case(*c)
