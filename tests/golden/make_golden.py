"""
Generate tests/golden/*.npz by running the reference's own, unmodified code
(/root/reference/pycc/{ccwfn,cctriples,utils}.py) on seeded synthetic inputs.

    python tests/golden/make_golden.py          # in the build container only

Each file stores the inputs (F, the six ERI blocks, t1/t2 test amplitudes) and the
reference's outputs, so the consumers never need /root/reference.
"""
import contextlib
import io
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from refharness import import_reference, reference_ccwfn_from_arrays   # noqa: E402
from pycc_b200.synthetic import SyntheticSystem                          # noqa: E402

CASES = [
    # name, no, nv, seed, fock_offdiag
    ("o3v5_offdiag", 3, 5, 11, True),
    ("o4v8", 4, 8, 1, False),
    ("o5v19_offdiag", 5, 19, 7, True),
    ("o6v13", 6, 13, 3, False),
]


def test_amplitudes(sys_, seed, symmetric=True):
    """MP2 amplitudes plus a seeded perturbation (so that every term of residuals() is exercised)."""
    no, nv = sys_.no, sys_.nv
    rng = np.random.default_rng(1000 + seed)
    eps = np.diag(sys_.F)
    eo, ev = eps[:no], eps[no:]
    D2 = eo[:, None, None, None] + eo[None, :, None, None] - ev[None, None, :, None] - ev[None, None, None, :]
    t2 = sys_.blocks["oovv"] / D2 + 0.02 * rng.standard_normal((no, no, nv, nv))
    if symmetric:
        t2 = 0.5 * (t2 + t2.transpose(1, 0, 3, 2))
    t1 = 0.05 * rng.standard_normal((no, nv))
    return t1, t2


def run_case(name, no, nv, seed, offdiag):
    _, tri, _ = import_reference()
    s = SyntheticSystem(no, nv, seed=seed, fock_offdiag=offdiag)
    ERI = s.full_eri()
    out = dict(no=no, nv=nv, seed=seed, scale=s.scale, F=s.F)
    for k, b in s.blocks.items():
        out["blk_" + k] = b
    for tag, symm in (("sym", True), ("gen", False)):
        w = reference_ccwfn_from_arrays(s.F, ERI, no, model="CCSD")
        t1, t2 = test_amplitudes(s, seed, symm)
        o, v, F, L = w.o, w.v, w.H.F, w.H.L
        out[tag + "_t1"], out[tag + "_t2"] = t1, t2
        out[tag + "_Fae"] = w.build_Fae(o, v, F, L, t1, t2)
        out[tag + "_Fmi"] = w.build_Fmi(o, v, F, L, t1, t2)
        out[tag + "_Fme"] = w.build_Fme(o, v, F, L, t1)
        out[tag + "_Wmnij"] = w.build_Wmnij(o, v, w.H.ERI, t1, t2)
        out[tag + "_Wmbej"] = w.build_Wmbej(o, v, w.H.ERI, L, t1, t2)
        out[tag + "_Wmbje"] = w.build_Wmbje(o, v, w.H.ERI, t1, t2)
        out[tag + "_Zmbij"] = w.build_Zmbij(o, v, w.H.ERI, t1, t2)
        r1, r2 = w.residuals(F, t1, t2)
        out[tag + "_r1"], out[tag + "_r2"] = r1, r2
        out[tag + "_tau_1_h"] = w.build_tau(t1, t2, 1.0, 0.5)
        out[tag + "_energy"] = w.cc_energy(o, v, F, L, t1, t2)
        # (T) on the (unconverged) test amplitudes: exercises t1- and Fov-dependent terms
        w.t1, w.t2 = t1.copy(), t2.copy()
        out[tag + "_et_tjl"] = tri.t_tjl(w)
        out[tag + "_et_vikings"] = tri.t_vikings(w)
        if nv <= 13:
            out[tag + "_et_vikings_inverted"] = tri.t_vikings_inverted(w)

    # full solve, DP, CCSD(T) (solve_cc calls t_tjl at convergence: ccwfn.py:313-320)
    w = reference_ccwfn_from_arrays(s.F, ERI, no, model="CCSD(T)")
    buf = io.StringIO()
    with contextlib.redirect_stdout(buf):
        e = w.solve_cc(e_conv=1e-12, r_conv=1e-12, maxiter=100)
    log = [l for l in buf.getvalue().split("\n") if l.startswith("CC Iter")]
    iters = []
    for l in log[1:]:
        p = l.replace("=", " ").split()
        iters.append((float(p[5]), float(p[7]), float(p[9])))
    out["solve_eccsd_t_total"] = e
    out["solve_iter_log"] = np.array(iters)        # (ecc, dE, rms) per iteration as printed
    out["solve_t1"], out["solve_t2"] = w.t1, w.t2
    w.model = "CCSD"
    out["solve_et"] = tri.t_tjl(w)
    out["solve_eccsd"] = e - out["solve_et"]
    out["solve_et_vikings"] = tri.t_vikings(w)

    # no-DIIS and late-DIIS variants pin the DIIS bookkeeping (utils.py:48-58, ccwfn.py:327-329)
    for tag, kw in (("nodiis", dict(max_diis=0)), ("diis3s2", dict(max_diis=3, start_diis=2))):
        w = reference_ccwfn_from_arrays(s.F, ERI, no, model="CCSD")
        buf = io.StringIO()
        with contextlib.redirect_stdout(buf):
            e = w.solve_cc(e_conv=1e-10, r_conv=1e-10, maxiter=200, **kw)
        log = [l for l in buf.getvalue().split("\n") if l.startswith("CC Iter")]
        out["solve_%s_ecc" % tag] = np.nan if e is None else e
        out["solve_%s_niter" % tag] = len(log) - 1

    # single precision mode (ccwfn.py:189-196): everything float32
    w = reference_ccwfn_from_arrays(s.F, ERI, no, model="CCSD", precision="SP")
    buf = io.StringIO()
    with contextlib.redirect_stdout(buf):
        e = w.solve_cc(e_conv=1e-6, r_conv=1e-6, maxiter=100)
    out["solve_sp_ecc"] = np.nan if e is None else float(e)
    np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)
    print(name, "E_CCSD", out["solve_eccsd"], "E_T", out["solve_et"], "niter", len(iters),
          "nodiis", out["solve_nodiis_niter"], "sp", out["solve_sp_ecc"])


if __name__ == "__main__":
    for c in CASES:
        run_case(*c)
