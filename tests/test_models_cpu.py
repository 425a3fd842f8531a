"""CCD and CC2 branches of the builders: host logic through the
NumPy interpreter back end against vectors produced by the reference's own code
(tests/golden/make_golden_models.py)."""
import contextlib
import io
import os

import numpy as np
import pytest

from tests.conftest import GOLDEN_DIR, Golden
from tests.interp_backend import InterpBackend
import pycc_b200

CASES = ["o3v5_offdiag", "o4v8", "o5v19_offdiag", "o6v13"]


def load(name):
    return Golden(name), np.load(os.path.join(GOLDEN_DIR, "models", name + ".npz"))


@pytest.mark.parametrize("model", ["CCD", "CC2"])
@pytest.mark.parametrize("name", CASES)
def test_model_residuals_and_solve(name, model):
    g, z = load(name)
    tag = model.lower()
    with contextlib.redirect_stdout(io.StringIO()):
        w = pycc_b200.ccwfn.from_arrays(g.F, g.blocks, g.no, model=model, _backend=InterpBackend())
    t1 = np.zeros_like(g["gen_t1"]) if model == "CCD" else g["gen_t1"]
    r1, r2 = w.residuals(g.F, t1, g["gen_t2"])
    assert np.abs(r1.numpy() - z[tag + "_r1"]).max() < 1e-12
    assert np.abs(r2.numpy() - z[tag + "_r2"]).max() < 1e-12
    assert abs(w.cc_energy(w.o, w.v, g.F, w.H.L, t1, g["gen_t2"]) - float(z[tag + "_energy"])) < 1e-12
    W = w.build_Wmnij(w.o, w.v, w.H.ERI, t1, g["gen_t2"])
    assert np.abs(W.numpy() - z[tag + "_Wmnij"]).max() < 1e-12
    if model == "CC2":
        Z = w.build_Zmbij(w.o, w.v, w.H.ERI, t1, g["gen_t2"])
        assert np.abs(Z.numpy() - z[tag + "_Zmbij"]).max() < 1e-12
        assert w.build_Wmbej(w.o, w.v, w.H.ERI, w.H.L, t1, g["gen_t2"]) is None
    else:
        assert w.build_Zmbij(w.o, w.v, w.H.ERI, t1, g["gen_t2"]) is None
        assert w.build_Fme(w.o, w.v, g.F, w.H.L, t1) is None
    with contextlib.redirect_stdout(io.StringIO()):
        w = pycc_b200.ccwfn.from_arrays(g.F, g.blocks, g.no, model=model, _backend=InterpBackend())
        e = w.solve_cc(e_conv=1e-12, r_conv=1e-12, maxiter=100)
    assert len(w.iteration_log) == int(z[tag + "_solve_niter"])
    assert abs(e - float(z[tag + "_solve_ecc"])) < 1e-12
    assert np.abs(w.t1.numpy() - z[tag + "_solve_t1"]).max() < 1e-11
    assert np.abs(w.t2.numpy() - z[tag + "_solve_t2"]).max() < 1e-11


@pytest.mark.parametrize("name", CASES)
def test_cc3_intermediates_and_residuals(name):
    """CC3: T1-dressed intermediates (pycc/ccwfn.py:625-710) and the triples contributions to r1 / r2."""
    g, z = load(name)
    with contextlib.redirect_stdout(io.StringIO()):
        w = pycc_b200.ccwfn.from_arrays(g.F, g.blocks, g.no, model="CC3", _backend=InterpBackend())
    r1, r2 = w.residuals(g.F, g["gen_t1"], g["gen_t2"])
    c = w.engine.cc3
    assert np.abs(c.Wmnij3.numpy() - z["cc3_Wmnij3"]).max() < 1e-12
    assert np.abs(c.Wo.permute(2, 3, 0, 1).numpy() - z["cc3_Wmbij"]).max() < 1e-12          # Wo[ijmb] = Wmbij[mbij]
    assert np.abs(c.Wmnie3.numpy() - z["cc3_Wmnie"]).max() < 1e-12
    assert np.abs(c.Wk.permute(1, 0, 2, 3).numpy() - z["cc3_Wamef"]).max() < 1e-12          # Wk[maef] = Wamef[amef]
    assert np.abs(c.Wv.permute(1, 2, 3, 0).numpy() - z["cc3_Wabei"]).max() < 1e-12          # Wv[iabe] = Wabei[abei]
    assert np.abs(r1.numpy() - z["cc3_r1"]).max() < 1e-12
    assert np.abs(r2.numpy() - z["cc3_r2"]).max() < 1e-12


@pytest.mark.parametrize("name", ["o3v5_offdiag", "o4v8"])
def test_cc3_solve(name):
    g, z = load(name)
    with contextlib.redirect_stdout(io.StringIO()):
        w = pycc_b200.ccwfn.from_arrays(g.F, g.blocks, g.no, model="CC3", _backend=InterpBackend())
        e = w.solve_cc(e_conv=1e-12, r_conv=1e-12, maxiter=100)
    assert len(w.iteration_log) == int(z["cc3_solve_niter"])
    assert abs(e - float(z["cc3_solve_ecc"])) < 1e-12
    assert np.abs(w.t2.numpy() - z["cc3_solve_t2"]).max() < 1e-11


@pytest.mark.parametrize("name", ["o3v5_offdiag", "o4v8"])
def test_cc3_real_time_field(name):
    """CC3 residuals with a field-dressed Fock matrix and real_time=True (t3_pert_ijk, pycc/ccwfn.py:378-386)."""
    import os
    from tests.conftest import GOLDEN_DIR
    g, _ = load(name)
    z = np.load(os.path.join(GOLDEN_DIR, "rtcc3", name + ".npz"))
    with contextlib.redirect_stdout(io.StringIO()):
        w = pycc_b200.ccwfn.from_arrays(g.F, g.blocks, g.no, model="CC3", _backend=InterpBackend())
    r1, r2 = w.residuals(z["F_field"], z["t1"], z["t2"], real_time=True)
    assert np.abs(z["r2"] - z["r2_plain"]).max() > 1e-6
    assert np.abs(r1.numpy() - z["r1"]).max() < 1e-12 and np.abs(r2.numpy() - z["r2"]).max() < 1e-12
    p1, p2 = w.residuals(z["F_field"], z["t1"], z["t2"])
    assert np.abs(p1.numpy() - z["r1_plain"]).max() < 1e-12 and np.abs(p2.numpy() - z["r2_plain"]).max() < 1e-12
    # the object's own Fock matrix and amplitudes are untouched
    assert np.abs(w.H.F.numpy() - g.F).max() == 0.0
