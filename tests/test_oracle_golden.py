"""The CPU oracle (oracle/) against the reference-generated golden vectors: this is what pins it."""
import numpy as np
import pytest

from oracle import ccsd_oracle as oc
from oracle import triples_oracle as ot

TOL = 1e-12


@pytest.mark.parametrize("tag", ["sym", "gen"])
def test_builders_and_residuals(golden, tag):
    g = golden
    blk = oc.Blocks(g.blocks)
    t1, t2 = g[tag + "_t1"], g[tag + "_t2"]
    r1, r2, im = oc.residuals(g.F, blk, t1, t2, want_intermediates=True)
    for k in ("Fae", "Fmi", "Fme", "Wmnij", "Wmbej", "Wmbje", "Zmbij"):
        assert np.abs(im[k] - g[tag + "_" + k]).max() < TOL, k
    assert np.abs(r1 - g[tag + "_r1"]).max() < TOL
    assert np.abs(r2 - g[tag + "_r2"]).max() < TOL
    assert np.abs(oc.tau(t1, t2, 1.0, 0.5) - g[tag + "_tau_1_h"]).max() < 1e-15
    assert abs(oc.cc_energy(g.F, blk, t1, t2) - float(g[tag + "_energy"])) < TOL


@pytest.mark.parametrize("tag", ["sym", "gen"])
def test_triples_on_test_amplitudes(golden, tag):
    g = golden
    t1, t2 = g[tag + "_t1"], g[tag + "_t2"]
    ref = float(g[tag + "_et_tjl"])
    assert abs(ot.t_tjl(g.F, g.blocks, t1, t2) - ref) < TOL
    if g.nv <= 13:
        assert abs(ot.t_vikings(g.F, g.blocks, t1, t2) - float(g[tag + "_et_vikings"])) < TOL


def test_solve_cc_matches_reference_iterations(golden):
    g = golden
    log = []
    ecc, t1, t2, niter = oc.solve_cc(g.F, g.blocks, 1e-12, 1e-12, maxiter=100, log=log)
    ref_log = g["solve_iter_log"]
    assert niter == ref_log.shape[0]
    assert abs(ecc - float(g["solve_eccsd"])) < 1e-12
    assert np.abs(t1 - g["solve_t1"]).max() < 1e-11
    assert np.abs(t2 - g["solve_t2"]).max() < 1e-11
    for (it, e, de, rms), row in zip(log[1:], ref_log):
        assert abs(e - row[0]) < 1e-12
    et = ot.t_tjl(g.F, g.blocks, t1, t2)
    assert abs(et - float(g["solve_et"])) < 1e-12
    # t_tjl == t_vikings on symmetric ERIs (reference test_005:38-40)
    assert abs(float(g["solve_et"]) - float(g["solve_et_vikings"])) < 1e-12


@pytest.mark.parametrize("tag,kw", [("nodiis", dict(max_diis=0)), ("diis3s2", dict(max_diis=3, start_diis=2))])
def test_diis_variants(golden, tag, kw):
    g = golden
    ecc, t1, t2, niter = oc.solve_cc(g.F, g.blocks, 1e-10, 1e-10, maxiter=200, **kw)
    assert niter == int(g["solve_%s_niter" % tag])
    assert abs(ecc - float(g["solve_%s_ecc" % tag])) < 1e-12


def test_lambda_oracle_against_reference_vectors(golden):
    """oracle/lambda_oracle.py (HBAR + Lambda, CCSD branch) against tests/golden/lambda/*.npz."""
    import os
    from oracle import lambda_oracle as ol
    from tests.conftest import GOLDEN_DIR
    from tests.lambda_checks import BLOCKS, HVVVV_SUB
    g = golden
    z = np.load(os.path.join(GOLDEN_DIR, "lambda", g.name + ".npz"))
    blk = oc.Blocks(g.blocks)
    t1, t2 = g["gen_t1"], g["gen_t2"]
    H = ol.build_hbar(g.F, blk, t1, t2)
    for b in BLOCKS:
        x = H[b][HVVVV_SUB] if (b == "Hvvvv" and g.nv > 8) else H[b]
        assert np.abs(x - z["ccsd_" + b]).max() < 1e-12, b
    l1, l2 = z["ccsd_l1"], z["ccsd_l2"]
    assert np.abs(ol.build_Goo(t2, l2) - z["ccsd_Goo"]).max() < 1e-12
    assert np.abs(ol.build_Gvv(t2, l2) - z["ccsd_Gvv"]).max() < 1e-12
    r1, r2 = ol.lambda_residuals(blk, H, t2, l1, l2)
    assert np.abs(r1 - z["ccsd_rl1"]).max() < 1e-12
    assert np.abs(r2 - z["ccsd_rl2"]).max() < 1e-12
    assert abs(ol.pseudoenergy(blk, l2) - float(z["ccsd_pseudo"])) < 1e-13
    # the converged Lambda amplitudes are a fixed point of the iteration
    ts1, ts2 = z["ccsd_solve_t1"], z["ccsd_solve_t2"]
    H = ol.build_hbar(g.F, blk, ts1, ts2)
    n1, n2, lecc, rms = ol.lambda_iteration(g.F, blk, H, ts2, z["ccsd_solve_l1"], z["ccsd_solve_l2"])
    assert rms < 1e-10 and abs(lecc - float(z["ccsd_solve_lecc"])) < 1e-11


# ---- the torch restatement (oracle/torch_oracle.py): same formulas on torch tensors, used by bench.py as
# the full-size checker on the GPU box (cuBLAS FP64); pinned here, on the CPU, to the same golden vectors
@pytest.mark.parametrize("tag", ["sym", "gen"])
def test_torch_oracle_residuals_and_energy(golden, tag):
    from oracle import torch_oracle as to
    g = golden
    blk = to.Blocks(g.blocks)
    t1, t2 = g[tag + "_t1"], g[tag + "_t2"]
    r1, r2 = to.residuals(g.F, blk, t1, t2)
    assert np.abs(r1.numpy() - g[tag + "_r1"]).max() < TOL
    assert np.abs(r2.numpy() - g[tag + "_r2"]).max() < TOL
    assert abs(to.cc_energy(g.F, blk, t1, t2) - float(g[tag + "_energy"])) < TOL
    # sharded ladder: rows [lo, hi) of <ab|ef> on each "rank"; pieces + the rest = the whole
    nv = g.nv
    cut = nv // 2
    _, rest = to.residuals(g.F, blk, t1, t2, skip_ladder=True)
    lad = 0.0
    for lo, hi in ((0, cut), (cut, nv)):
        sub = dict(g.blocks)
        sub["vvvv"] = g.blocks["vvvv"][lo:hi]
        _, piece = to.residuals(g.F, to.Blocks(sub, vv_rows=(lo, hi)), t1, t2, ladder_only=True)
        lad = lad + piece
    lad = lad + lad.permute(1, 0, 3, 2)
    assert np.abs((rest + lad).numpy() - g[tag + "_r2"]).max() < TOL


def test_torch_oracle_triplets(golden):
    from oracle import torch_oracle as to
    g = golden
    blk, tblk = oc.Blocks(g.blocks), to.Blocks(g.blocks)
    t1, t2 = g["gen_t1"], g["gen_t2"]
    no = g.no
    tot = 0.0
    for (i, j, k) in [(i, j, k) for i in range(no) for j in range(i + 1) for k in range(j + 1)]:
        e = to.et_triplet(g.F, tblk, t1, t2, i, j, k)
        assert abs(e - ot.et_triplet(g.F, blk, t1, t2, i, j, k)) < 1e-13
        tot += e
    assert abs(tot - float(g["gen_et_tjl"])) < TOL
