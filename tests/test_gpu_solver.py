"""GPU parity of the full path through the public API (ccwfn / cctriples) and the C ABI:
golden vectors produced by the reference itself, then larger seeded systems against the CPU
oracle (which is pinned to those golden vectors by tests/test_oracle_golden.py).
FP64 bars (BASELINE.json north_star): energies within 1e-10 Eh, T1/T2 within 1e-9 max-abs."""
import contextlib
import io

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

E_TOL, T_TOL = 1e-10, 1e-9


def _wfn(F, blocks, no, model="CCSD"):
    import pycc_b200
    with contextlib.redirect_stdout(io.StringIO()):
        return pycc_b200.ccwfn.from_arrays(F, blocks, no, model=model)


def test_golden_solve_and_triples(golden):
    import pycc_b200
    g = golden
    w = _wfn(g.F, g.blocks, g.no, "CCSD(T)")
    with contextlib.redirect_stdout(io.StringIO()):
        e = w.solve_cc(e_conv=1e-12, r_conv=1e-12, maxiter=100)
    ref_log = g["solve_iter_log"]
    assert len(w.iteration_log) == ref_log.shape[0]
    for (it, ecc, de, rms), row in zip(w.iteration_log, ref_log):
        assert abs(ecc - row[0]) < 1e-11
    assert abs(e - float(g["solve_eccsd_t_total"])) < E_TOL
    assert np.abs(w.t1.cpu().numpy() - g["solve_t1"]).max() < T_TOL
    assert np.abs(w.t2.cpu().numpy() - g["solve_t2"]).max() < T_TOL
    assert abs(pycc_b200.t_tjl(w) - float(g["solve_et"])) < E_TOL
    assert abs(pycc_b200.t_vikings(w) - float(g["solve_et_vikings"])) < E_TOL


@pytest.mark.parametrize("tag", ["sym", "gen"])
def test_golden_triples_on_test_amplitudes(golden, tag):
    import pycc_b200
    g = golden
    w = _wfn(g.F, g.blocks, g.no)
    w._load(None, g[tag + "_t1"], g[tag + "_t2"])
    assert abs(pycc_b200.t_tjl(w) - float(g[tag + "_et_tjl"])) < E_TOL


def test_golden_builder_api_and_operator(golden):
    import pycc_b200
    g = golden
    w = _wfn(g.F, g.blocks, g.no)
    t1, t2 = g["gen_t1"], g["gen_t2"]
    o, v, H, F = w.o, w.v, w.H, g.F
    got = dict(Fae=w.build_Fae(o, v, F, H.L, t1, t2), Fmi=w.build_Fmi(o, v, F, H.L, t1, t2),
               Fme=w.build_Fme(o, v, F, H.L, t1), Wmnij=w.build_Wmnij(o, v, H.ERI, t1, t2),
               Wmbej=w.build_Wmbej(o, v, H.ERI, H.L, t1, t2), Wmbje=w.build_Wmbje(o, v, H.ERI, t1, t2),
               Zmbij=w.build_Zmbij(o, v, H.ERI, t1, t2))
    for k, t in got.items():
        assert np.abs(t.cpu().numpy() - g["gen_" + k]).max() < 1e-12, k
    r1, r2 = w.residuals(F, t1, t2)
    assert np.abs(r1.cpu().numpy() - g["gen_r1"]).max() < 1e-12
    assert np.abs(r2.cpu().numpy() - g["gen_r2"]).max() < 1e-12
    assert float(w.t1.abs().max()) == 0.0 and torch.equal(w.H.F, w.engine.F)      # MP2 guess and H.F untouched
    # the builders above are pure (the object's own amplitudes were put back): the operator checks below want
    # the test amplitudes on the device
    w._load(None, t1, t2)
    c = w.contract
    x = c('ijef,abef->ijab', w.t2, H.ERI[v, v, v, v])
    assert np.abs(x.cpu().numpy() - np.einsum('ijef,abef->ijab', t2, g.blocks["vvvv"])).max() < 1e-12
    assert abs(float(c('ia,ia->', w.t1, w.t1)) - float((t1 ** 2).sum())) < 1e-13
    W = pycc_b200.t3c_ijk(o, v, g.no - 1, 1, 0, w.t2, H.ERI[v, v, v, o], H.ERI[o, v, o, o], H.F, c, False)
    from oracle import triples_oracle as ot
    from oracle.ccsd_oracle import Blocks
    assert np.abs(W.cpu().numpy() - ot.w_connected(Blocks(g.blocks), t2, g.no - 1, 1, 0)).max() < 1e-12
    from pycc_b200.cctriples import t3d_ijk
    D = t3d_ijk(o, v, g.no - 1, 1, 0, w.t1, w.t2, H.ERI[o, o, v, v], H.F, c, False)
    assert np.abs(D.cpu().numpy() - ot.v_disconnected(F, Blocks(g.blocks), t1, t2, g.no - 1, 1, 0)).max() < 1e-12
    Dd = t3d_ijk(o, v, g.no - 1, 1, 0, w.t1, w.t2, H.ERI[o, o, v, v], H.F, c, True)
    den = ot.triples_denominator(F, g.no, g.no - 1, 1, 0)
    assert np.abs(Dd.cpu().numpy() - ot.v_disconnected(F, Blocks(g.blocks), t1, t2, g.no - 1, 1, 0) / den).max() < 1e-11


@pytest.mark.parametrize("no,nv", [(1, 1), (1, 3), (2, 1), (3, 2), (1, 17), (17, 2)])
def test_degenerate_shapes_vs_oracle(no, nv):
    """Edge cases: a single occupied or virtual orbital, more occupied than virtual orbitals - every tile ragged, every
    index group of extent one somewhere.  Converged CCSD(T) against the oracle."""
    import pycc_b200
    from pycc_b200.synthetic import SyntheticSystem
    from oracle import ccsd_oracle as oc, triples_oracle as ot
    s = SyntheticSystem(no, nv, seed=3 * no + nv, fock_offdiag=True)
    w = _wfn(s.F, s.blocks, no)
    with contextlib.redirect_stdout(io.StringIO()):
        e = w.solve_cc(e_conv=1e-12, r_conv=1e-12, maxiter=200)
    eo, t1, t2, niter = oc.solve_cc(s.F, s.blocks, 1e-12, 1e-12, maxiter=200)
    if no * nv + (no * nv) ** 2 > 16:        # fewer unknowns than DIIS vectors: the singular DIIS system amplifies rounding
        assert len(w.iteration_log) == niter
    assert abs(e - eo) < E_TOL
    assert np.abs(w.t2.cpu().numpy() - t2).max() < T_TOL
    assert abs(pycc_b200.t_tjl(w) - ot.t_tjl(s.F, s.blocks, t1, t2)) < 1e-12


@pytest.mark.parametrize("no,nv,offdiag", [(8, 30, True), (10, 44, False), (7, 51, True)])
def test_seeded_systems_vs_oracle(no, nv, offdiag):
    """Sizes the reference itself would need minutes for; the oracle (pinned) does them in seconds."""
    import pycc_b200
    from pycc_b200.synthetic import SyntheticSystem
    from oracle import ccsd_oracle as oc, triples_oracle as ot
    s = SyntheticSystem(no, nv, seed=no + nv, fock_offdiag=offdiag)
    w = _wfn(s.F, s.blocks, no)
    with contextlib.redirect_stdout(io.StringIO()):
        e = w.solve_cc(e_conv=1e-11, r_conv=1e-11, maxiter=100)
    log = []
    eo, t1, t2, niter = oc.solve_cc(s.F, s.blocks, 1e-11, 1e-11, maxiter=100, log=log)
    assert len(w.iteration_log) == niter
    assert abs(e - eo) < E_TOL
    assert np.abs(w.t1.cpu().numpy() - t1).max() < T_TOL
    assert np.abs(w.t2.cpu().numpy() - t2).max() < T_TOL
    # (T): full sum at the small shape, a seeded sample of triplets otherwise
    tri = pycc_b200.cctriples._engine(w)
    rng = np.random.default_rng(0)
    pairs = tri.pairs()
    sample = [pairs[q] for q in sorted(rng.choice(len(pairs), size=min(6, len(pairs)), replace=False))]
    et, per, trip = tri.energy(sample)
    blk = oc.Blocks(s.blocks)
    for x, (i, j, k) in zip(per.tolist(), trip):
        assert abs(x - ot.et_triplet(s.F, blk, t1, t2, i, j, k)) < 1e-12
    # size-independent property: the sum over all pairs equals the sum of any partition of them
    et_all = pycc_b200.t_tjl(w)
    half = len(pairs) // 2
    et_a, _, _ = tri.energy(pairs[:half])
    et_b, _, _ = tri.energy(pairs[half:])
    assert abs(et_all - (et_a + et_b)) < 1e-13


def test_config2_o20_v150_converged_solve_vs_oracle():
    """BASELINE.json configs[1] ("synthetic RHF-CCSD o=20 v=150 FP64 on 1xB200 ... correctness vs CPU
    reference"), the bench's own seeded system: converged solve against the pinned NumPy oracle -
    same iteration count, |dE| <= 1e-10 Eh, max|dT1|, max|dT2| <= 1e-9 - and E(T) per triplet on a seeded
    sample against the oracle's t_tjl body.  About a minute of host time for the oracle."""
    import pycc_b200
    from pycc_b200.synthetic import SyntheticSystem
    from oracle import ccsd_oracle as oc, triples_oracle as ot
    no, nv = 20, 150
    s = SyntheticSystem(no, nv, seed=0)
    w = _wfn(s.F, s.blocks, no, "CCSD(T)")
    with contextlib.redirect_stdout(io.StringIO()):
        e = w.solve_cc(e_conv=1e-10, r_conv=1e-10, maxiter=60)
    et_gpu = e - w.iteration_log[-1][1]
    eo, t1, t2, niter = oc.solve_cc(s.F, s.blocks, 1e-10, 1e-10, maxiter=60)
    assert eo is not None and len(w.iteration_log) == niter
    assert abs(w.iteration_log[-1][1] - eo) < E_TOL
    assert np.abs(w.t1.cpu().numpy() - t1).max() < T_TOL
    assert np.abs(w.t2.cpu().numpy() - t2).max() < T_TOL
    tri = pycc_b200.cctriples._engine(w)
    rng = np.random.default_rng(20150)
    pairs = tri.pairs()
    sample = [pairs[q] for q in sorted(rng.choice(len(pairs), size=3, replace=False))]
    _, per, trip = tri.energy(sample)
    blk = oc.Blocks(s.blocks)
    pick = sorted(rng.choice(len(trip), size=min(6, len(trip)), replace=False))
    for q in pick:
        i, j, k = trip[q]
        assert abs(per[q].item() - ot.et_triplet(s.F, blk, t1, t2, i, j, k)) < 1e-12
    # the full (T) sum is bit-reproducible and additive over any partition of the pairs
    half = len(pairs) // 2
    assert abs(et_gpu - (tri.energy(pairs[:half])[0] + tri.energy(pairs[half:])[0])) < 1e-12


def test_residual_properties_at_scale():
    """(16, 100): no CPU oracle in the test budget -> size-independent properties.
    (1) pair symmetry r2[ijab] = r2[jiba]; (2) the ladder term equals the factorised checker
    sum_Q built from the density-fitting factor (O(o^2 v^3 naux) instead of o^2 v^4);
    (3) the tensor-core kernel agrees with the naive one-thread-per-element kernel."""
    from pycc_b200.synthetic import SyntheticSystem
    no, nv = 16, 100
    s = SyntheticSystem(no, nv, seed=5)
    w = _wfn(s.F, s.blocks, no)
    eng, be = w.engine, w.backend
    g = torch.Generator().manual_seed(1)
    t1 = 0.05 * torch.randn(no, nv, generator=g, dtype=torch.float64)
    t2 = 0.02 * torch.randn(no, no, nv, nv, generator=g, dtype=torch.float64)
    t2 = 0.5 * (t2 + t2.permute(1, 0, 3, 2))
    r1, r2 = w.residuals(s.F, t1, t2)
    assert (r2 - r2.permute(1, 0, 3, 2)).abs().max().item() < 1e-12
    be.use_naive_contract(True)
    try:
        n1, n2 = w.residuals(s.F, t1, t2)
    finally:
        be.use_naive_contract(False)
    assert (r1 - n1).abs().max().item() < 1e-11
    assert (r2 - n2).abs().max().item() < 1e-11
    # ladder through the factor: sum_ef tau[ijef] <ab|ef> = sum_Q (sum_e tau[ijef] B[Q,a,e]) B[Q,b,f]
    from pycc_b200.dsl import plan
    out = torch.zeros_like(eng.r2)
    eng.prepare()
    for L in plan("out[ijab] = tau[ijef]*vvvv[abef]", dict(out=out, tau=eng.tau, vvvv=eng.blk["vvvv"])):
        be.contract(L.desc)
    Bv = torch.as_tensor(s.B[:, no:, no:]).to(be.device)
    chk = torch.einsum('ijef,Qae,Qbf->ijab', eng.tau, Bv, Bv)
    assert (out - chk).abs().max().item() < 1e-11


def test_mixed_precision_mode(golden):
    """precision='SP': FP32 storage + 3xTF32 contractions, FP64 energy / norm / DIIS.
    Bar (BASELINE.json north_star): |E - E_DP(reference)| <= 1e-6 Eh; the reference's own
    all-float32 SP energy on the same input is reported next to it."""
    import pycc_b200
    g = golden
    with contextlib.redirect_stdout(io.StringIO()):
        w = pycc_b200.ccwfn.from_arrays(g.F, g.blocks, g.no, model="CCSD(T)", precision="SP")
        e = w.solve_cc(e_conv=1e-7, r_conv=1e-7, maxiter=100)
    assert e is not None
    assert w.t2.dtype == torch.float32
    assert abs(e - float(g["solve_eccsd_t_total"])) < 1e-6
    assert abs((e - pycc_b200.t_tjl(w)) - float(g["solve_sp_ecc"])) < 5e-6     # reference SP mode, all float32
    assert np.abs(w.t2.cpu().numpy().astype(np.float64) - g["solve_t2"]).max() < 1e-5


@pytest.mark.parametrize("no,nv,seed", [(8, 36, 11), (16, 24, 12)])
def test_mixed_precision_seeded(no, nv, seed):
    """(16,24): o^2 = 256 makes the K-contiguous o^4 v^2 term WmnijT[ijmn] tauT[abmn] eligible for the tcgen05 kernel."""
    import pycc_b200
    from pycc_b200.synthetic import SyntheticSystem
    from oracle import ccsd_oracle as oc
    s = SyntheticSystem(no, nv, seed=seed, fock_offdiag=True)
    with contextlib.redirect_stdout(io.StringIO()):
        w = pycc_b200.ccwfn.from_arrays(s.F, s.blocks, no, precision="SP")
        e = w.solve_cc(e_conv=1e-7, r_conv=1e-7, maxiter=100)
    eo, t1, t2, niter = oc.solve_cc(s.F, s.blocks, 1e-10, 1e-10)
    assert abs(e - eo) < 1e-6
    # the 3xTF32 kernel against the plain-FP32 naive kernel on one residual call
    be = w.backend
    r1, r2 = w.residuals(s.F, t1, t2)
    be.use_naive_contract(True)
    try:
        n1, n2 = w.residuals(s.F, t1, t2)
    finally:
        be.use_naive_contract(False)
    r1o, r2o = oc.residuals(s.F, oc.Blocks(s.blocks), t1, t2)
    assert np.abs(r2.cpu().numpy() - r2o).max() < 2e-6
    assert np.abs(n2.cpu().numpy() - r2o).max() < 2e-6
    assert np.abs(r1.cpu().numpy() - r1o).max() < 2e-6


@pytest.mark.parametrize("model", ["CCD", "CC2"])
@pytest.mark.parametrize("name", ["o4v8", "o5v19_offdiag", "o6v13"])
def test_ccd_cc2_models(name, model):
    """CCD / CC2 branches of the builders on the GPU against the reference's own vectors."""
    import os
    import pycc_b200
    from tests.conftest import GOLDEN_DIR, Golden
    g = Golden(name)
    z = np.load(os.path.join(GOLDEN_DIR, "models", name + ".npz"))
    tag = model.lower()
    w = _wfn(g.F, g.blocks, g.no, model)
    t1 = np.zeros_like(g["gen_t1"]) if model == "CCD" else g["gen_t1"]
    r1, r2 = w.residuals(g.F, t1, g["gen_t2"])
    assert np.abs(r1.cpu().numpy() - z[tag + "_r1"]).max() < 1e-12
    assert np.abs(r2.cpu().numpy() - z[tag + "_r2"]).max() < 1e-12
    w = _wfn(g.F, g.blocks, g.no, model)
    with contextlib.redirect_stdout(io.StringIO()):
        e = w.solve_cc(e_conv=1e-12, r_conv=1e-12, maxiter=100)
    assert len(w.iteration_log) == int(z[tag + "_solve_niter"])
    assert abs(e - float(z[tag + "_solve_ecc"])) < E_TOL
    assert np.abs(w.t1.cpu().numpy() - z[tag + "_solve_t1"]).max() < T_TOL
    assert np.abs(w.t2.cpu().numpy() - z[tag + "_solve_t2"]).max() < T_TOL


@pytest.mark.parametrize("name", ["o4v8", "o5v19_offdiag", "o6v13"])
def test_cc3_model(name):
    """CC3 (per-iteration triples loop with T1-dressed integrals) on the GPU against the reference's vectors."""
    import os
    from tests.conftest import GOLDEN_DIR, Golden
    g = Golden(name)
    z = np.load(os.path.join(GOLDEN_DIR, "models", name + ".npz"))
    w = _wfn(g.F, g.blocks, g.no, "CC3")
    r1, r2 = w.residuals(g.F, g["gen_t1"], g["gen_t2"])
    assert np.abs(r1.cpu().numpy() - z["cc3_r1"]).max() < 1e-12
    assert np.abs(r2.cpu().numpy() - z["cc3_r2"]).max() < 1e-12
    c = w.engine.cc3
    assert np.abs(c.Wv.permute(1, 2, 3, 0).cpu().numpy() - z["cc3_Wabei"]).max() < 1e-12
    w = _wfn(g.F, g.blocks, g.no, "CC3")
    with contextlib.redirect_stdout(io.StringIO()):
        e = w.solve_cc(e_conv=1e-12, r_conv=1e-12, maxiter=100)
    assert len(w.iteration_log) == int(z["cc3_solve_niter"])
    assert abs(e - float(z["cc3_solve_ecc"])) < E_TOL
    assert np.abs(w.t1.cpu().numpy() - z["cc3_solve_t1"]).max() < T_TOL
    assert np.abs(w.t2.cpu().numpy() - z["cc3_solve_t2"]).max() < T_TOL


@pytest.mark.parametrize("name", ["o3v5_offdiag", "o4v8", "o6v13"])
def test_cc3_real_time_field(name):
    """CC3 residuals with a field-dressed Fock matrix and real_time=True: the t3_pert_ijk term (pycc/ccwfn.py:378-386,
    cctriples.py:547-562) against the reference's own vectors; the same call without real_time stays the plain CC3."""
    import os
    from tests.conftest import GOLDEN_DIR, Golden
    g = Golden(name)
    z = np.load(os.path.join(GOLDEN_DIR, "rtcc3", name + ".npz"))
    w = _wfn(g.F, g.blocks, g.no, "CC3")
    r1, r2 = w.residuals(z["F_field"], z["t1"], z["t2"], real_time=True)
    assert np.abs(z["r2"] - z["r2_plain"]).max() > 1e-6
    assert np.abs(r1.cpu().numpy() - z["r1"]).max() < 1e-12
    assert np.abs(r2.cpu().numpy() - z["r2"]).max() < 1e-12
    p1, p2 = w.residuals(z["F_field"], z["t1"], z["t2"])
    assert np.abs(p1.cpu().numpy() - z["r1_plain"]).max() < 1e-12
    assert np.abs(p2.cpu().numpy() - z["r2_plain"]).max() < 1e-12


def test_amplitude_shard_roundtrip(golden):
    """Engine-only API used by bench.py's end-to-end leg: the amplitude vector goes to pinned host
    memory and back (one GPU: the shard is the whole vector) without changing a bit."""
    g = golden
    w = _wfn(g.F, g.blocks, g.no, "CCSD")
    w._load(None, g["gen_t1"], g["gen_t2"])
    lo, hi = w.amplitude_shard()
    assert (lo, hi) == (0, w.t1.numel() + w.t2.numel())
    host = torch.empty(hi - lo, dtype=torch.float64).pin_memory()
    w.store_amplitude_shard(host)
    torch.cuda.synchronize()
    assert np.array_equal(host[:w.t1.numel()].numpy().reshape(g.no, g.nv), g["gen_t1"])
    w.t1.zero_()
    w.t2.zero_()
    w.load_amplitude_shard(host)
    torch.cuda.synchronize()
    assert np.array_equal(w.t2.cpu().numpy(), g["gen_t2"]) and np.array_equal(w.t1.cpu().numpy(), g["gen_t1"])


def test_integral_front_end():
    """Hamiltonian.from_factor (DF block provider, per-rank rows) and from_ao (AO -> MO quarter transformations) on
    the device, against numpy and the oracle (pycc/hamiltonian.py:26-39 without psi4)."""
    from pycc_b200.backend import CudaBackend
    from tests import frontend_checks as fc
    be = CudaBackend()
    fc.check_factor_blocks(be)
    fc.check_ao_transform(be)
