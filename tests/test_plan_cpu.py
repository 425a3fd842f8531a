"""Host logic of the CCSD engine (contraction plans emitted by pycc_b200.dsl / engine) checked
on the CPU: the descriptors are interpreted with NumPy (tests/interp_backend.py) and compared
with the reference-generated golden vectors.  No GPU, no CUDA library involved."""
import numpy as np
import pytest
import torch

from tests.interp_backend import InterpBackend
from pycc_b200.engine import CCSDEngine

TOL = 1e-12


def _engine(g):
    return CCSDEngine(InterpBackend(), g.F, g.blocks, g.no, g.nv)


@pytest.mark.parametrize("tag", ["sym", "gen"])
def test_residual_plan_matches_reference(golden, tag):
    g = golden
    eng = _engine(g)
    eng.set_amplitudes(g[tag + "_t1"], g[tag + "_t2"])
    r1, r2 = eng.residuals()
    got = dict(Fae=eng.Fae, Fmi=eng.Fmi, Fme=eng.Fme, Wmnij=eng.Wmnij, Wmbej=eng.Wmbej_view(),
               Wmbje=eng.Wmbje_view(), Zmbij=eng.Zmbij_view())
    for k, t in got.items():
        assert np.abs(t.numpy() - g[tag + "_" + k]).max() < TOL, k
    assert np.abs(r1.numpy() - g[tag + "_r1"]).max() < TOL
    assert np.abs(r2.numpy() - g[tag + "_r2"]).max() < TOL
    assert abs(eng.energy() - float(g[tag + "_energy"])) < TOL


@pytest.mark.parametrize("tag", ["sym", "gen"])
def test_triples_plan_matches_reference(golden, tag):
    """The three batched W launches per occupied pair (pointer-patched templates) + energy."""
    from pycc_b200.triples_engine import TriplesEngine
    from oracle import triples_oracle as ot
    from oracle.ccsd_oracle import Blocks
    g = golden
    eng = _engine(g)
    eng.set_amplitudes(g[tag + "_t1"], g[tag + "_t2"])
    tri = TriplesEngine(eng)
    # W of the last launched pair against the oracle's connected triples
    et, per, trip = tri.energy()
    i, j = tri.pairs()[-1]
    blk = Blocks(g.blocks)
    for k in range(j + 1):
        Wref = ot.w_connected(blk, g[tag + "_t2"], i, j, k)
        assert np.abs(tri.W[k].numpy() - Wref).max() < 1e-12
    assert abs(et - float(g[tag + "_et_tjl"])) < 1e-12
    assert len(trip) == g.no * (g.no + 1) * (g.no + 2) // 6


def test_ring_terms_are_plain_matrix_products():
    """The ring-term statements over the permuted copies (engine.STATEMENTS_SP) must plan into launches whose
    operands are plain K-contiguous matrices in BOTH precisions - that is what makes them eligible for the
    TMA-staged FP64 kernel and the tcgen05 FP32 kernel (csrc/contract.cu tma_operand_ok, contract_tc05.cu
    operand_ok).  Guards the planner's tie-break (dsl._flatten prefer_flat) against silent regressions."""
    import contextlib
    import io
    import pycc_b200
    from tests.conftest import Golden
    from tests.interp_backend import InterpBackend

    def plain(c, ext):
        return c.div <= 1 or c.div >= ext or c.s_hi == c.div * c.s_lo

    def stride(c, ext):
        return c.s_hi if (c.div == 1 and ext > 1) else c.s_lo

    g = Golden("o6v13")
    for precision in ("DP", "SP"):
        with contextlib.redirect_stdout(io.StringIO()):
            w = pycc_b200.ccwfn.from_arrays(g.F, g.blocks, g.no, precision=precision, _backend=InterpBackend())
        assert w.engine.sp_plain
        big = 0
        for name in ("Wmbej", "Wmbje", "Zmbij", "Wmnij", "r_T2"):
            for L in w.engine.plans[name]:
                d = L.desc
                if max(d.seg[q].K for q in range(d.nseg)) < g.no * g.nv:      # the o^3 v^3 / o^2 v^4 class only
                    continue
                big += 1
                assert d.batch == 1, L.text
                assert len({d.seg[q].scale for q in range(d.nseg)}) == 1, L.text     # one scale per launch
                for q in range(d.nseg):
                    sg = d.seg[q]
                    assert plain(sg.am, d.M) and plain(sg.bn, d.N) and plain(sg.ak, sg.K) and plain(sg.bk, sg.K), L.text
                    assert stride(sg.ak, sg.K) == 1 and stride(sg.bk, sg.K) == 1, L.text
        assert big == 7
