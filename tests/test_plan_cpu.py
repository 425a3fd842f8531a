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
