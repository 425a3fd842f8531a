"""HBAR and Lambda host logic (plans, solver control flow) on the NumPy interpreter back end,
against reference-generated vectors."""
import numpy as np
import pytest

from tests.interp_backend import InterpBackend
from tests import lambda_checks as lc
import pycc_b200


def make_wfn(g, model):
    return pycc_b200.ccwfn.from_arrays(g.F, g.blocks, g.no, model=model, _backend=InterpBackend())


@pytest.mark.parametrize("model", lc.MODELS)
@pytest.mark.parametrize("name", lc.CASES)
def test_hbar_blocks_and_lambda_residuals(name, model):
    lc.check_blocks_and_residuals(make_wfn, name, model)


@pytest.mark.parametrize("model", lc.MODELS)
@pytest.mark.parametrize("name", ["o3v5_offdiag", "o4v8", "o6v13"])
def test_solve_lambda(name, model):
    lc.check_solve(make_wfn, name, model, chain=False)


def test_solve_chain():
    lc.check_solve(make_wfn, "o4v8", "CCSD", chain=True)


def test_unsupported_models():
    g, _ = lc.load("o3v5_offdiag")
    import contextlib, io
    with contextlib.redirect_stdout(io.StringIO()):
        w = make_wfn(g, "CCSD(T)")
    with pytest.raises(Exception, match="HBAR in the B200 engine"):
        pycc_b200.cchbar(w)
