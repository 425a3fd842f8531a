"""Complex-amplitude residuals: the plane expansion of the engines' statements (host logic) on the
NumPy interpreter back end, against reference-generated vectors."""
import pytest

from tests.interp_backend import InterpBackend
from tests import complex_checks as cc
import pycc_b200
from pycc_b200.complex_engine import expand


def make_wfn(g, model):
    return pycc_b200.ccwfn.from_arrays(g.F, g.blocks, g.no, model=model, _backend=InterpBackend())


@pytest.mark.parametrize("model", cc.MODELS)
@pytest.mark.parametrize("name", cc.CASES)
def test_complex_residuals(name, model):
    cc.check_complex_residuals(make_wfn, name, model)


@pytest.mark.parametrize("name", cc.CC3_CASES)
def test_complex_cc3_residuals(name):
    cc.check_complex_cc3_residuals(make_wfn, name)


@pytest.mark.parametrize("name", cc.CC3_CASES)
def test_complex_cc3_lambda_residuals(name):
    cc.check_complex_cc3_lambda_residuals(make_wfn, name)


def test_expand_rules():
    cplx = lambda n: n in ("C", "A", "B", "D")
    acts = expand("C[ij] = D[ij] + 0.5*A[ik]*R[jk] - A[ik]*B[jk] + S[ij]", cplx)
    assert acts[0] == ("zero", "C")
    texts = [a[1] for a in acts[1:]]
    assert texts[0].startswith("C[zij] +=") and "D[zij]" in texts[0] and "A[zik]*R[jk]" in texts[0]
    assert texts[1].startswith("C__r[ij] +=") and "S[ij]" in texts[1] and "A__i[ik]*B__i[jk]" in texts[1]
    assert texts[2].startswith("C__i[ij] +=") and "A__r[ik]*B__i[jk]" in texts[2] and "A__i[ik]*B__r[jk]" in texts[2]
    assert expand("C[ij] += R[ik]*S[jk]", cplx) == [("plan", "C__r[ij] += 1.0*R[ik]*S[jk]")]
