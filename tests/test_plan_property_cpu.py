"""
Property test of the contraction planner (pycc_b200/dsl.py): random pairwise contractions - random index sets,
random index ORDER inside every tensor, permuted (strided) storage of operands and output, an optional addend, accumulate forms, one or
two product terms - are planned into pccb_contraction descriptors, interpreted with NumPy
(tests/interp_backend.py) and compared with numpy.einsum, the semantics of the reference's operator seam
cc_contract (pycc/utils.py:162-187).  The planner may refuse a layout (PlanError: an index group that does not
fit two stride levels) but must never emit descriptors that compute something else.
"""
import numpy as np
import torch
from hypothesis import HealthCheck, given, settings, strategies as st

from pycc_b200 import _lib
from pycc_b200.dsl import PlanError, plan
from tests.interp_backend import InterpBackend

LETTERS = "abcdefgh"


@st.composite
def contractions(draw):
    n_m = draw(st.integers(0, 2))
    n_n = draw(st.integers(0, 2))
    n_k = draw(st.integers(0, 3))
    n_b = draw(st.integers(0, 1))
    if n_m + n_n + n_b == 0:
        n_m = 1
    pool = list(LETTERS)
    take = lambda n: [pool.pop(0) for _ in range(n)]
    M, N, K, Bt = take(n_m), take(n_n), take(n_k), take(n_b)
    ext = {c: draw(st.integers(1, 4)) for c in M + N + K + Bt}
    perm = lambda xs: draw(st.permutations(xs)) if xs else []
    a_idx = "".join(perm(M + K + Bt))
    b_idx = "".join(perm(N + K + Bt))
    c_idx = "".join(perm(M + N + Bt))
    if not a_idx or not b_idx:            # a factor without indices is not a tensor factor
        a_idx = a_idx or None
        b_idx = b_idx or None
    two_terms = draw(st.booleans())
    addend = draw(st.booleans())
    op = draw(st.sampled_from(["=", "+=", "-="]))
    coef = draw(st.sampled_from([1.0, -1.0, 0.5, 2.0]))
    store = {}
    for name, idx in (("A", a_idx), ("B", b_idx), ("A2", a_idx), ("B2", b_idx), ("D", c_idx), ("C", c_idx)):
        if idx is None:
            continue
        store[name] = "".join(draw(st.permutations(list(idx))))          # memory order of the tensor
    seed = draw(st.integers(0, 2 ** 31 - 1))
    return dict(a=a_idx, b=b_idx, c=c_idx, ext=ext, two=two_terms, addend=addend, op=op, coef=coef, store=store, seed=seed)


def _tensor(rng, idx, mem_order, ext):
    """A tensor indexed `idx` whose memory order is `mem_order` (a permuted view of a contiguous array)."""
    base = torch.from_numpy(rng.standard_normal([ext[c] for c in mem_order]))
    return base.permute([mem_order.index(c) for c in idx])


@settings(max_examples=400, deadline=None, derandomize=True, suppress_health_check=[HealthCheck.too_slow])
@given(contractions())
def test_planned_descriptors_compute_the_einsum(case):
    if case["a"] is None or case["b"] is None:
        return
    rng = np.random.default_rng(case["seed"])
    ext, st_ = case["ext"], case["store"]
    T = {"A": _tensor(rng, case["a"], st_["A"], ext), "B": _tensor(rng, case["b"], st_["B"], ext)}
    terms = [(case["coef"], "A", "B")]
    if case["two"]:
        T["A2"] = _tensor(rng, case["a"], st_["A2"], ext)
        T["B2"] = _tensor(rng, case["b"], st_["B2"], ext)
        terms.append((1.0, "A2", "B2"))
    rhs = ""
    if case["addend"]:
        T["D"] = _tensor(rng, case["c"], st_["D"], ext)
        rhs = "D[%s]" % case["c"]
    for c, x, y in terms:
        sign = (" - " if c < 0 else " + ") if rhs else ("-" if c < 0 else "")
        rhs += "%s%r*%s[%s]*%s[%s]" % (sign, abs(c), x, case["a"], y, case["b"])
    C = _tensor(rng, case["c"], st_["C"], ext)          # the output may be a permuted view as well
    T["C"] = C
    C0 = C.clone()
    stmt = "C[%s] %s %s" % (case["c"], case["op"], rhs)
    want = np.zeros(tuple(C.shape))
    for c, x, y in terms:
        want = want + c * np.einsum("%s,%s->%s" % (case["a"], case["b"], case["c"]), T[x].numpy(), T[y].numpy())
    if case["addend"]:
        want = want + T["D"].numpy()
    want = {"=": want, "+=": C0.numpy() + want, "-=": C0.numpy() - want}[case["op"]]
    try:
        launches = plan(stmt, T, _lib.F64)
    except PlanError as err:
        # a refused LAYOUT is fine (an index group that does not fit two stride levels); a wrong answer is not,
        # and neither is a statement the parser does not understand
        assert "more than two levels" in str(err) or "do not flatten" in str(err), (stmt, str(err))
        return
    be = InterpBackend()
    for L in launches:
        be.contract(L.desc)
    assert np.abs(C.numpy() - want).max() <= 1e-12 * max(1.0, np.abs(want).max()), stmt
