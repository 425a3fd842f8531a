"""
cctriples.py: (T) drivers, API-compatible with ``pycc/cctriples.py`` for the functions on the
CCSD(T) path: ``t_tjl(ccwfn)``, ``t_vikings(ccwfn)`` (drivers) and ``t3c_ijk`` / ``t3d_ijk``
(per-triplet builders taking the contraction operator, as in the reference).

For real orbitals the Lee-Rendell (``t_tjl``, :106-152) and "Vikings" (``t_vikings``, :156-186)
formulations give the same number (reference test ``test_005_ccsd_t_energy.py:38-40``); the
engine evaluates the Lee-Rendell form over i >= j >= k (minimum flops: 1/6.4 of ``t_vikings``)
and both names return it.
"""
import torch

from .triples_engine import TriplesEngine


def _engine(ccwfn):
    if getattr(ccwfn, "_triples", None) is None:
        ccwfn._triples = TriplesEngine(ccwfn.engine)
    return ccwfn._triples


def t_tjl(ccwfn):
    """Lee and Rendell's formulation (pycc/cctriples.py:106-152) on the amplitudes held by ``ccwfn``."""
    return _engine(ccwfn).energy_all_ranks()


def t_vikings(ccwfn):
    """Same energy as ``t_tjl`` for real orbitals (pycc/cctriples.py:156-186); see module docstring."""
    return t_tjl(ccwfn)


def _run_statements(contract, stmts, tensors, dtype):
    """Plan and launch statements over strided views (no temporaries): one pccb_contract launch each."""
    from . import _lib
    from .dsl import TRef, plan
    be = contract._backend()
    be.bind_stream()
    T = {n: TRef.of(t, n) for n, t in tensors.items()}
    dcode = _lib.F64 if dtype == torch.float64 else _lib.F32
    for st in stmts:
        for L in plan(st, T, dcode):
            be.contract(L.desc)


def _denominator(F, v, i, j, k, like):
    Fd = torch.diagonal(torch.as_tensor(F).to(like.device))
    Fv = Fd[v]
    return Fd[i] + Fd[j] + Fd[k] - (Fv.reshape(-1, 1, 1) + Fv.reshape(-1, 1) + Fv)


def t3c_ijk(o, v, i, j, k, t2, Wvvvo, Wovoo, F, contract, WithDenom=True):
    """Connected triples for one (i,j,k), same signature and result as pycc/cctriples.py:8-35.  The twelve
    terms run as the engine's three K-concatenated launches (triples_engine._STATEMENTS) over strided views of
    the arguments - Wvvvo[:,:,:,i] etc. are never copied, and no v^3 temporary is made per term."""
    from .triples_engine import _STATEMENTS
    dev = contract._backend().device
    t2 = torch.as_tensor(t2).to(dev)
    Wvvvo, Wovoo = torch.as_tensor(Wvvvo).to(dev), torch.as_tensor(Wovoo).to(dev)
    nv = t2.shape[2]
    W = torch.empty((nv, nv, nv), dtype=t2.dtype, device=dev)
    V = lambda x: Wvvvo[:, :, :, x]                      # [x,y,e] = <xy|e x>
    O = lambda x, y: Wovoo[:, :, x, y]                   # [m,c]   = <mc|xy>
    T = dict(W=W, Vi=V(i), Vj=V(j), Vk=V(k),
             tij=t2[i, j], tik=t2[i, k], tji=t2[j, i], tjk=t2[j, k], tki=t2[k, i], tkj=t2[k, j],
             ti=t2[i], tj=t2[j], tTj=t2[j].permute(0, 2, 1), tTk=t2[k].permute(0, 2, 1),
             Oij=O(i, j), Oik=O(i, k), Oji=O(j, i), Ojk=O(j, k), Oki=O(k, i), Okj=O(k, j))
    _run_statements(contract, _STATEMENTS, T, t2.dtype)
    if WithDenom is True:
        return W / _denominator(F, v, i, j, k, W)
    return W


# t3d_ijk (pycc/cctriples.py:69-74): six outer products, paired by their output structure
_DISCONNECTED = (
    "D[abc] = gij[ab]*t1k[c] + tij[ab]*fk[c]",
    "D[abc] += gik[ac]*t1j[b] + tik[ac]*fj[b]",
    "D[abc] += gjk[bc]*t1i[a] + tjk[bc]*fi[a]",
)


def t3d_ijk(o, v, i, j, k, t1, t2, Woovv, F, contract, WithDenom=True):
    """Disconnected triples for one (i,j,k) (pycc/cctriples.py:68-83): three launches of two outer products."""
    dev = contract._backend().device
    t1, t2 = torch.as_tensor(t1).to(dev), torch.as_tensor(t2).to(dev)
    Woovv = torch.as_tensor(Woovv).to(dev)
    F = torch.as_tensor(F).to(device=dev, dtype=t2.dtype)
    nv = t2.shape[2]
    D = torch.empty((nv, nv, nv), dtype=t2.dtype, device=dev)
    T = dict(D=D, gij=Woovv[i, j], gik=Woovv[i, k], gjk=Woovv[j, k], tij=t2[i, j], tik=t2[i, k], tjk=t2[j, k],
             t1i=t1[i], t1j=t1[j], t1k=t1[k], fi=F[i, v], fj=F[j, v], fk=F[k, v])
    _run_statements(contract, _DISCONNECTED, T, t2.dtype)
    if WithDenom is True:
        return D / _denominator(F, v, i, j, k, D)
    return D
