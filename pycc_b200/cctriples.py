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


def t3c_ijk(o, v, i, j, k, t2, Wvvvo, Wovoo, F, contract, WithDenom=True):
    """Connected triples for one (i,j,k) through the contraction operator, term by term as in
    pycc/cctriples.py:8-35 (each ``contract`` is one pccb_contract launch on strided views)."""
    t3 = contract('bae,ce->abc', Wvvvo[:, :, :, i], t2[k, j])
    t3 += contract('cae,be->abc', Wvvvo[:, :, :, i], t2[j, k])
    t3 += contract('ace,be->abc', Wvvvo[:, :, :, k], t2[j, i])
    t3 += contract('bce,ae->abc', Wvvvo[:, :, :, k], t2[i, j])
    t3 += contract('cbe,ae->abc', Wvvvo[:, :, :, j], t2[i, k])
    t3 += contract('abe,ce->abc', Wvvvo[:, :, :, j], t2[k, i])
    t3 -= contract('mc,mab->abc', Wovoo[:, :, j, k], t2[i])
    t3 -= contract('mb,mac->abc', Wovoo[:, :, k, j], t2[i])
    t3 -= contract('mb,mca->abc', Wovoo[:, :, i, j], t2[k])
    t3 -= contract('ma,mcb->abc', Wovoo[:, :, j, i], t2[k])
    t3 -= contract('ma,mbc->abc', Wovoo[:, :, k, i], t2[j])
    t3 -= contract('mc,mba->abc', Wovoo[:, :, i, k], t2[j])
    if WithDenom is True:
        Fd = torch.diagonal(torch.as_tensor(F).to(t3.device))
        Fv = Fd[v]
        denom = Fd[i] + Fd[j] + Fd[k] - (Fv.reshape(-1, 1, 1) + Fv.reshape(-1, 1) + Fv)
        return t3 / denom
    return t3


def t3d_ijk(o, v, i, j, k, t1, t2, Woovv, F, contract, WithDenom=True):
    """Disconnected triples for one (i,j,k) (pycc/cctriples.py:68-83)."""
    F = torch.as_tensor(F).to(t2.device)
    t3 = contract('ab,c->abc', Woovv[i, j], t1[k])
    t3 += contract('ac,b->abc', Woovv[i, k], t1[j])
    t3 += contract('bc,a->abc', Woovv[j, k], t1[i])
    t3 += contract('ab,c->abc', t2[i, j], F[k, v])
    t3 += contract('ac,b->abc', t2[i, k], F[j, v])
    t3 += contract('bc,a->abc', t2[j, k], F[i, v])
    if WithDenom is True:
        Fd = torch.diagonal(F)
        Fv = Fd[v]
        denom = Fd[i] + Fd[j] + Fd[k] - (Fv.reshape(-1, 1, 1) + Fv.reshape(-1, 1) + Fv)
        return t3 / denom
    return t3
