"""
ORACLE (test infrastructure, not product code): CPU/NumPy restatement of the reference's HBAR
builders (pycc/cchbar.py:90-375, CCSD branch) and Lambda residuals (pycc/cclambda.py:450-543,
CCSD branch), working from the six stored ERI blocks (see ``ccsd_oracle.Blocks``).

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import this package.  The engine (``pycc_b200``) never does.

Parity status: PINNED.  ``tests/test_oracle_golden.py`` checks every function here against
``tests/golden/lambda/*.npz``, produced by the reference's own, unmodified ``pycc/cchbar.py`` /
``pycc/cclambda.py`` on identical synthetic inputs (``tests/golden/make_golden_lambda.py``).

Further ERI slices used here, by the 8-fold symmetry of real integrals:
    ERI[v,o,v,v][a,m,e,f] = ovvv[m,a,f,e]     ERI[v,o,v,o][a,m,e,i] = ovov[m,a,i,e]
    ERI[v,o,o,v][b,m,i,e] = oovv[m,i,e,b]
"""
import numpy as np

from .ccsd_oracle import Blocks, tau, denominators, _ein


def vovv(blk):       # <am|ef>
    return blk.ovvv.transpose(1, 0, 3, 2)


def L_vovv(blk):     # 2<am|ef> - <am|fe>
    x = vovv(blk)
    return 2.0 * x - x.transpose(0, 1, 3, 2)


def vovo(blk):       # <am|ei>
    return blk.ovov.transpose(1, 0, 3, 2)


def voov(blk):       # <bm|ie>
    return blk.oovv.transpose(3, 0, 1, 2)


def build_hbar(F, blk, t1, t2):
    """All eleven HBAR blocks in the reference's index order (cchbar.py:73-83)."""
    no = blk.no
    Foo, Fov, Fvv = F[:no, :no], F[:no, no:], F[no:, no:]
    Loovv, Looov = blk.L_oovv(), blk.L_ooov()
    tt = tau(t1, t2)
    H = {}
    # build_Hov (:90-103)
    H["Hov"] = Fov + _ein('nf,mnef->me', t1, Loovv)
    # build_Hvv (:106-123)
    H["Hvv"] = Fvv - _ein('me,ma->ae', Fov, t1) + _ein('mf,amef->ae', t1, L_vovv(blk)) \
        - _ein('mnfa,mnfe->ae', tt, Loovv)
    # build_Hoo (:126-143)
    H["Hoo"] = Foo + _ein('ie,me->mi', t1, Fov) + _ein('ne,mnie->mi', t1, Looov) + _ein('inef,mnef->mi', tt, Loovv)
    # build_Hoooo (:146-167)
    tmp = _ein('je,mnie->mnij', t1, blk.ooov)
    H["Hoooo"] = blk.oooo + tmp + tmp.transpose(1, 0, 3, 2) + _ein('ijef,mnef->mnij', tt, blk.oovv)
    # build_Hvvvv (:170-191)
    tmp = _ein('mb,amef->abef', t1, vovv(blk))
    H["Hvvvv"] = blk.vvvv - tmp - tmp.transpose(1, 0, 3, 2) + _ein('mnab,mnef->abef', tt, blk.oovv)
    # build_Hvovv (:194-208), build_Hooov (:211-225)
    H["Hvovv"] = vovv(blk) - _ein('na,nmef->amef', t1, blk.oovv)
    H["Hooov"] = blk.ooov + _ein('if,nmef->mnie', t1, blk.oovv)
    # build_Hovvo (:228-249)
    H["Hovvo"] = blk.ovvo() + _ein('jf,mbef->mbej', t1, blk.ovvv) - _ein('nb,mnej->mbej', t1, blk.oovo()) \
        - _ein('jnfb,mnef->mbej', tt, blk.oovv) + _ein('njfb,mnef->mbej', t2, Loovv)
    # build_Hovov (:252-269)
    H["Hovov"] = blk.ovov + _ein('jf,bmef->mbje', t1, vovv(blk)) - _ein('nb,mnje->mbje', t1, blk.ooov) \
        - _ein('jnfb,nmef->mbje', tt, blk.oovv)
    # build_Hvvvo (:272-321)
    X = blk.vvvo() - _ein('me,miab->abei', H["Hov"], t2) + _ein('if,abef->abei', t1, H["Hvvvv"]) \
        + _ein('mnab,mnei->abei', tt, blk.oovo()) - _ein('imfa,bmfe->abei', t2, vovv(blk)) \
        - _ein('imfb,amef->abei', t2, vovv(blk)) + _ein('mifb,amef->abei', t2, L_vovv(blk))
    tmp = vovo(blk) - _ein('infa,mnfe->amei', t2, blk.oovv)
    X = X - _ein('mb,amei->abei', t1, tmp)
    tmp = voov(blk) - _ein('infb,mnef->bmie', t2, blk.oovv) + _ein('nifb,mnef->bmie', t2, Loovv)
    H["Hvvvo"] = X - _ein('ma,bmie->abei', t1, tmp)
    # build_Hovoo (:324-375)
    X = blk.ovoo() + _ein('me,ijeb->mbij', H["Hov"], t2) - _ein('nb,mnij->mbij', t1, H["Hoooo"]) \
        + _ein('ijef,mbef->mbij', tt, blk.ovvv) - _ein('ineb,nmje->mbij', t2, blk.ooov) \
        - _ein('jneb,mnie->mbij', t2, blk.ooov) + _ein('njeb,mnie->mbij', t2, Looov)
    tmp = blk.ovov - _ein('infb,mnfe->mbie', t2, blk.oovv)
    X = X + _ein('je,mbie->mbij', t1, tmp)
    tmp = voov(blk) - _ein('jnfb,mnef->bmje', t2, blk.oovv) + _ein('njfb,mnef->bmje', t2, Loovv)
    H["Hovoo"] = X + _ein('ie,bmje->mbij', t1, tmp)
    return H


def build_Goo(t2, l2):
    """cclambda.py:450-452."""
    return _ein('mjab,ijab->mi', t2, l2)


def build_Gvv(t2, l2):
    """cclambda.py:455-457."""
    return -1.0 * _ein('ijeb,ijab->ae', t2, l2)


def r_L1(H, l1, l2, Gvv, Goo):
    """cclambda.py:460-493 (CCSD branch)."""
    r = 2.0 * H["Hov"] + _ein('ie,ea->ia', l1, H["Hvv"]) - _ein('ma,im->ia', l1, H["Hoo"]) \
        + _ein('imef,efam->ia', l2, H["Hvvvo"]) - _ein('mnae,iemn->ia', l2, H["Hovoo"]) \
        + _ein('me,ieam->ia', l1, 2.0 * H["Hovvo"] - H["Hovov"].swapaxes(2, 3))
    r = r - 2.0 * _ein('ef,eifa->ia', Gvv, H["Hvovv"]) + _ein('ef,eiaf->ia', Gvv, H["Hvovv"]) \
        - 2.0 * _ein('mn,mina->ia', Goo, H["Hooov"]) + _ein('mn,imna->ia', Goo, H["Hooov"])
    return r


def r_L2(blk, H, l1, l2, Gvv, Goo):
    """cclambda.py:495-543 (CCSD branch), including the final r + r^T."""
    Loovv = blk.L_oovv()
    r = Loovv + 2.0 * _ein('ia,jb->ijab', l1, H["Hov"]) - _ein('ja,ib->ijab', l1, H["Hov"]) \
        + 2.0 * _ein('ie,ejab->ijab', l1, H["Hvovv"]) - _ein('ie,ejba->ijab', l1, H["Hvovv"]) \
        - 2.0 * _ein('mb,jima->ijab', l1, H["Hooov"]) + _ein('mb,ijma->ijab', l1, H["Hooov"])
    r = r + _ein('ijeb,ea->ijab', l2, H["Hvv"]) - _ein('mjab,im->ijab', l2, H["Hoo"]) \
        + 0.5 * _ein('mnab,ijmn->ijab', l2, H["Hoooo"]) + 0.5 * _ein('ijef,efab->ijab', l2, H["Hvvvv"]) \
        + _ein('mjeb,ieam->ijab', l2, 2.0 * H["Hovvo"] - H["Hovov"].swapaxes(2, 3)) \
        - _ein('mibe,jema->ijab', l2, H["Hovov"]) - _ein('mieb,jeam->ijab', l2, H["Hovvo"]) \
        + _ein('ae,ijeb->ijab', Gvv, Loovv) - _ein('mi,mjab->ijab', Goo, Loovv)
    return r + r.transpose(1, 0, 3, 2)


def lambda_residuals(blk, H, t2, l1, l2):
    """The loop body of solve_lambda (cclambda.py:146-151)."""
    Goo, Gvv = build_Goo(t2, l2), build_Gvv(t2, l2)
    return r_L1(H, l1, l2, Gvv, Goo), r_L2(blk, H, l1, l2, Gvv, Goo)


def pseudoenergy(blk, l2):
    """cclambda.py:579-581."""
    return 0.5 * float(_ein('ijab,ijab->', blk.oovv, l2))


def lambda_guess(t1, t2):
    """cclambda.py:57-58."""
    return 2.0 * t1, 2.0 * (2.0 * t2 - t2.swapaxes(2, 3))


def lambda_iteration(F, blk, H, t2, l1, l2):
    """One pass of the solve_lambda loop without DIIS (cclambda.py:141-151, 244-254):
    returns (l1_new, l2_new, pseudoenergy, rms)."""
    Dia, Dijab = denominators(F, blk.no)
    r1, r2 = lambda_residuals(blk, H, t2, l1, l2)
    i1, i2 = r1 / Dia, r2 / Dijab
    l1n, l2n = l1 + i1, l2 + i2
    rms = float(np.sqrt(np.sum(i1 * i1) + np.sum(i2 * i2)))
    return l1n, l2n, pseudoenergy(blk, l2n), rms
