"""
HBAR and Lambda engines: the one- and two-body blocks of the
similarity-transformed Hamiltonian (pycc/cchbar.py:90-375) and the Lambda residuals
(pycc/cclambda.py:450-543), planned once into ``pccb_contract`` launches over the six stored ERI
blocks exactly like the CCSD builders of ``engine.py``.

The reference slices ERI / L into sixteen differently ordered blocks; here every one of them is an
index relabelling of the six stored blocks by the 8-fold symmetry of real integrals
(<pq|rs> = <qp|sr> = <rs|pq> = <rq|ps>):

    <am|ef> = ovvv[mafe]   <mn|ej> = ooov[nmje]   <mb|ej> = oovv[mjeb]   <ab|ei> = ovvv[ieba]
    <am|ei> = ovov[maie]   <bm|ie> = oovv[mieb]   <mb|ij> = ooov[ijmb]
    L<am|ef> = 2 ovvv[mafe] - ovvv[maef]          L<mn|ie> = Looov[mnie]   L<mn|ef> = Loovv[mnef]

HBAR blocks are held in the reference's index order, except the two whose only Lambda consumer
contracts over their leading indices: Hvvvv is stored transposed, HvT[a,b,e,f] = Hvvvv[e,f,a,b], so
that the o^2 v^4 ladder ``l2[ijef] Hvvvv[efab]`` (cclambda.py:528) reads it with the contracted
pair contiguous, like the CCSD ladder reads <ab|ef>; and Hvvvo as Hs[e,i,a,b] = Hvvvo[a,b,e,i], so
that ``l2[imef] Hvvvo[efam]`` (:473) streams it.  ``Hvvvv`` / ``Hvvvo`` are exposed as permuted views.

CCD is the CCSD algebra with t1 = l1 = 0 (every reference CCD branch is the CCSD branch with its
t1 terms dropped: cchbar.py:92-96, 108-113, ..., cclambda.py:462-466, 497-512).  CC2 has its own
statements (cchbar.py:163, 186, 241, 263, 285-294, 337-346; cclambda.py:481-486, 522-524).

Multi-GPU (one process per GPU): the two big blocks are sharded by rows over the virtual slice of
the CCSD engine - HvT[A,b,e,f] and Hs[E,i,a,b] hold a, e in [vlo, vhi) only - everything else is
replicated; an upper-case A / B / E in a statement is the sliced index and ``X_sK`` the matching
view of X (as in ``engine.STATEMENTS``).  Every Lambda term then produces the rows a (or columns
b) of r1 / r2 that belong to the slice, and one allreduce per residual sums the partial results -
the same exchange step as the CCSD iteration.  CCSD, CCD and CC2.
"""
import re

import torch

from .dsl import TRef, plan


def resolve(T, name, eng):
    """Tensor name of a statement -> strided view.  ``X_sK`` is X restricted to this rank's slice of
    the virtual index along dimension K (the whole tensor on one GPU); the two blocks stored by
    rows, <ab|ef> and the local HvT / Hs / ovA / ovB, are addressed by their plain names."""
    if name == "vvvv_s0":
        return TRef.of(eng.vvvv_loc, name)
    if name in T:
        return TRef.of(T[name], name)
    base, dim = name[:-3], name[-1]
    if name[-3:-1] != "_s" or base not in T or not dim.isdigit():
        raise KeyError("unknown tensor %r" % name)
    return TRef.of(T[base].narrow(int(dim), eng.vlo, eng.vl), name)


class HbarEngine(object):
    ORDER = ("Hov", "Hvv", "Hoo", "Hoooo", "Hvvvv", "Hvovv", "Hooov", "Hovvo", "Hovov", "X", "Hvvvo", "Hovoo")

    STATEMENTS = {
        # build_Hov (cchbar.py:90-103)
        "Hov": ["Hov[me] = Fov[me] + t1[nf]*Loovv[mnef]"],
        # build_Hvv (:106-123)
        "Hvv": ["Hvv[ae] = Fvv[ae] - t1[ma]*Fov[me] - tau[mnfa]*Loovv[mnfe]",
                "Hvv[ae] += 2.0*t1[mf]*ovvv[mafe] - t1[mf]*ovvv[maef]"],
        # build_Hoo (:126-143)
        "Hoo": ["Hoo[mi] = Foo[mi] + Fov[me]*t1[ie] + Loovv[mnef]*tau[inef]",
                "Hoo[mi] += t1[ne]*Looov[mnie]"],
        # build_Hoooo (:146-167)
        "Hoooo": ["Hoooo[mnij] = oooo[mnij] + oovv[mnef]*tau[ijef]",
                  "Hoooo[mnij] += ooov[mnie]*t1[je]",
                  "Hoooo[mnij] += ooov[nmje]*t1[ie]"],
        # build_Hvvvv (:170-191), stored as HvT[abef] = Hvvvv[efab]; <ef|ab> = <ab|ef>
        # the two t1 terms (K = o only, bound by the v^4 read-modify-write) read <me|ba> and <mf|ab>
        # from the transposed copies ovA[mabe] = ovvv[meba], ovB[mabf] = ovvv[mfab] so that both
        # stream HvT and the integrals contiguously
        "Hvvvv": ["HvT[Abef] = vvvv_s0[Abef] + oovv_s2[mnAb]*tau[mnef]",
                  "HvT[Abef] -= ovA[mAbe]*t1[mf]",
                  "HvT[Abef] -= ovB[mAbf]*t1[me]"],
        # build_Hvovv (:194-208)
        "Hvovv": ["Hvovv[amef] = ovvv[mafe] - t1[na]*oovv[nmef]"],
        # build_Hooov (:211-225)
        "Hooov": ["Hooov[mnie] = ooov[mnie] + oovv[nmef]*t1[if]"],
        # build_Hovvo (:228-249)
        "Hovvo": ["Hovvo[mbej] = oovv[mjeb] - oovv[mnef]*tau[jnfb] + Loovv[mnef]*t2[njfb]",
                  "Hovvo[mbej] += ovvv[mbef]*t1[jf]",
                  "Hovvo[mbej] -= t1[nb]*ooov[nmje]"],
        # build_Hovov (:252-269)
        "Hovov": ["Hovov[mbje] = ovov[mbje] - oovv[nmef]*tau[jnfb]",
                  "Hovov[mbje] += ovvv[mbfe]*t1[jf]",
                  "Hovov[mbje] -= t1[nb]*ooov[mnje]"],
        # the two o^2 v^2 temporaries of build_Hvvvo (:303-316), which build_Hovoo forms again (:358-370)
        "X": ["Xa[mbie] = ovov[mbie] - oovv[mnfe]*t2[infb]",
              "Xb[bmje] = oovv[mjeb] - oovv[mnef]*t2[jnfb] + Loovv[mnef]*t2[njfb]"],
        # build_Hvvvo (:272-321), stored as Hs[eiab] = Hvvvo[abei]; t2x[mifb] = 2 t2[mifb] - t2[imfb] merges
        # the second and third o^2 v^4 terms (:300-301) into one K-concatenated launch with the fourth
        "Hvvvo": ["Hs[Eiab] = ovvv_s1[iEba] - t2[miab]*Hov_s1[mE]",
                  "Hs[Eiab] += HvT[Efab]*t1[if]",
                  "Hs[Eiab] += tau[mnab]*ooov_s3[nmiE]",
                  "Hs[Eiab] -= ovvv_s2[mbEf]*t2[imfa]",
                  "Hs[Eiab] += ovvv_s3[mafE]*t2x[mifb] - ovvv_s2[maEf]*t2[mifb]",
                  "Hs[Eiab] -= t1[mb]*Xa_s3[maiE]",
                  "Hs[Eiab] -= t1[ma]*Xb_s3[bmiE]"],
        # build_Hovoo (:324-375)
        "Hovoo": ["Hovoo[mbij] = ooov[ijmb] + Hov[me]*t2[ijeb]",
                  "Hovoo[mbij] -= t1[nb]*Hoooo[mnij]",
                  "Hovoo[mbij] += ovvv[mbef]*tau[ijef]",
                  "Hovoo[mbij] -= t2[ineb]*ooov[nmje]",
                  "Hovoo[mbij] -= t2[jneb]*ooov[mnie]",
                  "Hovoo[mbij] += t2[njeb]*Looov[mnie]",
                  "Hovoo[mbij] += Xa[mbie]*t1[je]",
                  "Hovoo[mbij] += Xb[bmje]*t1[ie]"],
    }

    # CC2 branches; T11[ijab] = t1[ia] t1[jb].  The reference's CC2 code reads ``self.ccwfn.H.F`` here
    # (cchbar.py:289, 341) - the object's own Fock matrix F0, not the F handed to
    # cclambda.residuals - and so does this.
    STATEMENTS_CC2 = {
        "X": [],
        "Hoooo": ["Hoooo[mnij] = oooo[mnij] + oovv[mnef]*T11[ijef]",
                  "Hoooo[mnij] += ooov[mnie]*t1[je]",
                  "Hoooo[mnij] += ooov[nmje]*t1[ie]"],
        "Hvvvv": ["HvT[Abef] = vvvv_s0[Abef] + oovv_s2[mnAb]*T11[mnef]",
                  "HvT[Abef] -= ovA[mAbe]*t1[mf]",
                  "HvT[Abef] -= ovB[mAbf]*t1[me]"],
        "Hovvo": ["Hovvo[mbej] = oovv[mjeb] + ovvv[mbef]*t1[jf]",
                  "Hovvo[mbej] -= t1[nb]*ooov[nmje]"],
        "Hovov": ["Hovov[mbje] = ovov[mbje] + ovvv[mbfe]*t1[jf]",
                  "Hovov[mbje] -= t1[nb]*ooov[mnje]"],
        "Hvvvo": ["Hs[Eiab] = ovvv_s1[iEba] - t2[miab]*F0ov_s1[mE]",
                  "Hs[Eiab] += HvT[Efab]*t1[if]",
                  "Hs[Eiab] += T11[mnab]*ooov_s3[nmiE]",
                  "Hs[Eiab] -= t1[mb]*ovov_s3[maiE]",
                  "Hs[Eiab] -= t1[ma]*oovv_s2[miEb]"],
        "Hovoo": ["Hovoo[mbij] = ooov[ijmb] + F0ov[me]*t2[ijeb]",
                  "Hovoo[mbij] -= t1[nb]*Hoooo[mnij]",
                  "Hovoo[mbij] += ovvv[mbef]*T11[ijef]",
                  "Hovoo[mbij] += ovov[mbie]*t1[je]",
                  "Hovoo[mbij] += oovv[mjeb]*t1[ie]"],
    }

    def __init__(self, eng):
        if eng.ovvv_full is None:
            raise ValueError("HBAR needs the full <mb|ef> block on every rank (ccwfn(..., full_ovvv=True))")
        if eng.model not in ('CCD', 'CC2', 'CCSD', 'CC3'):      # CC3 uses the CCSD HBAR (pycc/cchbar.py has no CC3 branch)
            raise ValueError("HBAR / Lambda: models CCD, CC2, CCSD, CCSD(T), CC3 (got %r)" % (eng.model,))
        self.eng, self.be = eng, eng.be
        o, v, vl, dt = eng.no, eng.nv, eng.vl, eng.dtype
        E = lambda *s: self.be.zeros(s, dt)
        # the v^4 and o v^3 blocks hold this rank's rows only (HvT[A,b,e,f], Hs[E,i,a,b]); the rest is replicated
        self.Hov, self.Hvv, self.Hoo = E(o, v), E(v, v), E(o, o)
        self.Hoooo, self.HvT = E(o, o, o, o), E(vl, v, v, v)
        self.Hvovv, self.Hooov = E(v, o, v, v), E(o, o, o, v)
        self.Hovvo, self.Hovov = E(o, v, v, o), E(o, v, o, v)
        self.Hs, self.Hovoo = E(vl, o, v, v), E(o, v, o, o)
        self.Xa, self.Xb = E(o, v, o, v), E(v, o, o, v)
        self.t2x, self.ovA, self.ovB = E(o, o, v, v), E(o, vl, v, v), E(o, vl, v, v)
        if eng.model == 'CC2':
            self.T11 = eng.T11
            self.t1_0, self.t2_0, self.tau_0 = E(o, v), E(o, o, v, v), E(o, o, v, v)
        st = dict(self.STATEMENTS)
        if eng.model == 'CC2':
            st.update(self.STATEMENTS_CC2)
        self._ov_ready = False
        T = self.tensors()
        self.plans = {}
        for name in self.ORDER:
            self.plans[name] = []
            for s in st[name]:
                names = set(re.findall(r"([A-Za-z_][A-Za-z_0-9]*)\[", s))
                self.plans[name] += plan(s, {k: resolve(T, k, eng) for k in names}, eng.dcode)

    def tensors(self):
        eng = self.eng
        o = eng.no
        b, F = eng.blk, eng.F
        d = dict(oooo=b["oooo"], ooov=b["ooov"], oovv=b["oovv"], ovov=b["ovov"], ovvv=eng.ovvv_full, vvvv=eng.vvvv_loc,
                 Foo=F[:o, :o], Fov=F[:o, o:], Fvv=F[o:, o:], Loovv=eng.Loovv, Looov=eng.Looov,
                 t1=eng.t1, t2=eng.t2, tau=eng.tau,
                 Hov=self.Hov, Hvv=self.Hvv, Hoo=self.Hoo, Hoooo=self.Hoooo, HvT=self.HvT, Hvovv=self.Hvovv,
                 Hooov=self.Hooov, Hovvo=self.Hovvo, Hovov=self.Hovov, Hs=self.Hs, Hovoo=self.Hovoo,
                 Xa=self.Xa, Xb=self.Xb, t2x=self.t2x, ovA=self.ovA, ovB=self.ovB)
        if eng.model == 'CC2':
            F0 = eng.F0
            d.update(T11=eng.T11, F0oo=F0[:o, :o], F0ov=F0[:o, o:], F0vv=F0[o:, o:],
                     t1_0=self.t1_0, t2_0=self.t2_0, tau_0=self.tau_0)
        return d

    def snapshot(self):
        """CC2: remember the object's own amplitudes (the engine's current ones) for the terms of
        the Lambda residual that the reference evaluates with ``self.ccwfn.t1 / .t2``."""
        if self.eng.model == 'CC2':
            eng = self.eng
            self.be.bind_stream()
            eng.prepare()
            self.t1_0.copy_(eng.t1)
            self.t2_0.copy_(eng.t2)
            self.tau_0.copy_(eng.tau)

    @property
    def Hvvvv(self):
        return self.HvT.permute(2, 3, 0, 1)

    @property
    def Hvvvo(self):
        return self.Hs.permute(2, 3, 0, 1)

    DEPS = dict(Hvvvo=("Hov", "Hvvvv", "X"), Hovoo=("Hov", "Hoooo", "X"))

    def build(self, only=None):
        """All blocks, in the reference's order (cchbar.py:73-83), or just the builder ``only`` and
        what it depends on, for the amplitudes and Fock matrix currently held by the CCSD engine."""
        eng = self.eng
        self.be.bind_stream()
        if eng.model == 'CCD':
            eng.t1.zero_()
        eng.prepare()
        if only in (None, "Hvvvo"):
            self.be.permute4(self.t2x, eng.t2, 2.0, 0.0)
            self.be.permute4(self.t2x, eng.t2.permute(1, 0, 2, 3), -1.0, 1.0)
        if not self._ov_ready:       # integrals only: formed once
            self.be.permute4(self.ovA, eng.ovvv_full.permute(0, 3, 2, 1).narrow(1, eng.vlo, eng.vl))
            self.be.permute4(self.ovB, eng.ovvv_full.permute(0, 2, 3, 1).narrow(1, eng.vlo, eng.vl))
            self._ov_ready = True
        todo = self.ORDER if only is None else [n for n in self.ORDER if n == only or n in self.DEPS.get(only, ())]
        for name in todo:
            for L in self.plans[name]:
                self.be.contract(L.desc)

    def flops_issued(self):
        tot = 0
        for ls in self.plans.values():
            for L in ls:
                d = L.desc
                tot += sum(2 * d.M * d.N * d.batch * d.seg[q].K for q in range(d.nseg))
        return tot


class LambdaEngine(object):
    """Lambda residuals r_L1 / r_L2 (pycc/cclambda.py:450-543) for the HBAR held by ``hbar``."""

    # r1 / r2 are zeroed first and every term accumulates into this rank's rows (A) or columns (B) of
    # the result, so that the per-rank partial residuals add up under the allreduce.
    STATEMENTS = {
        # build_Goo / build_Gvv (cclambda.py:450-457), replicated
        "G": ["Goo[mi] = t2[mjab]*l2[ijab]",
              "Gvv[ae] = -1.0*t2[ijeb]*l2[ijab]"],
        # r_L1 (:460-493); Hvvvo[efam] = Hs[amef]
        "r_L1": ["r1_s1[iA] += 2.0*Hov_s1[iA] + l1[ie]*Hvv_s1[eA] - l1_s1[mA]*Hoo[im] + l2[imef]*Hs[Amef]"
                 " - l2_s2[mnAe]*Hovoo[iemn]",
                 "r1_s1[iA] += 2.0*l1[me]*Hovvo_s2[ieAm] - l1[me]*Hovov_s3[iemA]",
                 "r1_s1[iA] += Gvv[ef]*Hvovv_s2[eiAf] - 2.0*Gvv[ef]*Hvovv_s3[eifA]",
                 "r1_s1[iA] += Goo[mn]*Hooov_s3[imnA] - 2.0*Goo[mn]*Hooov_s3[minA]"],
        # r_L2 (:495-543), before the final r + r^T; Hvvvv[efab] = HvT[abef]
        "r_L2": ["r2_s2[ijAb] += Loovv_s2[ijAb] + 2.0*l1_s1[iA]*Hov[jb] - l1_s1[jA]*Hov[ib]",
                 "r2_s2[ijAb] += 2.0*l1[ie]*Hvovv_s2[ejAb] - l1[ie]*Hvovv_s3[ejbA]",
                 "r2_s3[ijaB] += l1_s1[mB]*Hooov[ijma] - 2.0*l1_s1[mB]*Hooov[jima]",
                 "r2_s2[ijAb] += l2[ijeb]*Hvv_s1[eA] - l2_s2[mjAb]*Hoo[im]",
                 "r2_s2[ijAb] += 0.5*l2_s2[mnAb]*Hoooo[ijmn]",
                 "r2_s2[ijAb] += 0.5*l2[ijef]*HvT[Abef]",
                 "r2_s2[ijAb] += 2.0*l2[mjeb]*Hovvo_s2[ieAm] - l2[mjeb]*Hovov_s3[iemA]",
                 "r2_s2[ijAb] -= l2[mibe]*Hovov_s3[jemA]",
                 "r2_s2[ijAb] -= l2[mieb]*Hovvo_s2[jeAm]",
                 "r2_s2[ijAb] += Gvv_s0[Ae]*Loovv[ijeb] - Goo[mi]*Loovv_s2[mjAb]"],
    }

    # CC2 (:481-486, 522-524): Goo / Gvv are built but not used; Fa / Fi are the t1-dressed Fock blocks.
    # As in the reference these terms read the OBJECT's Fock matrix and amplitudes (self.ccwfn.H.F,
    # self.ccwfn.t1 / .t2): F0.., t1_0, t2_0, tau_0 - identical to the arguments inside solve_lambda,
    # different when cclambda.residuals is called with other amplitudes (real-time propagation).
    # Sliced like STATEMENTS (upper-case A / B = this rank's part of the virtual index); the small intermediates
    # Ya / Yb / Fa / Fi are formed in full on every rank.
    STATEMENTS_CC2 = {
        "r_L1": ["r1_s1[iA] += 2.0*Hov_s1[iA] + l1[ie]*Hvv_s1[eA] - l1_s1[mA]*Hoo[im] + l2[imef]*Hs[Amef]"
                 " - l2_s2[mnAe]*Hovoo[iemn]",
                 "r1_s1[iA] += 2.0*l1[me]*Hovvo_s2[ieAm] - l1[me]*Hovov_s3[iemA]",
                 "Ya[nf] = l1[me]*t2_0[nmfe]",
                 "Yb[nf] = l1[me]*tau_0[mnfe]",
                 "r1_s1[iA] += 2.0*Ya[nf]*Loovv_s2[inAf] - 2.0*Yb[nf]*oovv_s2[inAf] + Yb[nf]*oovv_s3[infA]"],
        "r_L2": ["r2_s2[ijAb] += Loovv_s2[ijAb] + 2.0*l1_s1[iA]*Hov[jb] - l1_s1[jA]*Hov[ib]",
                 "r2_s2[ijAb] += 2.0*l1[ie]*Hvovv_s2[ejAb] - l1[ie]*Hvovv_s3[ejbA]",
                 "r2_s3[ijaB] += l1_s1[mB]*Hooov[ijma] - 2.0*l1_s1[mB]*Hooov[jima]",
                 "Fa[ea] = F0vv[ea] - F0ov[ma]*t1_0[me]",      # (F_vv - F_ov^T t1)[e,a], index order as the reference uses it
                 "Fi[im] = F0oo[im] + t1_0[me]*F0ov[ie]",
                 "r2_s2[ijAb] += l2[ijeb]*Fa_s1[eA] - l2_s2[mjAb]*Fi[im]"],
    }

    def __init__(self, hbar):
        self.hbar, self.eng, self.be = hbar, hbar.eng, hbar.be
        eng = self.eng
        o, v, dt = eng.no, eng.nv, eng.dtype
        E = lambda *s: self.be.zeros(s, dt)
        self.l1, self.l2 = E(o, v), E(o, o, v, v)
        self.r1, self.r2 = E(o, v), E(o, o, v, v)
        self.Goo, self.Gvv = E(o, o), E(v, v)
        self.Ya, self.Yb, self.Fa, self.Fi = E(o, v), E(o, v), E(v, v), E(o, o)
        st = dict(self.STATEMENTS)
        if eng.model == 'CC2':
            st.update(self.STATEMENTS_CC2)
        T = hbar.tensors()
        for k in ("l1", "l2", "r1", "r2", "Goo", "Gvv", "Ya", "Yb", "Fa", "Fi"):
            T[k] = getattr(self, k)
        self.plans = {}
        for name, stmts in st.items():
            self.plans[name] = []
            for s in stmts:
                names = set(re.findall(r"([A-Za-z_][A-Za-z_0-9]*)\[", s))
                for L in plan(s, {k: resolve(T, k, eng) for k in names}, eng.dcode):
                    L.is_ladder = "HvT[" in s
                    self.plans[name].append(L)
        self.ladder_events = None
        self.t3_source = None        # (S1, S2) for model CCSD(T), set by cclambda
        self.cc3 = None              # cc3_lambda_engine.CC3LambdaEngine for model CC3, set by cclambda

    def guess(self):
        """l1 = 2 t1, l2 = 2 (2 t2 - t2^T) (cclambda.py:57-58)."""
        eng = self.eng
        self.be.bind_stream()
        self.be.axpby(self.l1, 2.0, eng.t1)
        self.be.pair_combine(self.l2, eng.t2, 4.0, -2.0, None, None, 0.0, eng.no, eng.nv)

    def _launch(self, name):
        for L in self.plans[name]:
            if self.ladder_events is not None and L.is_ladder:
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                self.be.contract(L.desc)
                e1.record()
                self.ladder_events.append((e0, e1))
            else:
                self.be.contract(L.desc)

    def run(self, name):
        """One of 'G', 'r_L1', 'r_L2' (the latter two read Goo / Gvv)."""
        self.be.bind_stream()
        if self.eng.model == 'CCD':
            self.l1.zero_()
        if name == "r_L1":
            self.r1.zero_()
            if self.eng.model == 'CCD':
                return               # cclambda.py:462-466
        elif name == "r_L2":
            self.r2.zero_()
        from .parallel import nvtx_range
        with nvtx_range("lambda/" + name):
            self._launch(name)
        # the one exchange step: sum the per-rank partial residuals
        if name == "r_L1":
            self.eng.comm.allreduce_(self.r1)
        elif name == "r_L2":
            self.eng.comm.allreduce_(self.r2)
        # CCSD(T): the (T) source terms of ccwfn.t3_density (cclambda.py:474-475, 519-520), added before the pair
        # symmetrisation of r2 exactly as the reference does (S2 is symmetric: 1/2 S2 + 1/2 S2^T = S2)
        src = getattr(self, "t3_source", None)
        if src is not None:
            if name == "r_L1":
                self.be.axpby(self.r1, 1.0, self.r1, 1.0, src[0])
            elif name == "r_L2":
                self.be.axpby(self.r2, 1.0, self.r2, 0.5, src[1])

    def residuals(self, symmetrize=True):
        """r1, r2 for the amplitudes in self.l1 / self.l2 (the loop body of cclambda.py:146-151)."""
        for name in ("G", "r_L1", "r_L2"):
            self.run(name)
        if getattr(self, "cc3", None) is not None:
            self.cc3.contribute()        # Y1, Y2 of the CC3 triples (cclambda.py:160-235), before r2 + r2^T
        if symmetrize:
            self.be.pair_symmetrize(self.r2, self.eng.no, self.eng.nv)     # :542
        return self.r1, self.r2

    def pseudoenergy(self):
        """1/2 <ij|ab> l2[ijab] (cclambda.py:579-581)."""
        self.be.bind_stream()
        return 0.5 * self.be.dot(self.eng.blk["oovv"], self.l2)

    def update(self, symmetrize_on_the_fly):
        """l += r/D; returns the squared increment norm (cclambda.py:244-247)."""
        eng = self.eng
        self.be.bind_stream()
        return self.be.update_amplitudes(self.l1, self.l2, self.r1, self.r2, eng.eps, eng.no, eng.nv,
                                         symmetrize_on_the_fly)

    def flops_issued(self):
        tot = 0
        for ls in self.plans.values():
            for L in ls:
                d = L.desc
                tot += sum(2 * d.M * d.N * d.batch * d.seg[q].K for q in range(d.nseg))
        return tot
