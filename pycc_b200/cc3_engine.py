"""
CC3 contributions to the singles and doubles residuals: the
reference's per-iteration loop over all occupied triplets (pycc/ccwfn.py:371-404) with T1-dressed
integrals (``build_cc3_*``, :625-710), on the engine's kernels:

  * the five dressed intermediates are planned statements over the stored ERI blocks; they are kept
    in the layouts the (T) machinery consumes, ``Wv[i][a][b][e] = Wabei[a,b,e,i]`` and
    ``Wo[i][j][m][b] = Wmbij[m,b,i,j]``, so the connected triples of a whole k-batch come from the
    same three batched ``pccb_contract`` launches per occupied pair (i, j) as in (T);
  * ``pccb_t3_combine`` divides by the triples denominator and forms ``t3 - t3^(ac)`` and
    ``2 t3 - t3^(bc) - t3^(ac)`` for the batch;
  * four contractions per pair, with k folded into the contracted index, add X1 / X2 (:395-398)
    straight into the unsymmetrised r1 / r2 (the final r2 + r2^T supplies X2^T, :401).

No o^3 v^3 tensor exists at any time (three o v^3 buffers).  Several GPUs: the occupied pairs are dealt over the
ranks, the v^4 term of the dressed <ab|ei> is formed from each rank's rows of <ab|ef> (one allreduce), and the
partial X1 / X2 are summed by the iteration's allreduce of r1 / r2.
"""
import torch

from .dsl import TRef, plan
from .triples_engine import TriplesEngine, TripletBatch


class CC3Engine(object):
    # dressed intermediates (index letters as in the reference)
    STATEMENTS = (
        # build_cc3_Wmnij (:625-635)
        "Wmnij3[mnij] = oooo[mnij] + oovv[mnef]*T11[ijef]",
        "Wmnij3[mnij] += ooov[mnie]*t1[je]",
        "Wmnij3[mnij] += ooov[nmje]*t1[ie]",
        # build_cc3_Wmbij (:637-650) as Wo[ijmb]; <mb|ij> = <ij|mb>, <mb|ej> = <mj|eb>
        "X[mbej] = ovvv[mbef]*t1[jf]",
        "Wo[ijmb] = ooov[ijmb] - Wmnij3[mnij]*t1[nb]",
        "Wo[ijmb] += ovov[mbie]*t1[je]",
        "Wo[ijmb] += t1[ie]*X[mbej]",
        "Wo[ijmb] += t1[ie]*oovv[mjeb]",
        # build_cc3_Wmnie (:652-659)
        "Wmnie3[mnie] = ooov[mnie] + t1[if]*oovv[mnfe]",
        # build_cc3_Wamef (:661-668) as Wk[maef]; <am|ef> = <ma|fe>  (the bare part is set by a permuted copy)
        "Wk[maef] -= t1[na]*oovv[nmef]",
        # build_cc3_Wabei (:670-710) as Wv[iabe] = Wabei[a,b,e,i]; P, Q, R start as permuted copies of
        # <ei|am> = ovov[i,e,m,a], <mn|ei> = ooov[n,m,i,e], <mb|ei> = oovv[m,i,e,b]
        "P[imae] += ovvv[mafe]*t1[if]",
        "Q[imne] += oovv[mnef]*t1[if]",
        "Qa[iane] = t1[ma]*Q[imne]",
        "R[imbe] += ovvv[mbef]*t1[if]",
        "Wv[iabe] += t1[if]*vvvv[abef]",
        "Wv[iabe] -= P[imae]*t1[mb]",
        "Wv[iabe] += Qa[iane]*t1[nb]",
        "Wv[iabe] -= t1[ma]*R[imbe]",
    )

    def __init__(self, eng):
        if eng.ovvv_full is None:
            raise ValueError("model='CC3' needs the full <mb|ef> block")
        self.eng, self.be = eng, eng.be
        o, v, dt = eng.no, eng.nv, eng.dtype
        E = lambda *s: self.be.zeros(s, dt)
        self.T11 = E(o, o, v, v)
        self.Wmnij3, self.Wo, self.Wmnie3 = E(o, o, o, o), E(o, o, o, v), E(o, o, o, v)
        self.X, self.P, self.R, self.Qa = E(o, v, v, o), E(o, o, v, v), E(o, o, v, v), E(o, v, o, v)
        self.Q = E(o, o, o, v)
        self.Wk, self.Wv = E(o, v, v, v), E(o, v, v, v)
        self.A1, self.A2 = E(o, v, v, v), E(o, v, v, v)
        b = eng.blk
        T = dict(oooo=b["oooo"], ooov=b["ooov"], oovv=b["oovv"], ovov=b["ovov"], ovvv=eng.ovvv_full, vvvv=eng.vvvv_loc,
                 t1=eng.t1, T11=self.T11, Wmnij3=self.Wmnij3, Wo=self.Wo, Wmnie3=self.Wmnie3, X=self.X, P=self.P,
                 R=self.R, Q=self.Q, Qa=self.Qa, Wk=self.Wk, Wv=self.Wv)
        T = {k: TRef.of(t, k) for k, t in T.items()}
        self.launches = []
        self.sharded = eng.comm.world > 1
        for s in self.STATEMENTS:
            if self.sharded and "vvvv[" in s:
                continue
            self.launches += plan(s, T, eng.dcode)
        if self.sharded:
            # several GPUs: a rank holds the rows a in [vlo, vhi) of <ab|ef>, so it forms those rows of the one
            # v^4 term of Wabei in a zero-initialised buffer; one allreduce of the disjoint slices is the gather.
            # Everything else is o-heavy and replicated; the occupied pairs of contribute() are dealt round-robin
            # and the partial X1 / X2 ride on the iteration's allreduce of r1 / r2 (engine.residuals).
            self.Wvv = E(o, v, v, v)
            Tl = dict(WvA=self.Wvv.narrow(1, eng.vlo, eng.vl), t1=eng.t1, vvvvA=eng.vvvv_loc)
            Tl = {k: TRef.of(t, k) for k, t in Tl.items()}
            self.launch_vvvv = plan("WvA[iabe] = t1[if]*vvvvA[abef]", Tl, eng.dcode)
        self.tri = TriplesEngine(eng, V=self.Wv, O=self.Wo)
        self._pair_plans = {}
        # real-time propagation with a field-dressed (real) Fock matrix: t3 -= t3_pert_ijk (pycc/ccwfn.py:378-386,
        # cctriples.py:547-562), i.e. one more term of W before the common denominator:
        #   W[abc] -= (sum_d V[l,d] t2[i,j,a,d]) t2[k,l,c,b],  V = F - H.F
        self.pert = False
        self.Vov, self.Tp = E(o, v), E(o, o, v, o)
        Tq = dict(Tp=self.Tp, t2=eng.t2, Vov=self.Vov)
        self.pert_tmp = plan("Tp[ijal] = t2[ijad]*Vov[ld]", {k: TRef.of(t, k) for k, t in Tq.items()}, eng.dcode)
        self.pert_w = TripletBatch(o, eng.dcode, ("W[abc] -= tmp[al]*t2k[lcb]",),
                                   lambda i, j, k: dict(W=self.tri.W[k], tmp=self.Tp[i, j], t2k=eng.t2[k]), self.be)

    def build_intermediates(self):
        eng, be = self.eng, self.be
        o, v = eng.no, eng.nv
        b = eng.blk
        be.pair_combine(self.T11, eng.t2, 0.0, 0.0, eng.t1, eng.t1, 1.0, o, v)
        # bare parts that are permuted views of the stored blocks
        be.permute4(self.Wk, eng.ovvv_full.permute(0, 1, 3, 2))        # <am|ef> = ovvv[m,a,f,e]
        be.permute4(self.P, b["ovov"].permute(0, 2, 3, 1))             # <ei|am> = ovov[i,e,m,a]
        be.permute4(self.Q, b["ooov"].permute(2, 1, 0, 3))             # <mn|ei> = ooov[n,m,i,e]
        be.permute4(self.R, b["oovv"].permute(1, 0, 3, 2))             # <mb|ei> = oovv[m,i,e,b]
        be.permute4(self.Wv, eng.ovvv_full.permute(0, 3, 2, 1))        # <ei|ab> = ovvv[i,e,b,a]
        for L in self.launches:
            be.contract(L.desc)
        if self.sharded:
            self.Wvv.zero_()             # the other ranks' rows hold the previous call's gathered values
            for L in self.launch_vvvv:
                be.contract(L.desc)
            eng.comm.allreduce_(self.Wvv)
            be.axpby(self.Wv, 1.0, self.Wv, 1.0, self.Wvv)

    def my_pairs(self):
        """The occupied pairs (i, j) this rank evaluates: all o^2 of them dealt round-robin (equal work each)."""
        o, c = self.eng.no, self.eng.comm
        return [(i, j) for i in range(o) for j in range(o)][c.rank::c.world]

    def _plans_for(self, i, j):
        key = (i, j)
        if key not in self._pair_plans:
            eng = self.eng
            T = dict(A1=self.A1, A2=self.A2, Lj=eng.Loovv[j], Fme=eng.Fme, Wk=self.Wk, Wj=self.Wmnie3[j],
                     r1i=eng.r1[i], r2ij=eng.r2[i, j], r2i=eng.r2[i])
            T = {k: TRef.of(t, k) for k, t in T.items()}
            ls = []
            for s in ("r1i[a] += A1[kabc]*Lj[kbc]",                    # X1 (:395)
                      "r2ij[ab] += A1[kabc]*Fme[kc]",                  # X2 (:396)
                      "r2ij[ad] += A2[kabc]*Wk[kdbc]",                 # X2 (:397)
                      "r2i[lab] -= A2[kabc]*Wj[klc]"):                 # X2 (:398)
                ls += plan(s, T, eng.dcode)
            self._pair_plans[key] = ls
        return self._pair_plans[key]

    def contribute(self):
        """r1 += X1, r2(unsymmetrised) += X2 for the amplitudes held by the engine; Fme must be current."""
        eng, be = self.eng, self.be
        o, v = eng.no, eng.nv
        self.build_intermediates()
        self.tri.refresh_t2T()
        if self.pert:
            self.Vov.copy_(eng.F[:o, o:] - eng.F0[:o, o:])
            for L in self.pert_tmp:
                be.contract(L.desc)
        for (i, j) in self.my_pairs():
            self.tri._launch_pair(i, j, nk=o)
            if self.pert:
                self.pert_w.launch_pair(i, j, o)
            be.t3_combine(self.tri.W, o, 0, i, j, eng.eps, o, v, self.A1, self.A2)
            for L in self._plans_for(i, j):
                be.contract(L.desc)
