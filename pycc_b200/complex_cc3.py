"""
CC3 triples contributions for COMPLEX amplitudes - the call the reference's real-time propagator makes for model
'CC3' (pycc/rt/rtcc.py:125 -> ccwfn.residuals(F, t1, t2, real_time=True), pycc/ccwfn.py:371-401) - on the same
kernels as the real path (cc3_engine.py), with every complex tensor held as two real planes (complex_engine.py):

  * the T1-dressed intermediates (``build_cc3_*``, pycc/ccwfn.py:625-710) are the statements of
    ``CC3Engine.STATEMENTS`` expanded plane by plane;
  * the connected triples of a whole k-batch of an occupied pair (i, j) are the twelve terms of ``t3c_ijk``
    (pycc/cctriples.py:8-35) expanded to their real and imaginary parts and launched through ``TripletBatch``
    (k = batch index of the launch), followed - with ``real_time`` - by the ``t3_pert_ijk`` term
    (cctriples.py:547-562) built from V = F - H.F;
  * the triples denominator is REAL (the diagonal of F must be real: a real field leaves it so), hence
    ``pccb_t3_combine`` - linear in W - is applied to each plane;
  * the four consumer contractions per pair add X1 / X2 (:395-398) into the planes of r1 / r2.

Several GPUs: the occupied pairs are dealt over the ranks as in cc3_engine.py.
"""
import torch

from .cc3_engine import CC3Engine
from .complex_engine import _NAME, expand
from .dsl import TRef, plan
from .triples_engine import TripletBatch

# t3c_ijk with the swapped-pair copies of t2 written as index order (tT_j[mab] = t_j[mba]): no helper copies
_W_STATEMENTS = (
    "W[abc] += Vi[bae]*tkj[ce] + Vj[abe]*tki[ce] - ti[mab]*Ojk[mc] - tj[mba]*Oik[mc]",
    "W[abc] += Vi[cae]*tjk[be] + Vk[ace]*tji[be] - ti[mac]*Okj[mb] - tk[mca]*Oij[mb]",
    "W[abc] += Vk[bce]*tij[ae] + Vj[cbe]*tik[ae] - tk[mcb]*Oji[ma] - tj[mbc]*Oki[ma]",
)
_CONSUMERS = (
    "r1i[a] += A1[kabc]*Lj[kbc]",                    # X1 (:395)
    "r2ij[ab] += A1[kabc]*Fme[kc]",                  # X2 (:396)
    "r2ij[ad] += A2[kabc]*Wk[kdbc]",                 # X2 (:397)
    "r2i[lab] -= A2[kabc]*Wj[klc]",                  # X2 (:398)
)


def _plane_refs(text, tensors, real=()):
    """TRefs for the names of an expanded statement: ``X`` = both planes (2, ...), ``X__r`` / ``X__i`` = one
    plane, names in ``real`` = real tensors."""
    out = {}
    for n in set(_NAME.findall(text)):
        if n in real:
            out[n] = TRef.of(tensors[n], n)
        elif n.endswith("__r") or n.endswith("__i"):
            out[n] = TRef.of(tensors[n[:-3]][0 if n.endswith("__r") else 1], n)
        else:
            out[n] = TRef.of(tensors[n], n)
    return out


def _expanded(statements, is_complex):
    texts = []
    for s in statements:
        for kind, arg in expand(s, is_complex):
            if kind != "plan":
                raise RuntimeError("accumulating statements only: %r" % s)
            texts.append(arg)
    return texts


class ComplexCC3(object):
    def __init__(self, cr):
        """``cr``: the ComplexResiduals of a CC3 engine (its plane set holds t1, t2, Fme, r1, r2, F)."""
        self.cr, self.eng, self.be, self.P = cr, cr.eng, cr.be, cr.P
        eng, P = self.eng, self.P
        if eng.ovvv_full is None:
            raise ValueError("model='CC3' needs the full <mb|ef> block")
        o, v = eng.no, eng.nv
        shapes = dict(T11=(o, o, v, v), Wmnij3=(o, o, o, o), Wo=(o, o, o, v), Wmnie3=(o, o, o, v), X=(o, v, v, o),
                      P=(o, o, v, v), R=(o, o, v, v), Qa=(o, v, o, v), Q=(o, o, o, v), Wk=(o, v, v, v), Wv=(o, v, v, v),
                      W3=(o, v, v, v), A1=(o, v, v, v), A2=(o, v, v, v), Vov=(o, v), Tp=(o, o, v, o))
        for k, s in shapes.items():
            P.add_complex(k, s)
        c = P.cplx
        stm = ["T11[ijab] = t1[ia]*t1[jb]"] + list(CC3Engine.STATEMENTS)
        self.sharded = cr.sharded
        if self.sharded:
            # several GPUs, as in cc3_engine: a rank holds the rows a in [vlo, vhi) of <ab|ef>, so it forms those rows of
            # the one v^4 term of Wabei in a zero-initialised buffer and one allreduce of the disjoint slices gathers
            # them; the occupied pairs are dealt over the ranks and the partial X1 / X2 ride on the allreduce of the
            # r1 / r2 planes in ComplexResiduals.residuals
            stm = [x for x in stm if "vvvv[" not in x]
            P.add_complex("WvG", (o, v, v, v))
            self.gather = P.compile(["WvG_s1[iAbe] += t1[if]*vvvv_s0[Abef]"])
        self.inter = P.compile(stm)
        self.pert_tmp = P.compile(["Tp[ijal] = t2[ijad]*Vov[ld]"])
        every = lambda name: True
        t2, V, O, W = c["t2"], c["Wv"], c["Wo"], c["W3"]

        def w_tensors(i, j, k):
            return dict(W=W[:, k], Vi=V[:, i], Vj=V[:, j], Vk=V[:, k], ti=t2[:, i], tj=t2[:, j], tk=t2[:, k],
                        tij=t2[:, i, j], tik=t2[:, i, k], tji=t2[:, j, i], tjk=t2[:, j, k], tki=t2[:, k, i], tkj=t2[:, k, j],
                        Oij=O[:, i, j], Oik=O[:, i, k], Oji=O[:, j, i], Ojk=O[:, j, k], Oki=O[:, k, i], Okj=O[:, k, j])

        def planes(fn):
            def g(i, j, k):
                out = {}
                for n, t in fn(i, j, k).items():
                    out[n + "__r"], out[n + "__i"] = t[0], t[1]
                return out
            return g

        self.w_batch = TripletBatch(o, eng.dcode, _expanded(_W_STATEMENTS, every), planes(w_tensors), self.be)
        Tp = c["Tp"]
        self.pert_batch = TripletBatch(o, eng.dcode, _expanded(("W[abc] -= tmp[al]*t2k[lcb]",), every),
                                       planes(lambda i, j, k: dict(W=W[:, k], tmp=Tp[:, i, j], t2k=t2[:, k])), self.be)
        self._pair_plans = {}

    def _consumers(self, i, j):
        key = (i, j)
        if key not in self._pair_plans:
            eng, c = self.eng, self.P.cplx
            T = dict(A1=c["A1"], A2=c["A2"], Lj=eng.Loovv[j], Fme=c["Fme"], Wk=c["Wk"], Wj=c["Wmnie3"][:, j],
                     r1i=c["r1"][:, i], r2ij=c["r2"][:, i, j], r2i=c["r2"][:, i])
            ls = []
            for text in _expanded(_CONSUMERS, lambda n: n != "Lj"):
                ls += plan(text, _plane_refs(text, T, real=("Lj",)), eng.dcode)
            self._pair_plans[key] = ls
        return self._pair_plans[key]

    def _bare(self, name, view):
        """Real integrals as the starting value of a complex intermediate."""
        p = self.P.cplx[name]
        self.be.permute4(p[0], view)
        p[1].zero_()

    def contribute(self, real_time=False):
        """r1 += X1, r2 (unsymmetrised) += X2 in the planes of the ComplexResiduals; Fme must be current."""
        eng, be, P = self.eng, self.be, self.P
        c = P.cplx
        o, v = eng.no, eng.nv
        b = eng.blk
        Fd = torch.diagonal(c["F"][1])
        if float(Fd.abs().max()) != 0.0:
            raise ValueError("complex-amplitude CC3: the diagonal of F must be real (the triples denominators are)")
        eps = torch.diagonal(c["F"][0]).contiguous()
        # dressed intermediates: bare parts are permuted views of the stored blocks (cc3_engine.build_intermediates)
        self._bare("Wk", eng.ovvv_full.permute(0, 1, 3, 2))
        self._bare("P", b["ovov"].permute(0, 2, 3, 1))
        self._bare("Q", b["ooov"].permute(2, 1, 0, 3))
        self._bare("R", b["oovv"].permute(1, 0, 3, 2))
        self._bare("Wv", eng.ovvv_full.permute(0, 3, 2, 1))
        P.run(self.inter)
        if self.sharded:
            c["WvG"].zero_()             # the other ranks' rows hold the previous call's gathered values
            P.run(self.gather)
            eng.comm.allreduce_(c["WvG"])
            c["Wv"].add_(c["WvG"])
        if real_time:
            c["Vov"].copy_(c["F"][:, :o, o:])
            c["Vov"][0].sub_(eng.F0[:o, o:])
            P.run(self.pert_tmp)
        W, A1, A2 = c["W3"], c["A1"], c["A2"]
        pairs = [(i, j) for i in range(o) for j in range(o)][eng.comm.rank::eng.comm.world]
        for (i, j) in pairs:
            W.zero_()
            self.w_batch.launch_pair(i, j, o)
            if real_time:
                self.pert_batch.launch_pair(i, j, o)
            for z in range(2):
                be.t3_combine(W[z], o, 0, i, j, eps, o, v, A1[z], A2[z])
            for L in self._consumers(i, j):
                be.contract(L.desc)
