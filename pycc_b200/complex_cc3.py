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

    def build_intermediates(self, real_time=False):
        """T1-dressed integrals (and the field term's Tp) for the amplitudes in the planes; returns the real
        orbital energies of the triples denominators."""
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
        return eps

    def launch_w(self, i, j, real_time=False):
        """W planes of the triplets (i, j, k), k = 0 .. o-1 (no denominator)."""
        self.P.cplx["W3"].zero_()
        self.w_batch.launch_pair(i, j, self.eng.no)
        if real_time:
            self.pert_batch.launch_pair(i, j, self.eng.no)

    def contribute(self, real_time=False):
        """r1 += X1, r2 (unsymmetrised) += X2 in the planes of the ComplexResiduals; Fme must be current."""
        eng, be, c = self.eng, self.be, self.P.cplx
        o, v = eng.no, eng.nv
        eps = self.build_intermediates(real_time)
        W, A1, A2 = c["W3"], c["A1"], c["A2"]
        pairs = [(i, j) for i in range(o) for j in range(o)][eng.comm.rank::eng.comm.world]
        for (i, j) in pairs:
            self.launch_w(i, j, real_time)
            for z in range(2):
                be.t3_combine(W[z], o, 0, i, j, eps, o, v, A1[z], A2[z])
            for L in self._consumers(i, j):
                be.contract(L.desc)


# ---- CC3 terms of the Lambda residuals with complex amplitudes ------------------------------------------------------
def compile_planes(statements, tensors, real, dcode):
    """Statements over named tensors -> [('zero', tensor) | ('launch', Launch)]: complex tensors are (2, ...) plane
    pairs, the names in ``real`` are real tensors; the plane expansion of complex_engine.expand, planned at once."""
    prog = []
    for s in statements:
        for kind, arg in expand(s, lambda n: n not in real):
            if kind == "zero":
                prog.append(("zero", tensors[arg]))
            else:
                prog += [("launch", L) for L in plan(arg, _plane_refs(arg, tensors, real), dcode)]
    return prog


def run_planes(prog, be):
    for kind, x in prog:
        if kind == "zero":
            x.zero_()
        else:
            be.contract(x.desc)


class ComplexCC3Lambda(object):
    """Y1 / Y2 of the CC3 triples in ``cclambda.residuals`` (pycc/cclambda.py:323-440: the t3 and l3 loops of the
    Lambda equations rebuilt for the call's own amplitudes) with complex t and l: the statements of
    cc3_lambda_engine.py plane by plane; ``pccb_t3_density_combine`` (linear, real denominators) per plane."""

    def __init__(self, cr):
        from . import cc3_lambda_engine as rl
        self.rl = rl
        self.cr, self.eng, self.be, self.P = cr, cr.eng, cr.be, cr.P
        eng, P = self.eng, self.P
        if cr.cc3 is None:
            cr.cc3 = ComplexCC3(cr)
        self.cc3 = cr.cc3
        o, v = eng.no, eng.nv
        E = lambda *s: self.be.zeros((2,) + s, eng.dtype)
        self.WK, self.LwK, self.Lo = E(o, v, v, v), E(o, v, v, v), E(o, o, o, v)
        self.L3, self.X, self.P1, self.P2 = E(o, v, v, v), E(o, v, v, v), E(o, v, v, v), E(o, v, v, v)
        self.l2T = E(o, o, v, v)
        self.Zb, self.Zmd = E(o, v, v, v), E(o, v, v, v)
        self.Zma, self.Zd1, self.Zmn = E(o, o, o, v), E(o, o, o, v), E(o, o, v, o)
        self.Znf = E(o, v)
        self.Xje, self.Xej = E(o, v, o, v), E(o, v, v, o)
        self.Q2, self.Q3 = E(o, o, o, v), E(o, o, o, v)
        self.WA, self.WB, self.WC, self.WD = E(o, o, v, v), E(o, o, v, v), E(o, o, v, v), E(o, o, v, v)
        # several GPUs, as in cc3_lambda_engine: the dressed <ab|ef> holds the rank's rows a only; the occupied pairs of
        # both loops are dealt round-robin; every Z intermediate and Y2 is a sum over pairs (one allreduce each), and the
        # one term that reads the dressed <ab|ef> gives a partial Y1 (allreduce).  One GPU: the same code path.
        vlo, vl = eng.vlo, eng.vl
        self.Wab = E(vl, v, v, v)
        self.Y1w, self.Y2 = E(o, v), E(o, o, v, v)
        c, b = P.cplx, eng.blk
        T = dict(ovvv=eng.ovvv_full, ooov=b["ooov"], oovv=b["oovv"], t1=c["t1"], T11=c["T11"], Xje=self.Xje, Xej=self.Xej,
                 Q2=self.Q2, Q3=self.Q3, Wab=self.Wab, t1A=c["t1"].narrow(2, vlo, vl),
                 ovvvA=eng.ovvv_full.narrow(1, vlo, vl), T11A=c["T11"].narrow(3, vlo, vl))
        self.dressed = compile_planes(rl._DRESSED, T, ("ovvv", "ooov", "oovv", "ovvvA"), eng.dcode)
        T = dict(r1=c["rl1"], l2=c["l2"], Zmd=self.Zmd, Loovv=eng.Loovv, Znf=self.Znf, Zmn=self.Zmn, Zb=self.Zb,
                 Wab=self.Wab, Y1w=self.Y1w, ZbD=self.Zb.narrow(3, vlo, vl), Woooo=c["Wmnij3"], Zma=self.Zma, Zd1=self.Zd1,
                 WA=self.WA, WB=self.WB, WC=self.WC, WD=self.WD)
        self.y1 = compile_planes(rl._Y1, T, ("Loovv",), eng.dcode)
        # l3 of one triplet: the spin-adapted integrals L[x,y] are the only real operands
        import re
        is_real = lambda n: re.match(r"^L[ijk][ijk]$", n) is not None
        texts = []
        for s in rl._L3:
            for kind, arg in expand(s.replace(" = ", " += ", 1), lambda n: not is_real(n)):
                texts.append(arg)
        self.l3 = TripletBatch(o, eng.dcode, texts, self._l3_tensors, self.be)
        self._pair, self._setup_plans = {}, {}

    def _l3_tensors(self, i, j, k):
        e, c = self.eng, self.P.cplx
        l1, l2, L, f, Wo3 = c["l1"], c["l2"], e.Loovv, c["Fme"], c["Wmnie3"]
        d = dict(W=self.L3[:, k])
        ix = dict(i=i, j=j, k=k)
        for x in "ijk":
            d["LwK" + x], d["WK" + x] = self.LwK[:, ix[x]], self.WK[:, ix[x]]
            d["l2" + x], d["l1" + x], d["f" + x] = l2[:, ix[x]], l1[:, ix[x]], f[:, ix[x]]
            d["l2T" + x] = self.l2T[:, ix[x]]
            for y in "ijk":
                if x != y:
                    d["l2" + x + y] = l2[:, ix[x], ix[y]]
                    d["l2T" + x + y] = self.l2T[:, ix[x], ix[y]]
                    d["Lo" + x + y] = self.Lo[:, ix[x], ix[y]]
                    d["Wo" + x + y] = Wo3[:, ix[x], ix[y]]
        out = {}
        for n, t in d.items():            # both planes under the plain name, one plane under __r / __i
            out[n], out[n + "__r"], out[n + "__i"] = t, t[0], t[1]
        for x in "ijk":
            for y in "ijk":
                if x != y:
                    out["L" + x + y] = L[ix[x], ix[y]]
        return out

    def _planes(self, fn, *args):
        for z in range(2):
            fn(z, *args)

    def setup(self, real_time):
        """Everything that depends on the T amplitudes only (cclambda.py:118-139), for the planes' (F, t1, t2);
        the Fme planes must be current."""
        e, be, cc3, c = self.eng, self.be, self.cc3, self.P.cplx
        o, v = e.no, e.nv
        b = e.blk
        eps = self.eps = cc3.build_intermediates(real_time)
        for z in range(2):
            be.permute4(self.WK[z], c["Wk"][z].permute(0, 2, 3, 1))
            be.permute4(self.LwK[z], self.WK[z].permute(0, 2, 1, 3))
            be.axpby(self.LwK[z], 2.0, self.WK[z], -1.0, self.LwK[z])
            be.permute4(self.Lo[z], c["Wmnie3"][z].permute(1, 0, 2, 3))
            be.axpby(self.Lo[z], 2.0, c["Wmnie3"][z], -1.0, self.Lo[z])
        self.Wab[0].copy_(e.vvvv_loc)
        self.Wab[1].zero_()
        run_planes(self.dressed, be)
        self.Xje[0].add_(b["ovov"])                                               # <mb|je>
        self.Xej[0].add_(b["oovv"].permute(0, 3, 2, 1))                           # <mb|ej> = <mj|eb>
        for z in range(2):
            be.permute4(self.WA[z], self.Xje[z].permute(2, 0, 1, 3))
            be.permute4(self.WB[z], self.Xje[z].permute(0, 2, 3, 1))
            be.permute4(self.WC[z], self.Xej[z].permute(3, 0, 1, 2))
            be.permute4(self.WD[z], self.Xej[z].permute(0, 3, 2, 1))
        self.Zmn.zero_()
        self.Zmd.zero_()
        W, A2 = c["W3"], c["A2"]
        for (l, m) in self.my_pairs():
            cc3.launch_w(l, m, real_time)
            for z in range(2):
                be.t3_density_combine(W[z], W[z], o, 0, l, m, eps, o, v, dict(M=A2[z]))
            run_planes(self._setup_for(l, m), be)
        if e.comm.world > 1:
            e.comm.allreduce_(self.Zmn)
            e.comm.allreduce_(self.Zmd)

    def my_pairs(self):
        o, cm = self.eng.no, self.eng.comm
        return [(i, j) for i in range(o) for j in range(o)][cm.rank::cm.world]

    def _setup_for(self, l, m):
        key = (l, m)
        if key not in self._setup_plans:
            e = self.eng
            T = dict(T=self.P.cplx["A2"], gl=e.blk["oovv"][:, l], Ll=e.Loovv[:, l], Zmn_m=self.Zmn[:, m], Zmd_m=self.Zmd[:, m])
            self._setup_plans[key] = compile_planes(self.rl._T3_SETUP, T, ("gl", "Ll"), e.dcode)
        return self._setup_plans[key]

    def _pair_plans(self, i, j):
        key = (i, j)
        if key not in self._pair:
            e, c = self.eng, self.P.cplx
            T = dict(X=self.X, P1=self.P1, P2=self.P2, t2=c["t2"], t2j=c["t2"][:, j], Zbi=self.Zb[:, i], Zmaij=self.Zma[:, i, j],
                     Zd1ij=self.Zd1[:, i, j], Y2ij=self.Y2[:, i, j], Y2i=self.Y2[:, i], Wv=c["Wv"], Woj=c["Wo"][:, j],
                     A1=c["A1"], l2ij=c["l2"][:, i, j], Znf=self.Znf)
            con = compile_planes(self.rl._L3_CONSUMERS, T, (), e.dcode)
            znf = compile_planes(("Znf[nf] += A1[ndef]*l2ij[de]",), T, (), e.dcode)
            self._pair[key] = (con, znf)
        return self._pair[key]

    def contribute(self, real_time=False):
        """rl1 += Y1, rl2 (unsymmetrised) += Y2 for the l1 / l2 planes (cclambda.py:160-235)."""
        e, be, cc3, c = self.eng, self.be, self.cc3, self.P.cplx
        o, v = e.no, e.nv
        self.setup(real_time)
        eps = self.eps
        for z in range(2):
            be.permute4(self.l2T[z], c["l2"][z].permute(0, 1, 3, 2))
        for t in (self.Znf, self.Zb, self.Zma, self.Zd1, self.Y2, self.Y1w):
            t.zero_()
        W, A1 = c["W3"], c["A1"]
        for (i, j) in self.my_pairs():
            con, znf = self._pair_plans(i, j)
            cc3.launch_w(i, j, real_time)                   # t3 of (l, m, n) = (i, j, batch): Znf
            for z in range(2):
                be.t3_density_combine(W[z], W[z], o, 0, i, j, eps, o, v, dict(A1=A1[z]))
            run_planes(znf, be)
            self.L3.zero_()                                 # l3 of (k, i, j), k = batch
            self.l3.launch_pair(i, j, o)
            for z in range(2):
                be.t3_density_combine(self.L3[z], self.L3[z], o, 0, i, j, eps, o, v, dict(M=self.X[z]))
                be.permute4(self.P1[z], self.X[z].permute(0, 3, 1, 2))
                be.permute4(self.P2[z], self.X[z].permute(0, 2, 1, 3))
            run_planes(con, be)
        if e.comm.world > 1:
            for t in (self.Znf, self.Zb, self.Zma, self.Zd1, self.Y2):
                e.comm.allreduce_(t)
        run_planes(self.y1, be)          # (rl1 holds the summed builders' part by now: complex_engine.lambda_residuals)
        if e.comm.world > 1:
            e.comm.allreduce_(self.Y1w)
        c["rl1"].add_(self.Y1w)
        c["rl2"].add_(self.Y2)
