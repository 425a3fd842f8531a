"""
CC3 contributions to the Lambda equations: the reference's loops over all occupied
triplets inside ``cclambda.solve_lambda`` (pycc/cclambda.py:118-139 setup, :160-235 per iteration) with
``cctriples.l3_ijk`` (:219-277), on the engine's kernels.

Per Lambda iteration the reference evaluates, for every (i, j, k): t3c_ijk (12 einsums), l3_ijk (24 einsums + 24
outer products) and eight consumer einsums.  Here ONE loop over the o^2 occupied pairs with k as the batch index:

  * l3 of all k: 12 batched ``pccb_contract`` launches per pair (6 with the 24 contractions K-concatenated four
    at a time by output structure, 6 with the 24 outer products as K = 1 segments), written for one triplet over
    slices of the stored tensors and batched by pointer differencing (triples_engine.TripletBatch);
    ``pccb_t3_density_combine`` applies the denominator, two permuted copies give the index orders the
    consumers contract over contiguously;
  * five consumer contractions per pair with k folded into the contracted index (Zbide, Zjlma, Zjlid, Y2 x 2);
    Zblad_1 / Zblad_2 / Zjlid_2 of the reference are index relabellings of Zbide / Zjlma and are not formed;
  * t3 of all k (for Znf) from the (T) machinery over the T1-dressed intermediates of cc3_engine.py;
  * the T-only intermediates Zmndi, Zmdfa and the three extra dressed integrals (Wmbje, Wmbej, Wabef,
    cclambda.py:546-577) once per solve.

No o^3 v^3 tensor exists.  Several GPUs: pairs dealt over the ranks, see __init__.  The amplitudes t1, t2 are those held by the CCSD engine (converged T).
"""
import torch

from .dsl import TRef, plan
from .triples_engine import TripletBatch

# l3_ijk(k, i, j, ...) for ONE triplet; virtual indices (a, b, c) belong to the occupied (k, i, j).
# LwK[m][x][y][e] = 2 Wamef[e,m,x,y] - Wamef[e,m,y,x], WK[m][x][y][e] = Wamef[e,m,x,y];
# Lo[m][n] = 2 Wmnie[m,n] - Wmnie[n,m], Wo = Wmnie;  Lxy = L[x,y,v,v], fx = Fme[x], l1x = l1[x].
_L3 = (
    # cctriples.py:237-250 (particle terms)
    "W[abc] = LwKi[abe]*l2jk[ce] + LwKk[bae]*l2ji[ce] - WKj[bae]*l2ik[ec] - WKj[abe]*l2ki[ec]",
    "W[abc] += LwKj[ace]*l2ik[be] + LwKk[cae]*l2ij[be] - WKi[ace]*l2kj[eb] - WKi[cae]*l2jk[eb]",
    "W[abc] += LwKj[bce]*l2ki[ae] + LwKi[cbe]*l2kj[ae] - WKk[bce]*l2ij[ea] - WKk[cbe]*l2ji[ea]",
    # :252-265 (hole terms); l2T = l2 with its virtual pair swapped, so that every operand walks (a, b, c) in the
    # output's order along a contiguous direction (16-byte staging instead of the element-wise gather)
    "W[abc] -= l2Ti[mab]*Lokj[mc] + l2k[mab]*Loij[mc] - l2Tj[mab]*Woki[mc] - l2j[mab]*Woik[mc]",
    "W[abc] -= l2Tj[mac]*Loki[mb] + l2k[mac]*Loji[mb] - l2Ti[mac]*Wokj[mb] - l2i[mac]*Wojk[mb]",
    "W[abc] -= l2Tj[mbc]*Loik[ma] + l2i[mbc]*Lojk[ma] - l2k[mbc]*Woji[ma] - l2Tk[mbc]*Woij[ma]",
    # :220-225 and :227-232 (disconnected terms), grouped by the index of the vector factor; L[x,y][q,p] = L[y,x][p,q]
    # (8-fold symmetry of the integrals) puts the L factors in the output's index order as well
    "W[abc] += Lki[ab]*l1j[c] - Lkj[ab]*l1i[c] + Lki[ab]*l1j[c] - Lji[ab]*l1k[c]",
    "W[abc] += fj[c]*l2ki[ab] - fk[c]*l2ji[ab] - fi[c]*l2Tjk[ab] + fj[c]*l2Tik[ab]",
    "W[abc] += Lkj[ac]*l1i[b] - Lki[ac]*l1j[b] + Lkj[ac]*l1i[b] - Lij[ac]*l1k[b]",
    "W[abc] += fi[b]*l2kj[ac] - fk[b]*l2ij[ac] + fi[b]*l2Tjk[ac] - fj[b]*l2Tik[ac]",
    "W[abc] += Lij[bc]*l1k[a] - Lik[bc]*l1j[a] + Lij[bc]*l1k[a] - Lkj[bc]*l1i[a]",
    "W[abc] += fk[a]*l2ij[bc] - fi[a]*l2kj[bc] + fk[a]*l2Tji[bc] - fj[a]*l2Tki[bc]",
)

# per occupied pair (i, j), k folded into the contraction; X[k][abc] = l3_kij, P1[k][bca] = X[k][cab],
# P2[k][bac] = X[k][abc] (first two virtual indices swapped)
_L3_CONSUMERS = (
    "Zbi[bde] += t2j[kbc]*X[kcde]",           # :208 Zbide[:,i]  (= Zblad_1; Zblad_2 is its transpose)
    "Zmaij[xa] += t2[xkbc]*P1[kbca]",         # :211 Zjlma[:,i,j]  (= Zjlid_2, :213)
    "Zd1ij[xd] += t2[xkbc]*P2[kbcd]",         # :212 Zjlid_1[:,i,j]
    "Y2ij[ab] += Wv[keda]*P2[kedb]",          # :215 Y2[i,j] += l3[deb] Wvvvo[e,d,a,k]
    "Y2i[lab] -= Woj[kld]*X[kdab]",           # :216 Y2[i]  -= l3[dab] Wovoo[l,d,j,k]
)

# once per solve, per pair (l, m) with n folded / batched; T[n][def] = t3_lmn (with denominator)
_T3_SETUP = (
    "Zmn_m[ndi] += T[ndef]*gl[ief]",          # :136 Zmndi[m,n] += t3[def] <il|ef>
    "Zmn_m[ndi] -= T[nfed]*Ll[ief]",          # :137             - t3[fed] L[i,l,e,f]
    "Zmd_m[dfa] += T[ndef]*gl[nea]",          # :138 Zmdfa[m] += t3[def] <nl|ea>
    "Zmd_m[dfa] -= T[ndfe]*Ll[nea]",          # :139           - t3[dfe] L[n,l,e,a]
)

# after the pair loops (cclambda.py:196-201, 218-230)
_Y1 = (
    "r1[ia] += l2[imdf]*Zmd[mdfa]",           # :197
    "r1[ia] += Loovv[imaf]*Znf[mf]",          # :198
    "r1[ia] += l2[mnad]*Zmn[mndi]",           # :200
    "Y1w[ia] += ZbD[ibde]*Wab[deab]",         # :218 Zbide . Wvvvv (the rank's rows d of the dressed <de|ab>)
    "r1[ia] += Woooo[ijlm]*Zma[lmja]",        # :219-222
    "r1[ia] -= Zd1[lijd]*WA[ljda]",           # :225 Zjlid_1[j,l] . Wovov[j,:,l,:]
    "r1[ia] -= Zma[lijd]*WC[ljda]",           # :226 Zjlid_2[j,l] . Wovvo[j,:,:,l]
    "r1[ia] -= WB[ilbd]*Zb[lbad]",            # :228 Zblad_1[:,l] . Wovov[:,:,l,:]
    "r1[ia] -= WD[ilbd]*Zb[lbda]",            # :229 Zblad_2[:,l] . Wovvo[:,:,:,l]
)

# the three additional dressed integrals (cclambda.py:546-577) in the layouts _Y1 reads:
#   WA[l][j][d][a] = Wmbje[j,d,l,a]   WB[i][l][b][d] = Wmbje[i,d,l,b]
#   WC[l][j][d][a] = Wmbej[j,d,a,l]   WD[i][l][b][d] = Wmbej[i,d,b,l]
#   Wab[a][b][e][f] = Wabef
_DRESSED = (
    "Xje[mbje] = ovvv[mbfe]*t1[jf]",
    "Xje[mbje] -= ooov[mnje]*t1[nb]",
    "Q2[jmne] = t1[jf]*oovv[mnfe]",           # <mn|fe> t1[jf]
    "Xje[mbje] -= Q2[jmne]*t1[nb]",
    "Xej[mbej] = ovvv[mbef]*t1[jf]",
    "Xej[mbej] -= ooov[nmje]*t1[nb]",         # <mn|ej> = <nm|je>
    "Q3[jmne] = t1[jf]*oovv[mnef]",
    "Xej[mbej] -= Q3[jmne]*t1[nb]",
    "Wab[abef] -= t1A[ma]*ovvv[mbef]",        # rows a of this rank's virtual slice (all of them on one GPU)
    "Wab[abef] -= ovvvA[mafe]*t1[mb]",
    "Wab[abef] += T11A[mnab]*oovv[mnef]",
)


class CC3LambdaEngine(object):
    def __init__(self, lam):
        """``lam``: hbar_engine.LambdaEngine of a CC3 wave function."""
        self.lam, self.eng, self.be = lam, lam.eng, lam.be
        eng = self.eng
        if eng.cc3 is None:
            from .cc3_engine import CC3Engine
            eng.cc3 = CC3Engine(eng)
        self.cc3 = cc3 = eng.cc3
        o, v, dt = eng.no, eng.nv, eng.dtype
        E = lambda *s: self.be.zeros(s, dt)
        self.WK, self.LwK = E(o, v, v, v), E(o, v, v, v)
        self.Lo = E(o, o, o, v)
        self.L3, self.X, self.P1, self.P2 = E(o, v, v, v), E(o, v, v, v), E(o, v, v, v), E(o, v, v, v)
        self.l2T = E(o, o, v, v)
        self.Zb, self.Zmd = E(o, v, v, v), E(o, v, v, v)
        self.Zma, self.Zd1, self.Zmn = E(o, o, o, v), E(o, o, o, v), E(o, o, v, o)
        self.Znf = E(o, v)
        self.Xje, self.Xej = E(o, v, o, v), E(o, v, v, o)
        self.Q2, self.Q3 = E(o, o, o, v), E(o, o, o, v)
        self.WA, self.WB, self.WC, self.WD = E(o, o, v, v), E(o, o, v, v), E(o, o, v, v), E(o, o, v, v)
        # several GPUs: the dressed <ab|ef> holds the rank's rows a only (like <ab|ef> itself); the occupied pairs of
        # both loops are dealt round-robin; every Z intermediate and Y2 is a sum over pairs (one allreduce each), and
        # the one term that reads the dressed <ab|ef> gives a partial Y1 (allreduce).  One GPU: the same code path.
        vlo, vl = eng.vlo, eng.vl
        self.Wab = E(vl, v, v, v)
        self.Y1w, self.Y2 = E(o, v), E(o, o, v, v)
        b = eng.blk
        self._b = b
        T = dict(ovvv=eng.ovvv_full, ooov=b["ooov"], oovv=b["oovv"], t1=eng.t1, T11=cc3.T11, Xje=self.Xje, Xej=self.Xej,
                 Q2=self.Q2, Q3=self.Q3, Wab=self.Wab, t1A=eng.t1.narrow(1, vlo, vl),
                 ovvvA=eng.ovvv_full.narrow(1, vlo, vl), T11A=cc3.T11.narrow(2, vlo, vl))
        T = {k: TRef.of(t, k) for k, t in T.items()}
        self.dressed = [L for s in _DRESSED for L in plan(s, T, eng.dcode)]
        self.l3 = TripletBatch(o, eng.dcode, _L3, self._l3_tensors, self.be)
        T = dict(r1=lam.r1, l2=lam.l2, Zmd=self.Zmd, Loovv=eng.Loovv, Znf=self.Znf, Zmn=self.Zmn, Zb=self.Zb, Wab=self.Wab,
                 Y1w=self.Y1w, ZbD=self.Zb.narrow(2, vlo, vl),
                 Woooo=cc3.Wmnij3, Zma=self.Zma, Zd1=self.Zd1, WA=self.WA, WB=self.WB, WC=self.WC, WD=self.WD)
        T = {k: TRef.of(t, k) for k, t in T.items()}
        self.y1 = [L for s in _Y1 for L in plan(s, T, eng.dcode)]
        self._pair = {}
        self._setup_plans = {}
        self.ready = False
        self.pert = False            # set by cclambda.residuals when ccwfn.real_time is on and F carries a field

    # ---- tensors of one triplet for the l3 statements -------------------------------------------------------
    def _l3_tensors(self, i, j, k):
        e, lam = self.eng, self.lam
        l1, l2, L, f = lam.l1, lam.l2, e.Loovv, e.Fme
        Wo3 = self.cc3.Wmnie3
        d = dict(W=self.L3[k])
        ix = dict(i=i, j=j, k=k)
        for x in "ijk":
            d["LwK" + x], d["WK" + x] = self.LwK[ix[x]], self.WK[ix[x]]
            d["l2" + x], d["l1" + x], d["f" + x] = l2[ix[x]], l1[ix[x]], f[ix[x]]
            d["l2T" + x] = self.l2T[ix[x]]
            for y in "ijk":
                if x != y:
                    d["l2" + x + y] = l2[ix[x], ix[y]]
                    d["l2T" + x + y] = self.l2T[ix[x], ix[y]]
                    d["L" + x + y] = L[ix[x], ix[y]]
                    d["Lo" + x + y] = self.Lo[ix[x], ix[y]]
                    d["Wo" + x + y] = Wo3[ix[x], ix[y]]
        return d

    def setup(self):
        """Everything that depends on the T amplitudes only (cclambda.py:118-139)."""
        e, be, cc3 = self.eng, self.be, self.cc3
        o, v = e.no, e.nv
        be.bind_stream()
        e.run_builder("Fme")                       # Fov = build_Fme(F, t1) for the Fock matrix the engine holds
        cc3.build_intermediates()                  # Wmnij, Wmbij, Wmnie, Wamef, Wabei at the converged t1
        b = self._b
        # Wvovv[e,m,x,y] = Wamef[e,m,x,y] = Wk[m][e][x][y]
        be.permute4(self.WK, cc3.Wk.permute(0, 2, 3, 1))
        be.permute4(self.LwK, self.WK.permute(0, 2, 1, 3))
        be.axpby(self.LwK, 2.0, self.WK, -1.0, self.LwK)
        be.permute4(self.Lo, cc3.Wmnie3.permute(1, 0, 2, 3))
        be.axpby(self.Lo, 2.0, cc3.Wmnie3, -1.0, self.Lo)
        # Wmbje / Wmbej / Wabef: bare parts + dressing, then the layouts the final contractions read
        self.Wab.copy_(e.vvvv_loc)
        for L in self.dressed:
            be.contract(L.desc)
        be.axpby(self.Xje, 1.0, self.Xje, 1.0, b["ovov"])                        # <mb|je>
        bare = self.WA.view(o, v, v, o)                                           # scratch (WA is filled below)
        be.permute4(bare, b["oovv"].permute(0, 3, 2, 1))                          # <mb|ej> = <mj|eb>
        be.axpby(self.Xej, 1.0, self.Xej, 1.0, bare)
        be.permute4(self.WA, self.Xje.permute(2, 0, 1, 3))       # WA[l,j,d,a] = Wmbje[j,d,l,a]
        be.permute4(self.WB, self.Xje.permute(0, 2, 3, 1))       # WB[i,l,b,d] = Wmbje[i,d,l,b]
        be.permute4(self.WC, self.Xej.permute(3, 0, 1, 2))       # WC[l,j,d,a] = Wmbej[j,d,a,l]
        be.permute4(self.WD, self.Xej.permute(0, 3, 2, 1))       # WD[i,l,b,d] = Wmbej[i,d,b,l]
        # Zmndi, Zmdfa from t3 of every (l, m, n)
        self.Zmn.zero_()
        self.Zmd.zero_()
        cc3.tri.refresh()
        outs = dict(M=cc3.A2)
        if self.pert:                              # ccwfn.real_time: t3 -= t3_pert_ijk (cclambda.py:350-356, 396-402)
            cc3.Vov.copy_(e.F[:o, o:] - e.F0[:o, o:])
            for L in cc3.pert_tmp:
                be.contract(L.desc)
        for (l, m) in self.my_pairs():
            cc3.tri._launch_pair(l, m, nk=o)
            if self.pert:
                cc3.pert_w.launch_pair(l, m, o)
            be.t3_density_combine(cc3.tri.W, cc3.tri.W, o, 0, l, m, e.eps, o, v, outs)
            for L in self._setup_for(l, m):
                be.contract(L.desc)
        if e.comm.world > 1:
            e.comm.allreduce_(self.Zmn)
            e.comm.allreduce_(self.Zmd)
        self.ready = True

    def my_pairs(self):
        o, c = self.eng.no, self.eng.comm
        return [(i, j) for i in range(o) for j in range(o)][c.rank::c.world]

    def _setup_for(self, l, m):
        key = (l, m)
        if key not in self._setup_plans:
            e = self.eng
            b = self._b
            T = dict(T=self.cc3.A2, gl=b["oovv"][:, l], Ll=e.Loovv[:, l], Zmn_m=self.Zmn[m], Zmd_m=self.Zmd[m])
            T = {k: TRef.of(t, k) for k, t in T.items()}
            self._setup_plans[key] = [L for s in _T3_SETUP for L in plan(s, T, e.dcode)]
        return self._setup_plans[key]

    def _pair_plans(self, i, j):
        key = (i, j)
        if key not in self._pair:
            e, lam, cc3 = self.eng, self.lam, self.cc3
            T = dict(X=self.X, P1=self.P1, P2=self.P2, t2=e.t2, t2j=e.t2[j], Zbi=self.Zb[i], Zmaij=self.Zma[i, j],
                     Zd1ij=self.Zd1[i, j], Y2ij=self.Y2[i, j], Y2i=self.Y2[i], Wv=cc3.Wv, Woj=cc3.Wo[j],
                     A1=cc3.A1, l2ij=lam.l2[i, j], Znf=self.Znf)
            T = {k: TRef.of(t, k) for k, t in T.items()}
            con = [L for s in _L3_CONSUMERS for L in plan(s, T, e.dcode)]
            znf = plan("Znf[nf] += A1[ndef]*l2ij[de]", T, e.dcode)     # :193, pair (l, m) = (i, j), batch n
            self._pair[key] = (con, znf)
        return self._pair[key]

    def contribute(self):
        """r1 += Y1, r2 (unsymmetrised) += Y2 for the Lambda amplitudes held by ``lam`` (cclambda.py:160-235);
        the final r2 + r2^T of the caller supplies Y2^T."""
        if not self.ready:
            self.setup()
        e, be, cc3, lam = self.eng, self.be, self.cc3, self.lam
        o, v = e.no, e.nv
        be.bind_stream()
        be.permute4(self.l2T, lam.l2.permute(0, 1, 3, 2))
        acc = (self.Znf, self.Zb, self.Zma, self.Zd1, self.Y2)
        for z in acc + (self.Y1w,):
            z.zero_()
        t_out, x_out = dict(A1=cc3.A1), dict(M=self.X)
        for (i, j) in self.my_pairs():
            con, znf = self._pair_plans(i, j)
            # t3 of (l, m, n) = (i, j, batch): Znf
            cc3.tri._launch_pair(i, j, nk=o)
            if self.pert:
                cc3.pert_w.launch_pair(i, j, o)
            be.t3_density_combine(cc3.tri.W, cc3.tri.W, o, 0, i, j, e.eps, o, v, t_out)
            for L in znf:
                be.contract(L.desc)
            # l3 of (k, i, j), k = batch
            self.l3.launch_pair(i, j, o)
            be.t3_density_combine(self.L3, self.L3, o, 0, i, j, e.eps, o, v, x_out)
            be.permute4(self.P1, self.X.permute(0, 3, 1, 2))       # P1[k][b][c][a] = X[k][c][a][b]
            be.permute4(self.P2, self.X.permute(0, 2, 1, 3))       # P2[k][b][a][c] = X[k][a][b][c]
            for L in con:
                be.contract(L.desc)
        if e.comm.world > 1:
            for z in acc:
                e.comm.allreduce_(z)
        for L in self.y1:
            be.contract(L.desc)
        if e.comm.world > 1:
            e.comm.allreduce_(self.Y1w)
        be.axpby(lam.r1, 1.0, lam.r1, 1.0, self.Y1w)
        be.axpby(lam.r2, 1.0, lam.r2, 1.0, self.Y2)
