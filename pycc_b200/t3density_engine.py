"""
(T) contributions to the Lambda equations and the one- / two-electron densities: ``ccwfn.t3_density``
(pycc/ccwfn.py:721-855), SURVEY.md section 8(f) rank 2.

The reference loops over all o^3 occupied triplets (a dozen einsums and v^3 temporaries each) and then, for the
occupied-occupied density, over all v^3 virtual triplets with o^3 tensors.  Here ONE loop over the o^2 occupied
pairs (i, j) does everything, on the (T) machinery of triples_engine.py:

  * the connected triples of all k come from the same three batched launches as (T), the disconnected ones
    (t3d_ijk) from three outer-product launches; ``pccb_t3_density_combine`` applies the denominator and forms
    the seven combinations the contractions below consume (M, A1, A2, B1, Z3, P, Q);
  * twelve contractions per pair, with k folded into the contracted index, accumulate X2 (the (T) energy in
    "Vikings" form), S1, S2, Dvv, Dov, Goovv, Gooov, Gvvvo;
  * Doo: the reference's second loop evaluates  Doo[i,j] = -1/2 sum_{kl,abc} M^{ikl}_{abc} (X+Y)^{jkl}_{abc}.
    t3 is invariant under SIMULTANEOUS permutation of its occupied and virtual triples, and X is a class function
    of the virtual permutation, so  M^{ikl}_{abc} = M^{kli}_{bca}  and likewise for X + Y: with the pair (k, l)
    fixed, the batch members x, y of this loop give  Doo[x,y] -= 1/2 <M[x], Q[y]>  - one more contraction.
    (The symmetry of the disconnected part's t2 . F_ov terms needs t2[ijab] = t2[jiba], which every iterate of
    solve_cc satisfies; for hand-made unsymmetric t2 AND a non-diagonal Fock matrix Doo differs from the reference's
    second loop at the 1e-3 relative level - every other quantity is exact for any input.)

No o^3 v^3 tensor exists (nine o v^3 buffers).  Several GPUs: pairs dealt over the ranks, one allreduce per accumulator.
"""
import torch

from .dsl import TRef, plan
from .triples_engine import TriplesEngine

_DISCONNECTED = (          # t3d_ijk for all k at once (pycc/cctriples.py:69-74)
    "Dd[kabc] = gij[ab]*t1[kc] + tij[ab]*fov[kc]",
    "Dd[kabc] += gi[kac]*t1j[b] + ti[kac]*fj[b]",
    "Dd[kabc] += gj[kbc]*t1i[a] + tj[kbc]*fi[a]",
)

_CONTRACTIONS = (          # pycc/ccwfn.py:757-778, k folded into the contraction
    "X2ij[ab] += A1[kabc]*fov[kc]",                 # :758
    "X2ij[ad] += A2[kabc]*ovT[kdbc]",               # :759   ERI[v,k,v,v][d,b,c] = <dk|bc> = <kd|cb> = ovT[k,d,b,c]
    "X2i[lab] -= A2[kabc]*ooovj[klc]",              # :760   ERI[j,k,o,v]
    "Dvv[ab] += 0.5*M[kacd]*Q[kbcd]",               # :763
    "Dovi[a] += A1[kabc]*t2xj[kbc]",                # :766   4 t2[j,k] - 2 t2[j,k]^T
    "Goovvij[ab] += 4.0*Z3[kabc]*t1[kc]",           # :770
    "Gooovji[la] -= P[kabc]*t2[lkbc]",              # :771
    "Gsj[abd] += P[kabc]*t2ci[kcd]",                # :772   Gvvvo[:,:,:,j], t2[k,i,c,d]
    "S1i[a] += B1[kabc]*Loovvj[kbc]",               # :775
    "S2i[lab] -= P[kabc]*ooovj[klc]",               # :777
    "S2ij[ad] += P[kabc]*ovT[kdbc]",                # :778   ERI[k,v,v,v][d,c,b] = ovT[k,d,b,c]
    "Doo[xy] -= 0.5*M[xabc]*Q[yabc]",               # :782-790 through the permutation symmetry (module docstring)
)


class T3DensityEngine(object):
    def __init__(self, eng):
        if eng.ovvv_full is None:
            raise ValueError("t3_density needs the full <mb|ef> block")
        self.eng, self.be = eng, eng.be
        o, v, dt = eng.no, eng.nv, eng.dtype
        E = lambda *s: self.be.zeros(s, dt)
        self.tri = TriplesEngine(eng)
        self.Dd = E(o, v, v, v)
        self.buf = {k: E(o, v, v, v) for k in ("M", "A1", "A2", "B1", "Z3", "P", "Q")}
        self.t2x = E(o, o, v, v)
        self.X2 = E(o, o, v, v)
        self.S1, self.S2 = E(o, v), E(o, o, v, v)
        self.Dvv, self.Doo, self.Dov = E(v, v), E(o, o), E(o, v)
        self.Goovv, self.Gooov = E(o, o, v, v), E(o, o, o, v)
        self.Gs = E(o, v, v, v)                       # Gs[j][a][b][d] = Gvvvo[a,b,d,j]
        # integrals only, formed once: ovT[k][d][b][c] = <kd|cb>, so that the contracted (k, b, c) of the two
        # o v^4 terms is a two-level index in both operands (the CC3 engine keeps the same copy)
        self.ovT = self.be.empty((o, v, v, v), dt)
        self.be.permute4(self.ovT, eng.ovvv_full.permute(0, 1, 3, 2))
        self._plans = {}

    @property
    def Gvvvo(self):
        return self.Gs.permute(1, 2, 3, 0)

    def _plans_for(self, i, j):
        key = (i, j)
        if key not in self._plans:
            e = self.eng
            o = e.no
            b = e.blk
            fov = e.F[:o, o:]
            T = dict(Dd=self.Dd, gij=b["oovv"][i, j], tij=e.t2[i, j], gi=b["oovv"][i], ti=e.t2[i], gj=b["oovv"][j],
                     tj=e.t2[j], t1=e.t1, fov=fov, t1i=e.t1[i], t1j=e.t1[j], fi=fov[i], fj=fov[j],
                     ovT=self.ovT, ooovj=b["ooov"][j], t2xj=self.t2x[j], t2=e.t2, t2ci=e.t2[:, i], Loovvj=e.Loovv[j],
                     X2ij=self.X2[i, j], X2i=self.X2[i], Dvv=self.Dvv, Dovi=self.Dov[i], Goovvij=self.Goovv[i, j],
                     Gooovji=self.Gooov[j, i], Gsj=self.Gs[j], S1i=self.S1[i], S2i=self.S2[i], S2ij=self.S2[i, j],
                     Doo=self.Doo)
            T.update(self.buf)
            T = {k: TRef.of(t, k) for k, t in T.items()}
            dis = [L for s in _DISCONNECTED for L in plan(s, T, e.dcode)]
            con = [L for s in _CONTRACTIONS for L in plan(s, T, e.dcode)]
            self._plans[key] = (dis, con)
        return self._plans[key]

    def run(self):
        """All (T) source terms and densities for the amplitudes held by the engine.  Returns E(T)
        (pycc/ccwfn.py:803-805: t1 . S1 + (4 t2 - 2 t2^T) . X2)."""
        e, be = self.eng, self.be
        o, v = e.no, e.nv
        be.bind_stream()
        e.set_fock(e.F0)
        for x in (self.X2, self.S1, self.S2, self.Dvv, self.Doo, self.Dov, self.Goovv, self.Gooov, self.Gs):
            x.zero_()
        be.pair_combine(self.t2x, e.t2, 4.0, -2.0, None, None, 0.0, o, v)
        self.tri.refresh()
        # several GPUs: the o^2 occupied pairs are dealt round-robin (equal work each), every accumulator is a sum
        # over pairs and is completed by one allreduce
        c = e.comm
        for (i, j) in [(i, j) for i in range(o) for j in range(o)][c.rank::c.world]:
            dis, con = self._plans_for(i, j)
            self.tri._launch_pair(i, j, nk=o)
            for L in dis:
                be.contract(L.desc)
            be.t3_density_combine(self.tri.W, self.Dd, o, 0, i, j, e.eps, o, v, self.buf)
            for L in con:
                be.contract(L.desc)
        if c.world > 1:
            for x in (self.X2, self.S1, self.S2, self.Dvv, self.Doo, self.Dov, self.Goovv, self.Gooov, self.Gs):
                c.allreduce_(x)
        be.pair_symmetrize(self.S2, o, v)                       # :780
        return be.dot(e.t1, self.S1) + be.dot(self.t2x, self.X2)
