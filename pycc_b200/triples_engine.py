"""
(T) engine: Lee-Rendell closed-shell perturbative triples (pycc/cctriples.py:106-152, t_tjl)
without ever holding an o^3 v^3 tensor.

For every occupied pair (i >= j) the connected W^{ijk}_{abc} of ALL k <= j are built in one
batched pass of three ``pccb_contract`` launches (the twelve einsums of t3c_ijk,
cctriples.py:10-22, K-concatenated by their output structure):

    N = c :  W[abc]  = Vi[bae] tkj[ce] + Vj[abe] tki[ce] - ti[mab] Ojk[mc] - tTj[mab] Oik[mc]
    N = b :  W[abc] += Vi[cae] tjk[be] + Vk[ace] tji[be] - ti[mac] Okj[mb] - tTk[mac] Oij[mb]
    N = a :  W[abc] += Vk[bce] tij[ae] + Vj[cbe] tik[ae] - tTk[mbc] Oji[ma] - tj[mbc] Oki[ma]

with Vi[x,y,e] = <xy|ei> = ovvv[i,x,y,e] (8-fold symmetry of the stored <mb|ef> block),
O_jk[m,c] = <mc|jk> = ooov[j,k,m,c], t_kj = t2[k,j], ti = t2[i] and tT = t2 with its virtual
pair swapped (so every operand tile is read along a contiguous direction).  The batch dimension
of the launch is k; the per-(i,j,k) operand addresses are affine in (i,j,k), so the three
descriptors are planned once and only their base pointers are patched per (i,j).
The W batch ([o][v][v][v], 8*o*v^3 bytes) is then consumed by ``pccb_triples_energy``, which
adds the disconnected part and reduces to one energy per triplet.
"""
import ctypes as C

import torch

from . import _lib
from .dsl import TRef, plan
from .parallel import deal_pairs

_STATEMENTS = (
    "W[abc] = Vi[bae]*tkj[ce] + Vj[abe]*tki[ce] - ti[mab]*Ojk[mc] - tTj[mab]*Oik[mc]",
    "W[abc] += Vi[cae]*tjk[be] + Vk[ace]*tji[be] - ti[mac]*Okj[mb] - tTk[mac]*Oij[mb]",
    "W[abc] += Vk[bce]*tij[ae] + Vj[cbe]*tik[ae] - tTk[mbc]*Oji[ma] - tj[mbc]*Oki[ma]",
)

# FP64: the same twelve terms over PLAIN K-contiguous matrices, so that all three launches run on the TMA-staged
# kernel (batched 3-D tensor maps) instead of the cp.async kernels, whose four producer warps of compound-index
# address arithmetic hold the 80-row tile at 81 % tensor-pipe activity (profiles/r02_ncu_triples_wse_o30v250.txt).
# Four helper copies, made once per (T) call (once per iteration in CC3, where V and O are dressed):
#   VT[i][y][x][e] = V[i][x][y][e]            tM[i][a][b][m]  = t2[i][m][a][b]
#   OT[j][k][c][m] = O[j][k][m][c]            tMT[j][a][b][m] = t2[j][m][b][a]
_STATEMENTS_PLAIN = (
    "W[abc] = VTi[abe]*tkj[ce] + Vj[abe]*tki[ce] - tMi[abm]*OTjk[cm] - tMTj[abm]*OTik[cm]",
    "W[abc] += VTi[ace]*tjk[be] + Vk[ace]*tji[be] - tMi[acm]*OTkj[bm] - tMTk[acm]*OTij[bm]",
    "W[abc] += Vk[bce]*tij[ae] + VTj[bce]*tik[ae] - tMTk[bcm]*OTji[am] - tMj[bcm]*OTki[am]",
)


def _ptr_fields(desc):
    """(getter, setter) pairs for every device pointer of a descriptor."""
    out = []
    for q in range(desc.nseg):
        out.append(("seg%dA" % q, q, "A_dev"))
        out.append(("seg%dB" % q, q, "B_dev"))
    out.append(("C", None, "C_dev"))
    return out


def _get(desc, f):
    _, q, name = f
    return getattr(desc.seg[q], name) if q is not None else getattr(desc, name)


def _set(desc, f, val):
    _, q, name = f
    if q is not None:
        setattr(desc.seg[q], name, val)
    else:
        setattr(desc, name, val)


class TripletBatch(object):
    """Statements written for ONE occupied triplet (i, j, k) over slices of stored tensors, launched for a pair
    (i, j) with k as the batch index of the launch: every operand address is affine in (i, j, k), so each
    statement is planned once at (0, 0, 0), the pointer derivatives come from the plans at the three unit steps,
    and a launch only patches base pointers and batch strides.  ``tensors_fn(i, j, k) -> {name: tensor}``."""

    def __init__(self, no, dcode, statements, tensors_fn, backend, prefer_flat=None):
        self.be = backend
        step = 1 if no > 1 else 0
        mk = lambda i, j, k: {n: TRef.of(t, n) for n, t in tensors_fn(i, j, k).items()}
        # a statement may plan into several launches (more than four product terms: the plane-expanded complex
        # statements of complex_cc3.py); the launch list has the same structure at every (i, j, k)
        base = [L for s in statements for L in plan(s, mk(0, 0, 0), dcode, prefer_flat=prefer_flat)]
        for L in base:
            if L.desc.batch != 1:
                raise RuntimeError("unexpected triplet-batch plan structure")
        shifted = [[L.desc for s in statements for L in plan(s, mk(*unit), dcode, prefer_flat=prefer_flat)]
                   for unit in ((step, 0, 0), (0, step, 0), (0, 0, step))]
        if any(len(x) != len(base) or any(a.nseg != b.desc.nseg for a, b in zip(x, base)) for x in shifted):
            raise RuntimeError("triplet-batch plans differ between triplets")
        self.templates = []
        for n, L in enumerate(base):
            d0 = L.desc
            fields = _ptr_fields(d0)
            deriv = {f[0]: tuple((_get(shifted[ax][n], f) - _get(d0, f)) if step else 0 for ax in range(3))
                     for f in fields}
            self.templates.append((L, fields, deriv))
        self._keep = base
        self._cache = {}
        self.itemsize = tensors_fn(0, 0, 0)[next(iter(tensors_fn(0, 0, 0)))].element_size()

    def launch_pair(self, i, j, nk):
        key = (i, j, nk)
        cached = self._cache.get(key)
        if cached is not None:           # the patched descriptors of a pair are reused by every later iteration
            for d in cached:
                self.be.contract(d)
            return
        made = []
        for (L, fields, deriv) in self.templates:
            d = _lib.Contraction()
            C.memmove(C.byref(d), C.byref(L.desc), C.sizeof(d))
            for f in fields:
                di, dj, dk = deriv[f[0]]
                _set(d, f, _get(L.desc, f) + i * di + j * dj)
                name, q, _ = f
                if q is None:
                    d.c_batch = dk // self.itemsize
                elif name.endswith("A"):
                    d.seg[q].a_batch = dk // self.itemsize
                else:
                    d.seg[q].b_batch = dk // self.itemsize
            d.batch = nk
            self.be.contract(d)
            made.append(d)
        if len(self._cache) < 4096:
            self._cache[key] = made


class TriplesEngine(object):
    def __init__(self, ccsd_engine, V=None, O=None):
        """``V`` [i][x][y][e] = Wvvvo[x,y,e,i] and ``O`` [j][k][m][c] = Wovoo[m,c,j,k] default to the bare
        integrals (views of the stored <mb|ef> / <mn|ie> blocks by the 8-fold symmetry); CC3 passes its
        T1-dressed intermediates in the same layouts (pycc/ccwfn.py:637-710)."""
        self.eng = eng = ccsd_engine
        self.be = eng.be
        o, v = eng.no, eng.nv
        if V is None and eng.ovvv_full is None:
            raise ValueError("(T) needs the full <mb|ef> block on every rank (only this rank's slice was given)")
        self.V = V if V is not None else eng.ovvv_full
        self.O = O if O is not None else eng.blk["ooov"]
        import os
        self.plain = eng.dtype == torch.float64 and os.environ.get("PCCB_T_PLAIN", "1") != "0"
        self.statements = _STATEMENTS_PLAIN if self.plain else _STATEMENTS
        if self.plain:
            self.VT = self.be.empty((o, v, v, v), eng.dtype)
            self.OT = self.be.empty((o, o, v, o), eng.dtype)
            self.tM = self.be.empty((o, v, v, o), eng.dtype)
            self.tMT = self.be.empty((o, v, v, o), eng.dtype)
            self.t2T = None
        else:
            self.t2T = self.be.empty((o, o, v, v), eng.dtype)
        self.W = self.be.empty((o, v, v, v), eng.dtype)          # batch over k
        self._plan()

    def _tensors(self, i, j, k):
        e = self.eng
        t2, V, O = e.t2, self.V, self.O
        d = dict(W=self.W[k], Vi=V[i], Vj=V[j], Vk=V[k],
                 tij=t2[i, j], tik=t2[i, k], tji=t2[j, i], tjk=t2[j, k], tki=t2[k, i], tkj=t2[k, j])
        if self.plain:
            VT, OT, tM, tMT = self.VT, self.OT, self.tM, self.tMT
            d.update(VTi=VT[i], VTj=VT[j], tMi=tM[i], tMj=tM[j], tMTj=tMT[j], tMTk=tMT[k],
                     OTij=OT[i, j], OTik=OT[i, k], OTji=OT[j, i], OTjk=OT[j, k], OTki=OT[k, i], OTkj=OT[k, j])
        else:
            tT = self.t2T
            d.update(ti=t2[i], tj=t2[j], tTj=tT[j], tTk=tT[k],
                     Oij=O[i, j], Oik=O[i, k], Oji=O[j, i], Ojk=O[j, k], Oki=O[k, i], Okj=O[k, j])
        return {n: TRef.of(t, n) for n, t in d.items()}

    def _plan(self):
        """Plan at (0,0,0) and difference the pointers at unit steps of i, j, k."""
        o = self.eng.no
        dcode = self.eng.dcode
        stm = self.statements
        pf = True if self.plain else None
        base = [plan(s, self._tensors(0, 0, 0), dcode, prefer_flat=pf) for s in stm]
        for L in base:
            if len(L) != 1 or L[0].desc.batch != 1:
                raise RuntimeError("unexpected (T) plan structure")
        self.templates = []
        step = 1 if o > 1 else 0
        shifted = []
        for unit in ((step, 0, 0), (0, step, 0), (0, 0, step)):
            shifted.append([plan(s, self._tensors(*unit), dcode, prefer_flat=pf)[0].desc for s in stm])
        for n, L in enumerate(base):
            d0 = L[0].desc
            fields = _ptr_fields(d0)
            deriv = {}
            for f in fields:
                p0 = _get(d0, f)
                deriv[f[0]] = tuple((_get(shifted[ax][n], f) - p0) if step else 0 for ax in range(3))
            self.templates.append((L[0], fields, deriv))
        self._keep = base
        self.itemsize = self.eng.t2.element_size()

    def refresh(self):
        """The amplitude- (and, for CC3, intermediate-) dependent helper layouts of the W build."""
        e, be = self.eng, self.be
        o, v = e.no, e.nv
        if self.plain:
            be.permute4(self.VT, self.V.permute(0, 2, 1, 3))
            be.permute4(self.OT, self.O.permute(0, 1, 3, 2))
            be.permute4(self.tM, e.t2.permute(0, 2, 3, 1))
            be.permute4(self.tMT, e.t2.permute(0, 3, 2, 1))
        else:
            be.pair_combine(self.t2T, e.t2, 0.0, 1.0, None, None, 0.0, o, v)

    refresh_t2T = refresh

    def _launch_pair(self, i, j, nk=None):
        """W[k] for k = 0..nk-1 (default: 0..j) of the pair (i, j): three batched contractions."""
        key = (i, j, nk)
        cache = self.__dict__.setdefault("_desc_cache", {})
        cached = cache.get(key)
        if cached is not None:
            for d in cached:
                self.be.contract(d)
            return
        made = []
        for (L, fields, deriv) in self.templates:
            d = _lib.Contraction()
            C.memmove(C.byref(d), C.byref(L.desc), C.sizeof(d))
            for f in fields:
                di, dj, dk = deriv[f[0]]
                _set(d, f, _get(L.desc, f) + i * di + j * dj)
                name, q, _ = f
                if q is None:
                    d.c_batch = dk // self.itemsize
                elif name.endswith("A"):
                    d.seg[q].a_batch = dk // self.itemsize
                else:
                    d.seg[q].b_batch = dk // self.itemsize
            d.batch = (j + 1) if nk is None else nk
            self.be.contract(d)
            made.append(d)
        if len(cache) < 4096:
            cache[key] = made

    def pairs(self):
        """All occupied pairs (i >= j)."""
        o = self.eng.no
        return [(i, j) for i in range(o) for j in range(i + 1)]

    def my_pairs(self):
        """The pairs this rank evaluates (all of them on one GPU)."""
        c = self.eng.comm
        return deal_pairs(self.eng.no, c.rank, c.world)

    def energy_all_ranks(self):
        """E(T): every rank sums its own pairs, then one allreduce of a single double."""
        et, _, _ = self.energy(self.my_pairs())
        return self.eng.comm.allreduce_scalar(et)

    def energy(self, pairs=None):
        """E(T) summed over the triplets (i, j, k<=j) of the given occupied pairs (default: all).
        Returns (total, per-triplet energies as a CPU tensor, triplet list)."""
        eng, be = self.eng, self.be
        o, v = eng.no, eng.nv
        be.bind_stream()
        if pairs is None:
            pairs = self.pairs()
        trip = [(i, j, k) for (i, j) in pairs for k in range(j + 1)]
        if not trip:
            return 0.0, torch.zeros(0, dtype=torch.float64), trip
        ijk = torch.tensor(trip, dtype=torch.int32).reshape(-1).to(be.device)
        et = be.zeros((len(trip),), torch.float64)
        self.refresh()
        fov = eng.F[:o, o:]
        pos = 0
        from .parallel import nvtx_range
        for (i, j) in pairs:
            n = j + 1
            with nvtx_range("(T)/W %d,%d" % (i, j)):
                self._launch_pair(i, j)
            be.triples_energy(self.W, ijk[3 * pos:], n, eng.t1, eng.t2, eng.blk["oovv"], fov, eng.F.stride(0),
                              eng.eps, o, v, et[pos:])
            pos += n
        be.synchronize()
        et_host = et.cpu()
        total = 0.0
        for x in et_host.tolist():      # fixed summation order
            total += x
        return total, et_host, trip
