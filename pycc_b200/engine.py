"""
CCSD residual engine: every builder of ``ccwfn.residuals`` (pycc/ccwfn.py:331-366) recast as
fused strided FP64 tensor-core contractions (``pccb_contract``) plus a handful of memory-bound
kernels, planned once per (o, v) and replayed every iteration.

Data layout in HBM (all C-contiguous, FP64 unless precision='SP'):
    constants   F (n,n), eps (n)                      Fock matrix and its diagonal
                oooo[mnij] ooov[mnie] oovv[ijab] ovov[mbje] ovvv[mbef] vvvv[abef]
                Loovv[mnef] = 2<mn|ef> - <mn|fe>      (pycc/hamiltonian.py:39, oovv block only)
                Looov[mnie] = 2<mn|ie> - <nm|ie>
    amplitudes  t1[ia], t2[ijab]
    per iteration (never bigger than o^2 v^2):
                tau = t2 + t1 t1, tauh = t2 + t1 t1/2, tauW = t2/2 + t1 t1, t2s = 2 t2 - t2^T
                Fae Fmi Fme Wmnij, and in "ring layout" [j,b,m,e]:
                Wr1[jbme] = Wmbej[mbej]   Wr2[jbme] = Wmbje[mbje]   Zr[mijb] = Zmbij[mbij]
The ring layout makes the five o^3 v^3 contractions of r_T2 (ccwfn.py:609-614) plain
(ov) x (ov) GEMMs whose K index (m,e) is contiguous in the W operand; the reference's other
slices of ERI / L (oovo, ovvo, vvvo, ovoo, L_ovvv, L_ovvo, L_oovo) are *views* of the six stored
blocks by the 8-fold symmetry and are never materialised.

Differences from the reference's operation order (results agree to rounding):
  * (t2 - t2^T) Wmbej + t2 (Wmbej + Wmbje^T)  is evaluated as  t2s Wmbej + t2 Wmbje^T;
  * the two o^3 v^3 terms built from tmp = t1 t1 (ccwfn.py:612-614) are factorised through
    o^3 v intermediates (2 o^3 v^2 flops each instead of 2 o^3 v^3);
  * -1/2 t2 (t1 Fme) terms (ccwfn.py:601-605) are folded into dressed Fae / Fmi matrices.
"""
import torch

from . import _lib
from .dsl import TRef, plan
from .parallel import Comm, virtual_slice, nvtx_range


def _unregister_lo(be, ptrs):
    try:
        if getattr(be, "h", None):
            for p in ptrs:
                be.unregister_lo_plane(p)
    except Exception:
        pass


class CCSDEngine(object):
    def __init__(self, backend, F, blocks, no, nv, dtype=torch.float64, comm=None, keep_full_ovvv=True,
                 model='CCSD'):
        """``comm``: pycc_b200.parallel.Comm (None = one GPU).  With more than one rank this rank
        works on the virtual slice [vlo, vhi); ``blocks['vvvv']`` may then be either the full block
        or just the rank's rows ``<ab|ef>, a in slice`` (shape (vhi-vlo, v, v, v)), and likewise
        ``blocks['ovvv']`` (full, or (o, vhi-vlo, v, v)); a full ovvv is needed for (T)."""
        self.be = be = backend
        self.model = 'CCSD' if model == 'CCSD(T)' else model
        if self.model not in ('CCD', 'CC2', 'CCSD', 'CC3'):
            raise ValueError("engine models: CCD, CC2, CCSD, CCSD(T), CC3 (got %r)" % (model,))
        self.no, self.nv, self.dtype = int(no), int(nv), dtype
        self.dcode = _lib.F64 if dtype == torch.float64 else _lib.F32
        self.comm = comm if comm is not None else Comm(None)
        o, v = self.no, self.nv
        n = o + v
        self.vlo, self.vhi = virtual_slice(v, self.comm.rank, self.comm.world)
        vl = self.vl = self.vhi - self.vlo
        # F0: the Hamiltonian's own Fock matrix (ccwfn.H.F); F: the one the kernels read, which
        # residuals(F, ...) may replace for one call (a field-dressed F in real-time propagation)
        self.F0 = be.to_device(F, dtype)
        if tuple(self.F0.shape) != (n, n):
            raise ValueError("F must be (%d,%d)" % (n, n))
        self.F = self.F0.clone()
        self.eps = torch.diagonal(self.F).contiguous()
        shapes = dict(oooo=(o, o, o, o), ooov=(o, o, o, v), oovv=(o, o, v, v), ovov=(o, v, o, v),
                      ovvv=(o, v, v, v), vvvv=(v, v, v, v))
        local = dict(ovvv=(o, vl, v, v), vvvv=(vl, v, v, v))
        self.blk = {}
        for k, shp in shapes.items():
            if k not in blocks:
                raise ValueError("missing ERI block %r" % k)
            src = blocks[k]
            if tuple(src.shape) == shp and k in local and self.comm.world > 1 and not (k == "ovvv" and keep_full_ovvv):
                # keep only this rank's rows on the device
                src = src[self.vlo:self.vhi] if k == "vvvv" else src[:, self.vlo:self.vhi]
            t = be.to_device(src, dtype)
            if tuple(t.shape) != shp and not (k in local and tuple(t.shape) == local[k]):
                raise ValueError("ERI block %s has shape %s, expected %s" % (k, tuple(t.shape), shp))
            self.blk[k] = t
        # this rank's rows of the two big blocks (views when the full block is resident)
        self.vvvv_loc = self.blk["vvvv"]
        self.ovvv_full = self.blk["ovvv"] if self.blk["ovvv"].shape[1] == v else None
        self.ovvv_loc = self.blk["ovvv"][:, self.vlo:self.vhi] if self.ovvv_full is not None else self.blk["ovvv"]
        # spin-adapted constants
        self.Loovv = be.empty((o, o, v, v), dtype)
        be.pair_combine(self.Loovv, self.blk["oovv"], 2.0, -1.0, None, None, 0.0, o, v)
        self.Looov = be.empty((o, o, o, v), dtype)
        be.permute4(self.Looov, self.blk["ooov"], 2.0, 0.0)
        be.permute4(self.Looov, self.blk["ooov"].permute(1, 0, 2, 3), -1.0, 1.0)
        # several GPUs: a rank's largest outputs (o^2 x vl v) are a few waves of long tiles; the tail-aware split-K of
        # pccb_contract needs up to 4 partial copies of one in the workspace (256 MB by default)
        if self.comm.world > 1 and hasattr(be, "ensure_workspace"):
            be.ensure_workspace(min(2 << 30, 4 * o * o * vl * v * (8 if dtype == torch.float64 else 4)))
        # amplitudes and per-iteration buffers (intermediates carrying the sliced index are local-sized)
        E = lambda *s: be.zeros(s, dtype)
        # Several GPUs: the amplitude update is split over the ranks by (i,j) blocks of t2 - rank g owns the
        # blocks [ij_lo, ij_hi), chunk_pairs = ceil(o^2 / world) of them - and the updated / extrapolated
        # chunks are allgathered IN PLACE: t2 is a view of a buffer padded to world * chunk_pairs blocks.
        W = self.comm.world
        self.chunk_pairs = -(-(o * o) // W)
        self.ij_lo = min(o * o, self.comm.rank * self.chunk_pairs)
        self.ij_hi = min(o * o, self.ij_lo + self.chunk_pairs)
        self._t2_store = be.zeros((W * self.chunk_pairs * v * v,), dtype)
        self.t1, self.t2 = E(o, v), self._t2_store[:o * o * v * v].view(o, o, v, v)
        self.t2_stale = False          # True between a split update and the allgather that completes t2
        self.tau, self.tauh, self.tauW, self.t2s = E(o, o, v, v), E(o, o, v, v), E(o, o, v, v), E(o, o, v, v)
        self._small = E(o * o + o ** 4)                      # [Fmi | Wmnij]: summed over the ranks in one allreduce
        self.Fae, self.Fmi, self.Fme = E(vl, v), self._small[:o * o].view(o, o), E(o, v)
        self.Fae2, self.Fmi2 = E(vl, v), E(o, o)
        self.Wmnij = self._small[o * o:].view(o, o, o, o)
        self.Wr1, self.Wr2 = E(o, vl, o, v), E(o, vl, o, v)      # [j,b,m,e]
        self.Zr = E(o, o, o, vl)                                 # [m,i,j,b]
        self.Y1, self.Y2 = E(o, o, o, vl), E(o, o, vl, o)        # [i,m,j,b], [i,m,a,j]
        # r2 is padded like t2: the iteration's exchange step is an in-place reduce-scatter over the same
        # (i,j)-block chunks, which leaves on each rank the finished residual of the blocks it updates
        self._r2_store = be.zeros((W * self.chunk_pairs * v * v,), dtype)
        self.r1, self.r2 = E(o, v), self._r2_store[:o * o * v * v].view(o, o, v, v)
        self._scatter_r2 = False       # set by residuals(symmetrize=False) on several GPUs
        self.r2_scattered = False      # True: r2 holds the symmetrised, summed residual of this rank's blocks only
        if self.model == 'CC2':      # t1 t1 products and the two factorised intermediates of the CC2 r_T2
            self.T11, self.Yw, self.Xv = E(o, o, v, v), E(o, o, o, vl), E(vl, v, o, v)
        # Mixed-precision mode.  (i) The ring-term operands are kept in PLAIN-MATRIX layouts as well
        # (permuted o^2 v^2 copies: constants once, amplitudes every iteration), so that the five
        # o^3 v^3 contractions are K-contiguous 2-D GEMMs like the ladder.  (ii) Every operand of those
        # GEMMs has a residual plane lo = x - trunc_tf32(x); the plain-matrix contractions then run as
        # 3xTF32 on the tcgen05 tensor cores (csrc/contract_tc05.cu), everything else on the mma.sync kernel.
        import os
        # (FP64: the same plain-matrix copies make the ring terms eligible for the TMA-staged kernel:
        #  28.35 -> 28.01 ms at (20,150), 0.4295 -> 0.4209 s at (30,250); PCCB_PLAIN_RING=0 disables)
        self.sp_plain = self.model != 'CC2' and (
            (dtype == torch.float32 and os.environ.get("PCCB_SP_PLAIN", "1") != "0") or
            (dtype == torch.float64 and os.environ.get("PCCB_PLAIN_RING", "1") != "0"))
        self._lo = None
        if self.sp_plain:
            self.oovvP, self.oovvX, self.LoovvP = E(o, v, o, v), E(o, v, o, v), E(o, v, o, v)      # [m,e,n,f]
            be.permute4(self.oovvP, self.blk["oovv"].permute(0, 2, 1, 3))       # <mn|ef>
            be.permute4(self.oovvX, self.blk["oovv"].permute(0, 3, 1, 2))       # <mn|fe>
            be.permute4(self.LoovvP, self.Loovv.permute(0, 2, 1, 3))
            self.t2sP, self.t2P, self.t2Q = E(o, v, o, v), E(o, v, o, v), E(o, v, o, v)            # [i,a,m,e]
            self.tauWPn, self.t2Rh = E(o, vl, o, v), E(o, vl, o, v)                                # [j,B,n,f]
        # tcgen05 mode only: K-contiguous copies for the o^4 v^2 term of r_T2, tauT[Abmn] = tau[mnAb] and
        # WmnijT[ijmn] = Wmnij[mnij] (12.1 -> 2.x ms at (40,300)); t1, t2 and Fae2 get residual planes so that
        # t1[ie] <ja|be> and t2[ijae] Fae2[be] run there as well
        self.sp_tc = self.sp_plain and dtype == torch.float32 and hasattr(be, "split_lo") and \
            os.environ.get("PCCB_SP_TC_EXTRA", "1") != "0"
        if self.sp_tc:
            self.tauT, self.WmnijT = E(vl, v, o, o), E(o, o, o, o)
        if dtype == torch.float32 and hasattr(be, "split_lo"):
            self._lo = {}
            const = [("vvvv", self.vvvv_loc), ("ovvv", self.blk["ovvv"]), ("oovv", self.blk["oovv"])]
            if self.sp_plain:
                const += [("oovvP", self.oovvP), ("oovvX", self.oovvX), ("LoovvP", self.LoovvP)]
            for name, x in const:
                lo = be.empty(tuple(x.shape), dtype)
                be.split_lo(x, lo)
                be.register_lo_plane(x, lo)
                self._lo[name] = lo
            varying = ["tau"] + (["t2sP", "t2P", "t2Q", "tauWPn", "t2Rh", "Wr1", "Wr2"] if self.sp_plain else [])
            if self.sp_tc:
                varying += ["tauT", "WmnijT", "t1", "t2", "Fae2"]
            for name in varying:
                x = getattr(self, name)
                self._lo[name] = be.empty(tuple(x.shape), dtype)
                # t1 / t2 are also written OUTSIDE prepare() (update, DIIS, set_amplitudes), after which their planes
                # are stale until the next prepare(): they are registered only while r_T2 runs (_run), so that no
                # other engine's launch over t1 / t2 (HBAR, (T), CC3, Lambda) can pick up a stale plane
                if name not in ("t1", "t2"):
                    be.register_lo_plane(x, self._lo[name])
        if self._lo is not None:
            # the registry maps raw address ranges to lo planes: drop this engine's ranges when it dies, so
            # that memory the caching allocator recycles can never match a stale entry
            import weakref
            ranges = [x.data_ptr() for x in
                      [self.vvvv_loc, self.blk["ovvv"], self.blk["oovv"]] + [getattr(self, n) for n in self._lo
                                                                            if hasattr(self, n)]]
            weakref.finalize(self, _unregister_lo, be, sorted(set(ranges)))
        self.cc3 = None
        self.ladder_events = None    # bench.py: list that receives (start, end) CUDA events per ladder launch
        self._build_plans()

    # ---------------------------------------------------------------------------------------
    def _tensors(self, names):
        """Name -> strided view.  ``X_sK`` is tensor X restricted to this rank's virtual slice
        along dimension K (the whole tensor on one GPU)."""
        o = self.no
        F = self.F
        d = dict(oooo=self.blk["oooo"], ooov=self.blk["ooov"], oovv=self.blk["oovv"], ovov=self.blk["ovov"])
        d.update(Foo=F[:o, :o], Fov=F[:o, o:], Fvv=F[o:, o:], Loovv=self.Loovv, Looov=self.Looov,
                 t1=self.t1, t2=self.t2, tau=self.tau, tauh=self.tauh, tauW=self.tauW, t2s=self.t2s,
                 Fae=self.Fae, Fmi=self.Fmi, Fme=self.Fme, Fae2=self.Fae2, Fmi2=self.Fmi2, Wmnij=self.Wmnij,
                 Wr1=self.Wr1, Wr2=self.Wr2, Zr=self.Zr, Y1=self.Y1, Y2=self.Y2, r1=self.r1, r2=self.r2)
        if self.model == 'CC2':
            d.update(T11=self.T11, Yw=self.Yw, Xv=self.Xv)
        if self.sp_plain:
            d.update(oovvP=self.oovvP, oovvX=self.oovvX, LoovvP=self.LoovvP, t2sP=self.t2sP, t2P=self.t2P, t2Q=self.t2Q,
                     tauWPn=self.tauWPn, t2Rh=self.t2Rh)
        if getattr(self, "sp_tc", False):
            d.update(tauT=self.tauT, WmnijT=self.WmnijT)
        out = {}
        for nm in names:
            if nm == "vvvv_s0":
                t = self.vvvv_loc
            elif nm == "ovvv_s1":
                t = self.ovvv_loc
            elif nm in d:
                t = d[nm]
            else:
                base, dim = nm[:-3], int(nm[-1])
                if nm[-3:-1] != "_s" or base not in d:
                    raise KeyError("unknown tensor %r" % nm)
                t = d[base].narrow(dim, self.vlo, self.vl)
            out[nm] = TRef.of(t, nm)
        return out

    # Statements, grouped per reference builder.  Index letters as in the reference; an upper-case
    # A / B is the virtual index restricted to this rank's slice (extent vl), and ``X_sK`` the
    # matching view of X.  r1 / r2 are zeroed first and every term accumulates, so that the
    # per-rank partial results add up under the allreduce.
    STATEMENTS = {
        # build_Fae (pycc/ccwfn.py:421-428), rows A only; L[o,v,v,v] = 2<ma|fe> - <ma|ef>
        "Fae": ["Fae[Ae] = Fvv_s0[Ae] - 0.5*t1_s1[mA]*Fov[me] - tauh_s2[mnAf]*Loovv[mnef]",
                "Fae[Ae] += 2.0*t1[mf]*ovvv_s1[mAfe] - t1[mf]*ovvv_s1[mAef]"],
        # build_Fmi (:441-447); L[o,o,o,v] = Looov
        "Fmi": ["Fmi[mi] = Foo[mi] + 0.5*Fov[me]*t1[ie] + Loovv[mnef]*tauh[inef]",
                "Fmi[mi] += t1[ne]*Looov[mnie]"],
        # build_Fme (:456-460)
        "Fme": ["Fme[me] = Fov[me] + Loovv[mnef]*t1[nf]"],
        # build_Wmnij (:473-482); <mn|ej> = <nm|je>
        "Wmnij": ["Wmnij[mnij] = oooo[mnij] + oovv[mnef]*tau[ijef]",
                  "Wmnij[mnij] += ooov[mnie]*t1[je]",
                  "Wmnij[mnij] += ooov[nmje]*t1[ie]"],
        # build_Wmbej (:498-505) in ring layout Wr1[jbme], b in slice; <mb|ej> = <mj|eb>, <mn|ej> = <nm|je>
        "Wmbej": ["Wr1[jBme] = oovv_s3[mjeB] - tauW_s3[jnfB]*oovv[mnef] + 0.5*t2_s3[njfB]*Loovv[mnef]",
                  "Wr1[jBme] += t1[jf]*ovvv_s1[mBef]",
                  "Wr1[jBme] -= t1_s1[nB]*ooov[nmje]"],
        # build_Wmbje (:520-526) in ring layout Wr2[jbme] = Wmbje[mbje], b in slice
        "Wmbje": ["Wr2[jBme] = -1.0*ovov_s1[mBje] + tauW_s3[jnfB]*oovv[mnfe]",
                  "Wr2[jBme] -= t1[jf]*ovvv_s1[mBfe]",
                  "Wr2[jBme] += t1_s1[nB]*ooov[mnje]"],
        # build_Zmbij (:537) as Zr[mijb], b in slice
        "Zmbij": ["Zr[mijB] = ovvv_s1[mBef]*tau[ijef]"],
        # r_T1 (:548-557), columns A; L[o,v,v,o][nafi] = 2<ni|fa> - <na|if>; L[o,o,v,o][nmei] = Looov[mnie]
        "r_T1": ["r1_s1[iA] += Fov_s1[iA] + t1[ie]*Fae[Ae] - Fmi[mi]*t1_s1[mA] + t2s[mief]*ovvv_s1[mAef]"
                 " - Looov[mnie]*t2_s2[mnAe]",
                 "r1_s1[iA] += t2s_s2[imAe]*Fme[me]",
                 "r1_s1[iA] += 2.0*t1[nf]*oovv_s3[nifA] - t1[nf]*ovov_s1[nAif]"],
        # r_T2 (:596-616), unsymmetrised; <ab|ej> = <ja|be>, <mb|ij> = <ij|mb>
        "r_T2": ["Fae2[Be] = Fae[Be] - 0.5*t1_s1[mB]*Fme[me]",
                 "Fmi2[mj] = Fmi[mj] + 0.5*Fme[me]*t1[je]",
                 "r2_s2[ijAb] += 0.5*oovv_s2[ijAb] + 0.5*tau[ijef]*vvvv_s0[Abef]",
                 "r2_s2[ijAb] += 0.5*Wmnij[mnij]*tau_s2[mnAb]",
                 "r2_s3[ijaB] += t2s[imae]*Wr1[jBme] + t2[imae]*Wr2[jBme]",
                 "r2_s3[ijaB] += t2[mjae]*Wr2[iBme]",
                 "r2_s3[ijaB] += t2[ijae]*Fae2[Be]",
                 "r2_s2[ijAb] -= Fmi2[mj]*t2_s2[imAb]",
                 "r2_s3[ijaB] -= t1[ma]*Zr[mijB]",
                 "Y1[imjB] = t1[ie]*oovv_s3[mjeB]",
                 "r2_s3[ijaB] -= t1[ma]*Y1[imjB]",
                 "Y2[imAj] = t1[ie]*ovov_s1[mAje]",
                 "r2_s2[ijAb] -= Y2[imAj]*t1[mb]",
                 "r2_s2[ijAb] += t1[ie]*ovvv_s1[jAbe]",
                 "r2_s2[ijAb] -= t1_s1[mA]*ooov[ijmb]"],
    }

    # Several GPUs: the two builders every rank would otherwise compute in full (2 o^4 v^2 + 2 o^3 v^2 flops,
    # 3 ms of a 54 ms rank share at (30,250) x 8) are sliced over the CONTRACTED virtual index E instead; the
    # partial sums share one buffer and one small allreduce, then the cheap unsliced terms are added.
    SLICED_K = {
        "Fmi": (["Fmi[mi] = Loovv_s2[mnEf]*tauh_s2[inEf]"],
                ["Fmi[mi] += Foo[mi] + 0.5*Fov[me]*t1[ie]", "Fmi[mi] += t1[ne]*Looov[mnie]"]),
        "Wmnij": (["Wmnij[mnij] = oovv_s2[mnEf]*tau_s2[ijEf]"],
                  ["Wmnij[mnij] += oooo[mnij] + ooov[mnie]*t1[je]", "Wmnij[mnij] += ooov[nmje]*t1[ie]"]),
    }
    SLICED_K_CC2 = {
        "Wmnij": (["Wmnij[mnij] = oovv_s2[mnEf]*T11_s2[ijEf]"],
                  ["Wmnij[mnij] += oooo[mnij] + ooov[mnie]*t1[je]", "Wmnij[mnij] += ooov[nmje]*t1[ie]"]),
    }

    # CC2 branch of the same builders (pycc/ccwfn.py:478-479, 495-496, 517-518, 534-535, 575-592):
    # Fae / Fmi / Fme / r_T1 as in CCSD; Wmnij and Zmbij with t1 t1 in place of tau; no Wmbej / Wmbje;
    # r_T2 with bare Fock matrices and the doubles-into-doubles terms dropped.  T11[ijef] = t1[ie] t1[jf].
    STATEMENTS_CC2 = {
        "Wmnij": ["Wmnij[mnij] = oooo[mnij] + oovv[mnef]*T11[ijef]",
                  "Wmnij[mnij] += ooov[mnie]*t1[je]",
                  "Wmnij[mnij] += ooov[nmje]*t1[ie]"],
        "Zmbij": ["Zr[mijB] = ovvv_s1[mBef]*T11[ijef]"],
        # (:575-592); the two terms in t2 Fvv-like / Foo-like matrices are merged:
        #   Fae2[be] = Fvv[be] - t1[mb] Fov[me],  Fmi2[mj] = Foo[mj] + Fov[me] t1[je]
        "r_T2": ["Fae2[Be] = Fvv_s0[Be] - t1_s1[mB]*Fov[me]",
                 "Fmi2[mj] = Foo[mj] + Fov[me]*t1[je]",
                 "r2_s2[ijAb] += 0.5*oovv_s2[ijAb]",
                 "r2_s3[ijaB] += t2[ijae]*Fae2[Be]",
                 "r2_s2[ijAb] -= Fmi2[mj]*t2_s2[imAb]",
                 "Yw[mijB] = Wmnij[mnij]*t1_s1[nB]",
                 "r2_s3[ijaB] += 0.5*t1[ma]*Yw[mijB]",
                 "Xv[Abif] = t1[ie]*vvvv_s0[Abef]",
                 "r2_s2[ijAb] += 0.5*Xv[Abif]*t1[jf]",
                 "r2_s3[ijaB] -= t1[ma]*Zr[mijB]",
                 "Y1[imjB] = t1[ie]*oovv_s3[mjeB]",
                 "r2_s3[ijaB] -= t1[ma]*Y1[imjB]",
                 "Y2[imAj] = t1[ie]*ovov_s1[mAje]",
                 "r2_s2[ijAb] -= Y2[imAj]*t1[mb]",
                 "r2_s2[ijAb] += t1[ie]*ovvv_s1[jAbe]",
                 "r2_s2[ijAb] -= t1_s1[mA]*ooov[ijmb]"],
    }

    # Mixed-precision mode: the ring terms over plain-matrix copies (see __init__), same algebra:
    #   oovvP[menf] = <mn|ef>, oovvX[menf] = <mn|fe>, LoovvP[menf] = Loovv[mnef],
    #   t2sP[iame] = t2s[imae], t2P[iame] = t2[imae], t2Q[jame] = t2[mjae],
    #   tauWPn[jBnf] = -tauW[jnfB], t2Rh[jBnf] = 0.5 t2[njfB]   (factors folded in: one scale per launch)
    STATEMENTS_SP = {
        "Wmbej": ["Wr1[jBme] = oovv_s3[mjeB] + tauWPn[jBnf]*oovvP[menf] + t2Rh[jBnf]*LoovvP[menf]",
                  "Wr1[jBme] += t1[jf]*ovvv_s1[mBef]",
                  "Wr1[jBme] -= t1_s1[nB]*ooov[nmje]"],
        "Wmbje": ["Wr2[jBme] = -1.0*ovov_s1[mBje] - tauWPn[jBnf]*oovvX[menf]",
                  "Wr2[jBme] -= t1[jf]*ovvv_s1[mBfe]",
                  "Wr2[jBme] += t1_s1[nB]*ooov[mnje]"],
    }
    R_T2_SP = {"r2_s3[ijaB] += t2s[imae]*Wr1[jBme] + t2[imae]*Wr2[jBme]": "r2_s3[ijaB] += t2sP[iame]*Wr1[jBme] + t2P[iame]*Wr2[jBme]",
               "r2_s3[ijaB] += t2[mjae]*Wr2[iBme]": "r2_s3[ijaB] += t2Q[jame]*Wr2[iBme]"}
    R_T2_TC = {"r2_s2[ijAb] += 0.5*Wmnij[mnij]*tau_s2[mnAb]": "r2_s2[ijAb] += 0.5*WmnijT[ijmn]*tauT[Abmn]"}

    def builder_order(self):
        """Builders run by residuals(), in the reference's order (pycc/ccwfn.py:357-366).
        CCD (:415-420, 434-439, 466-471, 488-496, 511-518, 563-573) is the CCSD algebra with
        t1 = 0 and r1 = 0; its t1-carrying launches then contribute exact zeros."""
        if self.model == 'CC2':
            return ("Fae", "Fmi", "Fme", "Wmnij", "Zmbij", "r_T1", "r_T2")
        if self.model == 'CCD':
            return ("Fae", "Fmi", "Fme", "Wmnij", "Wmbej", "Wmbje", "Zmbij", "r_T2")
        return ("Fae", "Fmi", "Fme", "Wmnij", "Wmbej", "Wmbje", "Zmbij", "r_T1", "r_T2")

    def _statements(self, plain=None):
        """``plain=False``: the layout-independent statement set (complex_engine expands that one)."""
        own = plain is None
        plain = self.sp_plain if plain is None else plain
        st = dict(self.STATEMENTS)
        if self.model == 'CC2':
            st.update(self.STATEMENTS_CC2)
            del st["Wmbej"], st["Wmbje"]
        elif plain:
            st.update(self.STATEMENTS_SP)
            st["r_T2"] = [self.R_T2_SP.get(x, x) for x in st["r_T2"]]
            if getattr(self, "sp_tc", False) and own:
                st["r_T2"] = [self.R_T2_TC.get(x, x) for x in st["r_T2"]]
        return st

    def _plan_list(self, stmts):
        import re
        launches = []
        for s in stmts:
            T = self._tensors(set(re.findall(r"([A-Za-z_][A-Za-z_0-9]*)\[", s)))
            mine = plan(s, T, self.dcode, prefer_flat=True if self.sp_plain else None)
            for L in mine:
                L.is_ladder = "vvvv_s0" in s
                L.post = None
                launches.append(L)
            if getattr(self, "sp_tc", False) and s.startswith("Fae2[") and mine:
                mine[-1].post = "Fae2"          # _launch: refresh its residual plane right after
        return launches

    def _build_plans(self):
        self.plans = {}
        self.pre_plans = {}         # world > 1: K-sliced partial sums that precede the small allreduce
        for name, stmts in self._statements().items():
            self.plans[name] = self._plan_list(stmts)
        if self.comm.world > 1:
            sk = dict(self.SLICED_K)
            if self.model == 'CC2':
                sk.update(self.SLICED_K_CC2)
            for name, (pre, post) in sk.items():
                self.pre_plans[name] = self._plan_list(pre)
                self.plans[name] = self._plan_list(post)

    def _run_small(self, names):
        """Fmi / Wmnij on several GPUs: partial sums over this rank's slice of the contracted index, ONE
        allreduce of the shared buffer, then the unsliced terms."""
        for name in names:
            for L in self.pre_plans[name]:
                self.be.contract(L.desc)
        self.comm.allreduce_(self._small)
        for name in names:
            self._launch(name)

    def _run(self, name):
        if name in self.pre_plans:
            return self._run_small([name])
        if name == "r_T1":
            self.r1.zero_()
        elif name == "r_T2":
            self.r2.zero_()
            if self._lo is not None and self.sp_plain:      # the ring intermediates are operands now
                self.be.split_lo(self.Wr1, self._lo["Wr1"])
                self.be.split_lo(self.Wr2, self._lo["Wr2"])
            if getattr(self, "sp_tc", False):
                self.be.permute4(self.WmnijT, self.Wmnij.permute(2, 3, 0, 1))
                self.be.split_lo(self.WmnijT, self._lo["WmnijT"])
        amp_planes = name == "r_T2" and getattr(self, "sp_tc", False)
        if amp_planes:                # fresh since this call's prepare()
            self.be.register_lo_plane(self.t1, self._lo["t1"])
            self.be.register_lo_plane(self.t2, self._lo["t2"])
        try:
            with nvtx_range("ccsd/" + name):
                self._launch(name)
        finally:
            if amp_planes:
                self.be.unregister_lo_plane(self.t1.data_ptr())
                self.be.unregister_lo_plane(self.t2.data_ptr())
        # the one exchange step of the iteration: sum the per-rank partial residuals
        if getattr(self, "_defer_reduce", False):
            return                   # CC3 on several GPUs: after the triples terms (residuals())
        if name == "r_T1":
            self.comm.allreduce_(self.r1)
        elif name == "r_T2":
            if self._scatter_r2:
                # r2 + r2^T (ccwfn.py:621) is linear: symmetrise the partial sums locally, then reduce-scatter
                # by (i,j) blocks - half the bytes of an allreduce, and no partner read in the update
                self.be.pair_symmetrize(self.r2, self.no, self.nv)
                n = self.chunk_pairs * self.nv * self.nv
                self.comm.reduce_scatter_(self._r2_store, self.comm.rank * n, n)
                self.r2_scattered = True
            else:
                self.comm.allreduce_(self.r2)
                self.r2_scattered = False

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
            if L.post is not None:
                self.be.split_lo(getattr(self, L.post), self._lo[L.post])

    # ---- amplitude-dependent element-wise preparations ---------------------------------------
    def prepare(self):
        """tau variants, the spin-adapted t2 (build_tau, pycc/ccwfn.py:408-410; :554) and the plain-matrix
        layouts of the ring-term operands: ONE pass over t2 (pccb_prepare_amplitudes)."""
        be, o, v = self.be, self.no, self.nv
        outs = dict(tau=self.tau, tauh=self.tauh, tauW=self.tauW, t2s=self.t2s)
        if self.model == 'CC2':
            outs["t11"] = self.T11
        if self.sp_plain:
            outs.update(t2sP=self.t2sP, t2P=self.t2P, t2Q=self.t2Q, tauWPn=self.tauWPn, t2Rh=self.t2Rh)
        be.prepare_amplitudes(self.t1, self.t2, o, v, self.vlo, self.vl, **outs)
        if self._lo is not None:
            if self.sp_tc:
                be.permute4(self.tauT, self.tau.narrow(2, self.vlo, self.vl).permute(2, 3, 0, 1))
            for name in ["tau"] + (["t2sP", "t2P", "t2Q", "tauWPn", "t2Rh"] if self.sp_plain else []) + \
                    (["tauT", "t1", "t2"] if self.sp_tc else []):
                be.split_lo(getattr(self, name), self._lo[name])

    def set_amplitudes(self, t1, t2):
        self.t1.copy_(torch.as_tensor(t1).to(self.t1.device, self.dtype))
        self.t2.copy_(torch.as_tensor(t2).to(self.t2.device, self.dtype))

    def set_fock(self, F):
        """A different Fock operator for the next residuals() call (ccwfn.residuals takes F)."""
        self.F.copy_(torch.as_tensor(F).to(self.F.device, self.dtype))
        self.eps.copy_(torch.diagonal(self.F))

    def residuals(self, symmetrize=True):
        """r1, r2 for the amplitudes held in self.t1 / self.t2 (pycc/ccwfn.py:357-366)."""
        self.be.bind_stream()
        if self.model == 'CCD':
            self.t1.zero_()          # CCD carries no singles (ccwfn.py:543-546): before tau is formed
            self.r1.zero_()
        self.prepare()
        # the solver's iteration (symmetrize=False) may leave r2 scattered over the ranks; callers that want the
        # whole residual get the allreduce.  CC3 adds its triples terms after the builders: allreduce there.
        self._scatter_r2 = (not symmetrize) and self.comm.world > 1 and self.model != 'CC3'
        self._defer_reduce = self.model == 'CC3' and self.comm.world > 1
        order = list(self.builder_order())
        if self.pre_plans:
            self._run_small([n for n in order if n in self.pre_plans])
            order = [n for n in order if n not in self.pre_plans]
        for name in order:
            self._run(name)
        if self.model == 'CC3':      # triples contributions X1, X2 (pycc/ccwfn.py:371-401)
            if self.cc3 is None:
                from .cc3_engine import CC3Engine
                self.cc3 = CC3Engine(self)
            self.cc3.contribute()
            if self._defer_reduce:       # builders + this rank's pairs of the triples terms: one sum over the ranks
                self.comm.allreduce_(self.r1)
                self.comm.allreduce_(self.r2)
                self.r2_scattered = False
        self._defer_reduce = False
        self._scatter_r2 = False
        if symmetrize:
            self.be.pair_symmetrize(self.r2, self.no, self.nv)       # ccwfn.py:621
        return self.r1, self.r2

    def run_builder(self, name):
        self.be.bind_stream()
        self.prepare()
        deps = {"Fae": [], "Fmi": [], "Fme": [], "Wmnij": [], "Wmbej": [], "Wmbje": [], "Zmbij": [],
                "r_T1": ["Fae", "Fmi", "Fme"],
                "r_T2": [b for b in self.builder_order() if b not in ("r_T1", "r_T2")]}
        for d in deps[name]:
            self._run(d)
        self._run(name)

    # reference-ordered views of the ring-layout intermediates
    def Wmbej_view(self):
        return self.Wr1.permute(2, 1, 3, 0)      # [j,b,m,e] -> [m,b,e,j]

    def Wmbje_view(self):
        return self.Wr2.permute(2, 1, 0, 3)      # [j,b,m,e] -> [m,b,j,e]

    def Zmbij_view(self):
        return self.Zr.permute(0, 3, 1, 2)       # [m,i,j,b] -> [m,b,i,j]

    def energy(self):
        """cc_energy (pycc/ccwfn.py:717-718)."""
        self.be.bind_stream()
        o = self.no
        fov = self.F[:o, o:]
        return self.be.cc_energy(fov, self.F.stride(0), self.t1, self.t2, self.blk["oovv"], self.no, self.nv)

    def update(self, symmetrize_on_the_fly):
        """t += r/D; returns the squared increment norm (pycc/ccwfn.py:287-290)."""
        self.be.bind_stream()
        return self.be.update_amplitudes(self.t1, self.t2, self.r1, self.r2, self.eps, self.no, self.nv,
                                         symmetrize_on_the_fly)

    def update_energy(self):
        """t += r/D (the pair symmetrisation of r2, ccwfn.py:621, on the fly), the increment norm and the
        energy of the new amplitudes in ONE pass (pccb_update_energy).  Several GPUs: each rank updates the
        (i,j) blocks it owns - whose residual the reduce-scatter of residuals(symmetrize=False) delivered
        already summed and symmetrised - (t1: all of it, every rank holds the summed r1), the two scalars are
        summed over the ranks, and t2 stays incomplete until ``sync_amplitudes`` or the DIIS allgather.
        Returns (squared increment norm, energy)."""
        self.be.bind_stream()
        o = self.no
        fov = self.F[:o, o:]
        rms2, e = self.be.update_energy(self.t1, self.t2, self.r1, self.r2, self.eps, fov, self.F.stride(0),
                                        self.blk["oovv"], o, self.nv, not self.r2_scattered, self.ij_lo, self.ij_hi,
                                        self.comm.rank == 0)
        if self.comm.world > 1:
            rms2, e = self.comm.allreduce_scalars([rms2, e])
            self.t2_stale = True
        return rms2, e

    def sync_amplitudes(self):
        """Complete t2 on every rank after a split update (in-place allgather of the ranks' chunks)."""
        if self.t2_stale:
            n = self.chunk_pairs * self.nv * self.nv
            self.comm.allgather_into(self._t2_store, self._t2_store[self.comm.rank * n:(self.comm.rank + 1) * n],
                                     in_place=True)
            self.t2_stale = False

    def flops_issued(self):
        """Flops the planned launches actually issue per residuals() call (2*M*N*K summed)."""
        tot = 0
        for launches in list(self.plans.values()) + list(self.pre_plans.values()):
            for L in launches:
                d = L.desc
                for q in range(d.nseg):
                    tot += 2 * d.M * d.N * d.batch * d.seg[q].K
        return tot
