"""
ORACLE (test infrastructure, not product code): the restatement of ``oracle/ccsd_oracle.py`` /
``oracle/triples_oracle.py`` with ``torch.einsum`` in place of ``numpy.einsum``, term for term.

Why it exists: at the BASELINE shapes (30,250), (40,300), (40,400) the NumPy oracle needs tens of
minutes of host time and a host-resident ``<ab|ef>`` block (31 - 205 GB), so ``bench.py`` could
not attach a parity check to those timings.  This file runs the SAME formulas on whatever device
its input tensors live on: on the GPU box that is cuBLAS DGEMM / eager PyTorch in FP64 - a code
path that shares nothing with ``libpyccb200.so`` - and one full-size ``residuals()`` then takes
about a second.  Sharded ``<ab|ef>`` rows are supported (``vv_rows``), so every rank can check its
own ladder slice.

Import rules as for the rest of ``oracle/``: only ``tests/`` and ``bench.py``'s checker legs.
Parity status: PINNED - ``tests/test_oracle_golden.py::test_torch_oracle_*`` holds it to the
reference-generated golden vectors (and to the NumPy oracle) on the CPU.

Reference lines: pycc/ccwfn.py:408-622, 712-719 (builders, residuals, energy);
pycc/cctriples.py:8-35, 68-83, 106-152 (t3c_ijk, t3d_ijk, body of t_tjl).
"""
import torch


def _t(x, like=None, dtype=torch.float64, device=None):
    if isinstance(x, torch.Tensor):
        return x.to(dtype=dtype, device=device if device is not None else x.device)
    return torch.as_tensor(x, dtype=dtype, device=device)


class Blocks(object):
    """The reference's ERI / L slices through the six stored blocks (see ccsd_oracle.Blocks).

    The two big blocks are only ever touched row by row, so that they may be PROVIDED instead of stored:
    ``ovvv`` is a tensor [o,v,v,v] or a callable m -> <mb|ef> [v,v,v]; ``vvvv`` is a tensor holding the
    rows a in [vv_rows[0], vv_rows[1]) of <ab|ef>, or a callable (lo, hi) -> those rows.  ``chunk`` bounds
    the rows requested at once (bench.py regenerates them from the factor B: SURVEY.md 8(d) allows the
    factor form for exactly this, an independent checker)."""

    def __init__(self, blocks, device=None, vv_rows=None, chunk=None):
        g = lambda k: _t(blocks[k], device=device) if blocks.get(k) is not None else None
        self.oooo, self.ooov, self.oovv, self.ovov = g("oooo"), g("ooov"), g("oovv"), g("ovov")
        self._ovvv = blocks["ovvv"] if callable(blocks.get("ovvv")) else g("ovvv")
        self._vvvv = blocks["vvvv"] if callable(blocks.get("vvvv")) else g("vvvv")
        self.no, self.nv = self.oovv.shape[0], self.oovv.shape[2]
        self.vv_rows = vv_rows if vv_rows is not None else (0, self.nv)
        self.chunk = chunk

    def ovvv_row(self, m):
        return self._ovvv(m) if callable(self._ovvv) else self._ovvv[m]

    def vvvv_chunks(self):
        """Yields (lo, hi, <ab|ef> rows a in [lo, hi)) covering this object's vv_rows."""
        lo0, hi0 = self.vv_rows
        step = self.chunk or (hi0 - lo0)
        for lo in range(lo0, hi0, max(step, 1)):
            hi = min(hi0, lo + step)
            yield lo, hi, (self._vvvv(lo, hi) if callable(self._vvvv) else self._vvvv[lo - lo0:hi - lo0])

    def oovo(self):   # <mn|ej> = <nm|je>
        return self.ooov.permute(1, 0, 3, 2)

    def ovvo(self):   # <mb|ej> = <mj|eb>
        return self.oovv.permute(0, 3, 2, 1)

    def ovoo(self):   # <mb|ij> = <ij|mb>
        return self.ooov.permute(2, 3, 0, 1)

    def L_oovv(self):
        return 2.0 * self.oovv - self.oovv.permute(0, 1, 3, 2)

    def L_ooov(self):
        return 2.0 * self.ooov - self.ooov.permute(1, 0, 2, 3)

    def L_ovvo(self):
        return 2.0 * self.ovvo() - self.ovov.permute(0, 1, 3, 2)

    def L_oovo(self):
        return 2.0 * self.oovo() - self.ooov.permute(0, 1, 3, 2)


_ein = torch.einsum


def tau(t1, t2, f1=1.0, f2=1.0):
    """build_tau (pycc/ccwfn.py:408-410)."""
    return f1 * t2 + f2 * t1[:, None, :, None] * t1[None, :, None, :]


def _ladder(blk, tt, out):
    """out[:, :, a, :] += 1/2 sum_ef tau[ijef] <ab|ef> for the rows a this object holds (ccwfn.py:607)."""
    for lo, hi, vv in blk.vvvv_chunks():
        out[:, :, lo:hi, :] += 0.5 * _ein('ijef,abef->ijab', tt, vv)
    return out


def residuals(F, blk, t1, t2, ladder_only=False, skip_ladder=False):
    """ccwfn.residuals, CCSD branch (pycc/ccwfn.py:357-366, 413-622), symmetrised r2 included.
    ``ladder_only``: just the (unsymmetrised) o^2 v^4 term for the <ab|ef> rows this object holds, in a
    full-size zero array (the rank-local piece of a sharded check); ``skip_ladder``: everything else.
    Every <mb|ef> term is evaluated row by row over m (the terms are sums over m, or carry m as an
    output index), which is the reference's contraction split into o independent pieces."""
    F = _t(F, device=blk.oovv.device)
    t1, t2 = _t(t1, device=F.device), _t(t2, device=F.device)
    no = blk.no
    Foo, Fov, Fvv = F[:no, :no], F[:no, no:], F[no:, no:]
    tt = tau(t1, t2)
    if ladder_only:
        return None, _ladder(blk, tt, torch.zeros_like(t2))
    th = tau(t1, t2, 1.0, 0.5)
    tw = tau(t1, t2, 0.5, 1.0)
    t2s = 2.0 * t2 - t2.permute(0, 1, 3, 2)
    Loovv = blk.L_oovv()
    # build_Fae / Fmi / Fme (ccwfn.py:421-428, 441-447, 456-460) without their <mb|ef> terms
    Fae = Fvv - 0.5 * (t1.T @ Fov) - _ein('mnaf,mnef->ae', th, Loovv)
    Fmi = Foo + 0.5 * (Fov @ t1.T) + _ein('ne,mnie->mi', t1, blk.L_ooov()) + _ein('inef,mnef->mi', th, Loovv)
    Fme = Fov + _ein('nf,mnef->me', t1, Loovv)
    # build_Wmnij (:473-482)
    Wmnij = blk.oooo + _ein('je,mnie->mnij', t1, blk.ooov) + _ein('ie,mnej->mnij', t1, blk.oovo()) \
        + _ein('ijef,mnef->mnij', tt, blk.oovv)
    # build_Wmbej / Wmbje (:498-505, 520-526) without their <mb|ef> terms
    Wmbej = blk.ovvo() - _ein('nb,mnej->mbej', t1, blk.oovo()) \
        - _ein('jnfb,mnef->mbej', tw, blk.oovv) + 0.5 * _ein('njfb,mnef->mbej', t2, Loovv)
    Wmbje = -1.0 * blk.ovov + _ein('nb,mnje->mbje', t1, blk.ooov) + _ein('jnfb,mnfe->mbje', tw, blk.oovv)
    Zmbij = torch.zeros((no, blk.nv, no, no), dtype=t2.dtype, device=t2.device)
    r1_ovvv = torch.zeros_like(t1)
    r2_vvvo = torch.zeros_like(t2)
    for m in range(no):
        ov = blk.ovvv_row(m)                                            # <mb|ef> [b,e,f]
        # Fae += t_mf L_mafe, L[o,v,v,v][m,a,f,e] = 2<ma|fe> - <ma|ef>      (:425)
        Fae = Fae + 2.0 * _ein('f,afe->ae', t1[m], ov) - _ein('f,aef->ae', t1[m], ov)
        Wmbej[m] += _ein('jf,bef->bej', t1, ov)                         # + t_jf <mb|ef>  (:500)
        Wmbje[m] -= _ein('jf,bfe->bje', t1, ov)                         # - t_jf <mb|fe>  (:522)
        Zmbij[m] = _ein('bef,ijef->bij', ov, tt)                        # (:537)
        r1_ovvv += _ein('ief,aef->ia', t2s[m], ov)                      # (2t2-t2^T)_mief <ma|ef>  (:555)
        r2_vvvo[:, m] = _ein('ie,eba->iab', t1, ov)                     # t_ie <ab|ej>, <ab|ej> = <je|ba>  (:615)
        del ov
    # r_T1 (:548-557)
    r1 = Fov + t1 @ Fae.T - Fmi.T @ t1 + _ein('imae,me->ia', t2s, Fme) + _ein('nf,nafi->ia', t1, blk.L_ovvo()) \
        + r1_ovvv - _ein('mnae,nmei->ia', t2, blk.L_oovo())
    # r_T2 (:596-621)
    r = 0.5 * blk.oovv
    r = r + _ein('ijae,be->ijab', t2, Fae)
    r = r - 0.5 * _ein('ijae,be->ijab', t2, _ein('mb,me->be', t1, Fme))
    r = r - _ein('imab,mj->ijab', t2, Fmi)
    r = r - 0.5 * _ein('imab,jm->ijab', t2, _ein('je,me->jm', t1, Fme))
    r = r + 0.5 * _ein('mnab,mnij->ijab', tt, Wmnij)
    if not skip_ladder:
        r = _ladder(blk, tt, r)
    r = r - _ein('ma,mbij->ijab', t1, Zmbij)
    r = r + _ein('imae,mbej->ijab', t2 - t2.permute(0, 1, 3, 2), Wmbej)
    r = r + _ein('imae,mbej->ijab', t2, Wmbej + Wmbje.permute(0, 1, 3, 2))
    r = r + _ein('mjae,mbie->ijab', t2, Wmbje)
    tmp = t1[:, None, :, None] * t1[None, :, None, :]        # tmp[i,m,e,a] = t_ie t_ma
    r = r - _ein('imea,mbej->ijab', tmp, blk.ovvo())
    r = r - _ein('imeb,maje->ijab', tmp, blk.ovov)
    r = r + r2_vvvo
    r = r - _ein('ma,mbij->ijab', t1, blk.ovoo())
    return r1, r + r.permute(1, 0, 3, 2)


def cc_energy(F, blk, t1, t2):
    """pycc/ccwfn.py:717-718."""
    F = _t(F, device=blk.oovv.device)
    t1, t2 = _t(t1, device=F.device), _t(t2, device=F.device)
    no = blk.no
    return float(2.0 * (F[:no, no:] * t1).sum() + (tau(t1, t2) * blk.L_oovv()).sum())


def w_connected(blk, t2, i, j, k):
    """t3c_ijk with WithDenom=False (pycc/cctriples.py:10-22).  Needs only the rows i, j, k of <mb|ef>."""
    vi, vj, vk = (blk.ovvv_row(x).permute(2, 1, 0) for x in (i, j, k))    # [x,y,e] = Wvvvo[x,y,e,i]
    W = _ein('bae,ce->abc', vi, t2[k, j])
    W = W + _ein('cae,be->abc', vi, t2[j, k])
    W = W + _ein('ace,be->abc', vk, t2[j, i])
    W = W + _ein('bce,ae->abc', vk, t2[i, j])
    W = W + _ein('cbe,ae->abc', vj, t2[i, k])
    W = W + _ein('abe,ce->abc', vj, t2[k, i])
    oo = blk.ooov
    W = W - _ein('mc,mab->abc', oo[j, k], t2[i])
    W = W - _ein('mb,mac->abc', oo[k, j], t2[i])
    W = W - _ein('mb,mca->abc', oo[i, j], t2[k])
    W = W - _ein('ma,mcb->abc', oo[j, i], t2[k])
    W = W - _ein('ma,mbc->abc', oo[k, i], t2[j])
    W = W - _ein('mc,mba->abc', oo[i, k], t2[j])
    return W


def et_triplet(F, blk, t1, t2, i, j, k):
    """Energy contribution of one (i>=j>=k) triplet: body of t_tjl (pycc/cctriples.py:121-150)."""
    F = _t(F, device=blk.oovv.device)
    t1, t2 = _t(t1, device=F.device), _t(t2, device=F.device)
    no, nv = blk.no, blk.nv
    Fov, g = F[:no, no:], blk.oovv
    W = w_connected(blk, t2, i, j, k)
    D = g[i, j][:, :, None] * t1[k][None, None, :] + g[i, k][:, None, :] * t1[j][None, :, None] \
        + g[j, k][None, :, :] * t1[i][:, None, None] + t2[i, j][:, :, None] * Fov[k][None, None, :] \
        + t2[i, k][:, None, :] * Fov[j][None, :, None] + t2[j, k][None, :, :] * Fov[i][:, None, None]
    idx = torch.arange(nv, device=F.device)
    a, b, c = idx[:, None, None], idx[None, :, None], idx[None, None, :]
    V = (D + W) / (1.0 + (a == b).double() + (a == c).double() + (b == c).double())
    perms = ((0, 2, 1), (1, 0, 2), (1, 2, 0), (2, 0, 1), (2, 1, 0))
    X = W * V
    for p in perms:
        X = X + W.permute(p) * V.permute(p)
    Y = V + V.permute(1, 2, 0) + V.permute(2, 0, 1)
    Z = V.permute(0, 2, 1) + V.permute(1, 0, 2) + V.permute(2, 1, 0)
    Wc = W + W.permute(1, 2, 0) + W.permute(2, 0, 1)
    Wa = W.permute(0, 2, 1) + W.permute(1, 0, 2) + W.permute(2, 1, 0)
    ev = torch.diagonal(F)[no:]
    denom = (F[i, i] + F[j, j] + F[k, k]) - (ev[:, None, None] + ev[None, :, None] + ev[None, None, :])
    occ = 2.0 - ((i == j) + (i == k) + (j == k))
    term = ((Y - 2.0 * Z) * Wc + (Z - 2.0 * Y) * Wa + 3.0 * X) * occ / denom
    mask = (a >= b) & (b >= c)
    return float(term[mask].sum())
