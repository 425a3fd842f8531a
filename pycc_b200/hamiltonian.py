"""
hamiltonian.py: array-layout contract of ``pycc/hamiltonian.py:26-39``.

The reference's Hamiltonian holds ``F (n,n)``, the full Dirac array ``ERI[p,q,r,s] = <pq|rs>``
(``mo_eri(...).swapaxes(1,2)``, :36-38) and ``L = 2*ERI - ERI.swapaxes(2,3)`` (:39): two n^4
arrays (600 GB at o=40, v=400).  The CCSD / (T) path only ever slices ten orbital-space blocks
out of ``ERI`` and five out of ``L`` (pycc/ccwfn.py:413-616, pycc/cctriples.py:121-122,176-181),
and by the 8-fold symmetry of real orbitals all of them are views or two-term combinations of
SIX blocks.  This class stores exactly those six, on the device:

    oooo[m,n,i,j] = <mn|ij>   ooov[m,n,i,e] = <mn|ie>   oovv[i,j,a,b] = <ij|ab>
    ovov[m,b,j,e] = <mb|je>   ovvv[m,b,e,f] = <mb|ef>   vvvv[a,b,e,f] = <ab|ef>

``Hamiltonian.from_full(F, ERI, no)`` accepts the reference's full array (small systems, drop-in
use); ``Hamiltonian.from_blocks`` accepts the blocks directly (large / synthetic systems, where
the n^4 array cannot exist).  ``H.ERI`` and ``H.L`` are block-provider objects supporting the
slice forms the reference uses, e.g. ``H.ERI[o,o,v,v]`` or ``H.L[o,v,v,v]``, returning device
views / tensors in the reference's index order.
"""
import numpy as np
import torch

BLOCKS = ("oooo", "ooov", "oovv", "ovov", "ovvv", "vvvv")


def blocks_from_full(ERI, no):
    """Slice the six stored blocks out of a full (n,n,n,n) Dirac array (NumPy or torch)."""
    n = ERI.shape[0]
    o, v = slice(0, no), slice(no, n)
    return {"oooo": ERI[o, o, o, o], "ooov": ERI[o, o, o, v], "oovv": ERI[o, o, v, v],
            "ovov": ERI[o, v, o, v], "ovvv": ERI[o, v, v, v], "vvvv": ERI[v, v, v, v]}


class _BlockProvider(object):
    """``provider[o,o,v,v]``-style access to ERI (spin_adapt=False) or L (spin_adapt=True)."""

    # reference slice pattern -> (stored block, torch.permute argument giving the reference index order)
    _MAP = {
        "oooo": ("oooo", (0, 1, 2, 3)),
        "ooov": ("ooov", (0, 1, 2, 3)),
        "oovo": ("ooov", (1, 0, 3, 2)),     # <mn|ej> = <nm|je>
        "ovoo": ("ooov", (2, 3, 0, 1)),     # <mb|ij> = <ij|mb>
        "vooo": ("ooov", (3, 2, 1, 0)),     # <am|ij> = <ji|ma>
        "oovv": ("oovv", (0, 1, 2, 3)),
        "vvoo": ("oovv", (2, 3, 0, 1)),
        "ovvo": ("oovv", (0, 3, 2, 1)),     # <mb|ej> = <mj|eb>
        "voov": ("oovv", (3, 0, 1, 2)),     # <am|ie> = <mi|ea>  -> oovv[m,i,e,a]
        "ovov": ("ovov", (0, 1, 2, 3)),
        "vovo": ("ovov", (1, 0, 3, 2)),
        "ovvv": ("ovvv", (0, 1, 2, 3)),
        "vovv": ("ovvv", (1, 0, 3, 2)),     # <am|ef> = <ma|fe>
        "vvov": ("ovvv", (2, 3, 0, 1)),     # <ab|ie> = <ie|ab>
        "vvvo": ("ovvv", (3, 2, 1, 0)),     # <ab|ej> = <je|ba>
        "vvvv": ("vvvv", (0, 1, 2, 3)),
    }

    def __init__(self, H, spin_adapt):
        self.H, self.spin_adapt = H, spin_adapt

    def _space(self, s):
        H = self.H
        if s == H.o or (isinstance(s, slice) and (s.start or 0) == 0 and s.stop == H.no):
            return "o"
        if s == H.v or (isinstance(s, slice) and s.start == H.no and s.stop in (None, H.no + H.nv)):
            return "v"
        raise IndexError("only whole occupied / virtual slices are supported by the block store (got %r)" % (s,))

    def _block(self, pattern):
        """The reference-ordered view: reference[x0..x3] = stored.permute(perm)[x0..x3]."""
        name, perm = self._MAP[pattern]
        return self.H.blocks[name].permute(*perm)

    def __getitem__(self, key):
        if not (isinstance(key, tuple) and len(key) == 4):
            raise IndexError("ERI / L are accessed by four orbital-space slices")
        pattern = "".join(self._space(s) for s in key)
        if pattern not in self._MAP:
            raise IndexError("block %s is not reachable from the stored blocks" % pattern)
        g = self._block(pattern)
        if not self.spin_adapt:
            return g
        # L[pqrs] = 2<pq|rs> - <pq|sr>: the exchange partner is the block with the last two spaces swapped
        swapped = pattern[:2] + pattern[3] + pattern[2]
        if swapped not in self._MAP:
            raise IndexError("L block %s is not reachable from the stored blocks" % pattern)
        x = self._block(swapped).transpose(2, 3)
        return 2.0 * g - x


class Hamiltonian(object):
    def __init__(self, F, blocks, no, nv, device, dtype=torch.float64):
        self.no, self.nv = int(no), int(nv)
        self.o, self.v = slice(0, self.no), slice(self.no, self.no + self.nv)
        self.device, self.dtype = device, dtype
        self.F = F
        self.blocks = blocks
        self.ERI = _BlockProvider(self, False)
        self.L = _BlockProvider(self, True)

    @staticmethod
    def _to_dev(x, device, dtype):
        if isinstance(x, torch.Tensor):
            return x.to(device=device, dtype=dtype).contiguous()
        return torch.as_tensor(np.ascontiguousarray(x), dtype=dtype).to(device)

    @classmethod
    def from_blocks(cls, F, blocks, no, device, dtype=torch.float64, vslice=None, keep_full_ovvv=True):
        """``vslice=(lo, hi)`` (multi-GPU): only rows lo..hi of the leading virtual index of ``vvvv``
        (and of the ``b`` index of ``ovvv`` unless ``keep_full_ovvv``) are uploaded; blocks that are
        already given in that local shape are taken as they are.  ``H.ERI[v,v,v,v]`` then yields
        this rank's rows only."""
        nv = blocks["oovv"].shape[2]
        dev_blocks = {}
        for k in BLOCKS:
            if k not in blocks:
                raise ValueError("missing ERI block %r (need %s)" % (k, ", ".join(BLOCKS)))
            src = blocks[k]
            if vslice is not None and tuple(vslice) != (0, nv):
                lo, hi = vslice
                if k == "vvvv" and src.shape[0] == nv:
                    src = src[lo:hi]
                elif k == "ovvv" and not keep_full_ovvv and src.shape[1] == nv:
                    src = src[:, lo:hi]
            dev_blocks[k] = cls._to_dev(src, device, dtype)
        H = cls(cls._to_dev(F, device, dtype), dev_blocks, no, nv, device, dtype)
        H.vslice = tuple(vslice) if vslice is not None else (0, nv)
        return H

    @classmethod
    def from_full(cls, F, ERI, no, device, dtype=torch.float64, vslice=None, keep_full_ovvv=True):
        return cls.from_blocks(F, blocks_from_full(ERI, no), no, device, dtype, vslice, keep_full_ovvv)


# ---- integral front end without psi4 ------------------------------------------
# The reference builds  F = C^T Fa C  and  ERI = mo_eri(C, C, C, C).swapaxes(1, 2)  through psi4's MintsHelper
# (pycc/hamiltonian.py:32-38).  Without psi4 the same arrays come from (a) AO-basis integrals and the MO
# coefficients, by the O(n^5) quarter transformations, or (b) a density-fitting / Cholesky factor
# B[Q,p,q], (pq|rs) = sum_Q B[Q,p,q] B[Q,r,s]; both run on the device as pccb_contract launches, and (b)
# never forms an n^4 array: each rank manufactures only the blocks (and <ab|ef> / <mb|ef> rows) it owns.
def _run(backend, stmt, tensors, dtype):
    from . import _lib
    from .dsl import TRef, plan
    backend.bind_stream()
    T = {k: TRef.of(t, k) for k, t in tensors.items()}
    for L in plan(stmt, T, _lib.F64 if dtype == torch.float64 else _lib.F32):
        backend.contract(L.desc)


def blocks_from_factor(backend, B, no, dtype=torch.float64, vslice=None, keep_full_ovvv=True, row_chunk=None):
    """The six stored Dirac blocks <pq|rs> = (pr|qs) = sum_Q B[Q,p,r] B[Q,q,s] from a factor B[Q,n,n] (symmetric
    in its orbital indices), each as one pccb_contract launch per chunk of leading rows (``row_chunk`` bounds
    nothing but the launch size: the blocks are written in place, no temporary).  ``vslice`` as in from_blocks."""
    Bd = B.to(device=backend.device, dtype=dtype).contiguous() if isinstance(B, torch.Tensor) else \
        torch.as_tensor(np.ascontiguousarray(B), dtype=dtype).to(backend.device)
    n = Bd.shape[1]
    nv = n - no
    lo, hi = (0, nv) if vslice is None else vslice
    sp = {"o": (0, no), "v": (no, n)}

    def build(pattern, p_rng=None, q_rng=None):
        rng = [sp[c] for c in pattern]
        if p_rng is not None:
            rng[0] = p_rng
        if q_rng is not None:
            rng[1] = q_rng
        (p0, p1), (q0, q1), (r0, r1), (s0, s1) = rng
        out = torch.empty((p1 - p0, q1 - q0, r1 - r0, s1 - s0), dtype=dtype, device=backend.device)
        step = row_chunk or (p1 - p0)
        for a in range(p0, p1, max(step, 1)):
            b = min(p1, a + step)
            _run(backend, "X[pqrs] = Y[Qpr]*Z[Qqs]",
                 dict(X=out[a - p0:b - p0], Y=Bd[:, a:b, r0:r1], Z=Bd[:, q0:q1, s0:s1]), dtype)
        return out

    return {"oooo": build("oooo"), "ooov": build("ooov"), "oovv": build("oovv"), "ovov": build("ovov"),
            "ovvv": build("ovvv", q_rng=None if keep_full_ovvv else (no + lo, no + hi)),
            "vvvv": build("vvvv", p_rng=(no + lo, no + hi), )}


def mo_transform(backend, Fao, ao_eri, C, dtype=torch.float64):
    """F = C^T Fa C and the Dirac array ERI[p,q,r,s] = <pq|rs> = (pr|qs) from AO-basis integrals in chemists'
    order ao_eri[mu,nu,la,si] = (mu nu|la si) and MO coefficients C[mu,p]: what MintsHelper.mo_eri(C,C,C,C)
    .swapaxes(1,2) returns (pycc/hamiltonian.py:36-38), as four quarter transformations = four pccb_contract
    launches.  The result is a permuted VIEW of the chemists'-order MO tensor, like the reference's."""
    dev = backend.device
    to = lambda x: (x if isinstance(x, torch.Tensor) else torch.as_tensor(np.ascontiguousarray(x))).to(device=dev, dtype=dtype).contiguous()
    Fao, g, C = to(Fao), to(ao_eri), to(C)
    nao, nmo = C.shape
    E = lambda *s: torch.empty(s, dtype=dtype, device=dev)
    h = E(nao, nmo)
    F = E(nmo, nmo)
    _run(backend, "h[mq] = Fa[mn]*C[nq]", dict(h=h, Fa=Fao, C=C), dtype)
    _run(backend, "F[pq] = C[mp]*h[mq]", dict(F=F, C=C, h=h), dtype)
    t1, t2 = E(nmo, nao, nao, nao), E(nmo, nmo, nao, nao)
    _run(backend, "t[pnls] = C[mp]*g[mnls]", dict(t=t1, C=C, g=g), dtype)
    _run(backend, "t[pqls] = C[nq]*u[pnls]", dict(t=t2, C=C, u=t1), dtype)
    del t1
    t3 = E(nmo, nmo, nmo, nao)
    _run(backend, "t[pqrs] = C[lr]*u[pqls]", dict(t=t3, C=C, u=t2), dtype)
    del t2
    mo = E(nmo, nmo, nmo, nmo)
    _run(backend, "t[pqrs] = C[ms]*u[pqrm]", dict(t=mo, C=C, u=t3), dtype)
    return F, mo.permute(0, 2, 1, 3)            # (pq|rs) chemists' -> <pr|qs>: ERI[p,q,r,s] = mo[p,r,q,s]


def _from_factor(cls, F, B, no, backend, dtype=torch.float64, vslice=None, keep_full_ovvv=True):
    """Hamiltonian from a DF / Cholesky factor B[Q,n,n] in the MO basis (never an n^4 array)."""
    blocks = blocks_from_factor(backend, B, no, dtype, vslice, keep_full_ovvv, row_chunk=8)
    n = blocks["oovv"].shape[2] + no
    H = cls(cls._to_dev(F, backend.device, dtype), blocks, no, n - no, backend.device, dtype)
    H.vslice = tuple(vslice) if vslice is not None else (0, n - no)
    return H


def _from_ao(cls, Fao, ao_eri, C, no, backend, dtype=torch.float64, vslice=None, keep_full_ovvv=True):
    """Hamiltonian from AO-basis Fock / two-electron integrals and MO coefficients (small systems: n^4)."""
    F, ERI = mo_transform(backend, Fao, ao_eri, C, dtype)
    return cls.from_full(F, ERI, no, backend.device, dtype, vslice, keep_full_ovvv)


Hamiltonian.from_factor = classmethod(_from_factor)
Hamiltonian.from_ao = classmethod(_from_ao)
