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
