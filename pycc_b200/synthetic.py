"""
Seeded synthetic, symmetry-correct RHF inputs for the CCSD / (T) hot path.

The reference obtains ``F`` and ``ERI`` from psi4 (``pycc/hamiltonian.py:32-39``), which is
not available offline; the parity runs and the benchmark therefore use a density-fitting-like
factor ``B[Q,p,q]`` (symmetric in p,q) so that the chemists' integrals
``(pq|rs) = sum_Q B[Q,p,q] B[Q,r,s]`` carry the full 8-fold permutational symmetry of real
orbitals, and the Dirac array is ``<pq|rs> = (pr|qs)`` exactly as ``hamiltonian.py:38`` builds
it (``mo_eri(...).swapaxes(1,2)``).  Recipe: SURVEY.md section 8(d).

The engine and the timed kernels always see the blocks as *dense* data; the factor form is
only used to manufacture them (and, in tests, as an independent checker).
"""
import numpy as np

BLOCK_NAMES = ("oooo", "ooov", "oovv", "ovov", "ovvv", "vvvv")


def make_factor(no, nv, seed=0, naux=None, dtype=np.float64):
    """B[Q,p,q] (unit scale), orbital energies eps[n] and a symmetric Fock perturbation P[n,n]."""
    n = no + nv
    naux = 3 * n if naux is None else naux
    rng = np.random.default_rng(seed)
    G = rng.standard_normal((naux, n, n))
    B = (0.5 / np.sqrt(naux)) * (G + G.transpose(0, 2, 1))
    eps = np.concatenate((np.sort(-2.0 + rng.random(no)), np.sort(0.5 + 2.0 * rng.random(nv))))
    P = rng.standard_normal((n, n))
    P = 1e-3 * (P + P.T)
    return B.astype(dtype), eps.astype(dtype), P.astype(dtype)


def _blk(B, p, q, r, s):
    """<pq|rs> = (pr|qs) = sum_Q B[Q,p,r] B[Q,q,s] for orbital ranges p,q,r,s (slices)."""
    X = B[:, p, r]
    Y = B[:, q, s]
    naux = B.shape[0]
    out = X.reshape(naux, -1).T @ Y.reshape(naux, -1)          # [(p r), (q s)]
    out = out.reshape(X.shape[1], X.shape[2], Y.shape[1], Y.shape[2])
    return np.ascontiguousarray(out.transpose(0, 2, 1, 3))     # [p,q,r,s]


def mp2_energy_from_oovv(oovv, eps, no):
    eo, ev = eps[:no], eps[no:]
    D = eo[:, None, None, None] + eo[None, :, None, None] - ev[None, None, :, None] - ev[None, None, None, :]
    t2 = oovv / D
    return float(np.sum(t2 * (2.0 * oovv - oovv.transpose(0, 1, 3, 2))))


def calibrated_scale(B, eps, no, target_emp2=-0.10):
    """Scale s for B so that the MP2 correlation energy of the synthetic system equals
    ``target_emp2`` (E_MP2 is homogeneous of degree 4 in s)."""
    o = slice(0, no); v = slice(no, B.shape[1])
    e1 = mp2_energy_from_oovv(_blk(B, o, o, v, v), eps, no)
    return float((target_emp2 / e1) ** 0.25)


class SyntheticSystem(object):
    """Container for one synthetic closed-shell system.

    Attributes: no, nv, eps, F (n,n), B (naux,n,n; already scaled), blocks (dict of the six
    distinct Dirac blocks, each C-contiguous): oooo <mn|ij>, ooov <mn|ie>, oovv <ij|ab>,
    ovov <mb|je>, ovvv <mb|ef>, vvvv <ab|ef>.
    """

    def __init__(self, no, nv, seed=0, scale=None, target_emp2=-0.10, fock_offdiag=False,
                 with_vvvv=True):
        self.no, self.nv, self.seed = no, nv, seed
        B, eps, P = make_factor(no, nv, seed)
        if scale is None:
            scale = calibrated_scale(B, eps, no, target_emp2)
        self.scale = scale
        self.B = B * scale
        self.eps = eps
        self.F = np.diag(eps).copy()
        if fock_offdiag:
            self.F += P
        self.fock_offdiag = fock_offdiag
        o = slice(0, no); v = slice(no, no + nv)
        B = self.B
        self.blocks = {
            "oooo": _blk(B, o, o, o, o), "ooov": _blk(B, o, o, o, v), "oovv": _blk(B, o, o, v, v),
            "ovov": _blk(B, o, v, o, v), "ovvv": _blk(B, o, v, v, v),
        }
        if with_vvvv:
            self.blocks["vvvv"] = _blk(B, v, v, v, v)

    def full_eri(self):
        """The full Dirac (n,n,n,n) array the reference's Hamiltonian would hold
        (``hamiltonian.py:36-38``). Small systems only."""
        n = self.no + self.nv
        a = slice(0, n)
        return _blk(self.B, a, a, a, a)


def device_blocks(B, no, device, vslice=None, keep_full_ovvv=True, dtype=None):
    """The six blocks built ON THE DEVICE from the (scaled) factor ``B`` (set-up only, untimed:
    plain ``torch`` GEMMs; at o=40, v=400 the host cannot hold the 205 GB ``vvvv`` block, each rank
    manufactures just its own rows).  ``vslice=(lo, hi)`` restricts ``vvvv`` to rows a in [lo, hi)
    (and ``ovvv`` to b in [lo, hi) unless ``keep_full_ovvv``).  Same numbers as ``_blk``."""
    import torch
    dtype = dtype or torch.float64
    Bd = torch.as_tensor(B).to(device=device, dtype=torch.float64)
    n = Bd.shape[1]
    nv = n - no
    o, v = slice(0, no), slice(no, n)
    lo, hi = (0, nv) if vslice is None else vslice
    vs = slice(no + lo, no + hi)

    def blk(p, q, r, s, chunk=None):
        X = Bd[:, p, r]
        Y = Bd[:, q, s]
        naux = Bd.shape[0]
        P, R, Q, S = X.shape[1], X.shape[2], Y.shape[1], Y.shape[2]
        out = torch.empty((P, Q, R, S), dtype=dtype, device=device)
        Ym = Y.reshape(naux, -1)
        step = chunk or P
        for p0 in range(0, P, step):        # chunked over p to bound the temporary
            p1 = min(P, p0 + step)
            tmp = (X[:, p0:p1].reshape(naux, -1).T @ Ym).reshape(p1 - p0, R, Q, S)
            out[p0:p1].copy_(tmp.permute(0, 2, 1, 3))
        return out

    blocks = {"oooo": blk(o, o, o, o), "ooov": blk(o, o, o, v), "oovv": blk(o, o, v, v),
              "ovov": blk(o, v, o, v),
              "ovvv": blk(o, v if keep_full_ovvv else vs, v, v, chunk=4),
              "vvvv": blk(vs, v, v, v, chunk=8)}
    return blocks
