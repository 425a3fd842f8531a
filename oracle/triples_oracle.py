"""
ORACLE (test infrastructure, not product code): CPU/NumPy restatement of the reference's (T)
correction, from the stored ERI blocks.  See ``oracle/ccsd_oracle.py`` for the import rules
and the parity status (PINNED against golden vectors produced by the reference's own
``t_tjl`` / ``t_vikings`` / ``t_vikings_inverted``).

The reference's ``t_tjl`` (pycc/cctriples.py:106-152) has two interpreted v^3 loops per
(i,j,k); here the same per-triplet algebra is evaluated with array masks so that it stays
usable at v ~ 100+.  ``et_triplet`` exposes the single-triplet energy for sampled parity
checks at sizes where the whole sum is too expensive on a CPU.
"""
import numpy as np

from .ccsd_oracle import Blocks, _ein


def w_connected(blk, t2, i, j, k):
    """t3c_ijk with WithDenom=False (pycc/cctriples.py:10-22): connected W[a,b,c] for one triplet.
    Wvvvo[x,y,e,i] = <xy|ei> = ovvv[i,e,y,x];  Wovoo[m,c,j,k] = <mc|jk> = ooov[j,k,m,c]."""
    vi = blk.ovvv[i].transpose(2, 1, 0)      # [x,y,e] = Wvvvo[x,y,e,i]
    vj = blk.ovvv[j].transpose(2, 1, 0)
    vk = blk.ovvv[k].transpose(2, 1, 0)
    W = _ein('bae,ce->abc', vi, t2[k, j])
    W += _ein('cae,be->abc', vi, t2[j, k])
    W += _ein('ace,be->abc', vk, t2[j, i])
    W += _ein('bce,ae->abc', vk, t2[i, j])
    W += _ein('cbe,ae->abc', vj, t2[i, k])
    W += _ein('abe,ce->abc', vj, t2[k, i])
    oo = blk.ooov                            # Wovoo[m,c,j,k] = oo[j,k,m,c]
    W -= _ein('mc,mab->abc', oo[j, k], t2[i])
    W -= _ein('mb,mac->abc', oo[k, j], t2[i])
    W -= _ein('mb,mca->abc', oo[i, j], t2[k])
    W -= _ein('ma,mcb->abc', oo[j, i], t2[k])
    W -= _ein('ma,mbc->abc', oo[k, i], t2[j])
    W -= _ein('mc,mba->abc', oo[i, k], t2[j])
    return W


def v_disconnected(F, blk, t1, t2, i, j, k):
    """t3d_ijk with WithDenom=False (pycc/cctriples.py:69-74)."""
    no = blk.no
    Fov = F[:no, no:]
    g = blk.oovv
    D = g[i, j][:, :, None] * t1[k][None, None, :]
    D += g[i, k][:, None, :] * t1[j][None, :, None]
    D += g[j, k][None, :, :] * t1[i][:, None, None]
    D += t2[i, j][:, :, None] * Fov[k][None, None, :]
    D += t2[i, k][:, None, :] * Fov[j][None, :, None]
    D += t2[j, k][None, :, :] * Fov[i][:, None, None]
    return D


def triples_denominator(F, no, i, j, k):
    """pycc/cctriples.py:139-142."""
    ev = np.diag(F)[no:]
    return (F[i, i] + F[j, j] + F[k, k]) - (ev[:, None, None] + ev[None, :, None] + ev[None, None, :])


def et_triplet(F, blk, t1, t2, i, j, k):
    """Energy contribution of one (i>=j>=k) triplet, body of t_tjl (pycc/cctriples.py:121-150)."""
    nv = blk.nv
    W = w_connected(blk, t2, i, j, k)
    V = v_disconnected(F, blk, t1, t2, i, j, k) + W
    idx = np.arange(nv)
    a, b, c = idx[:, None, None], idx[None, :, None], idx[None, None, :]
    V = V / (1.0 + (a == b) + (a == c) + (b == c))
    p_acb, p_bac, p_bca, p_cab, p_cba = (0, 2, 1), (1, 0, 2), (1, 2, 0), (2, 0, 1), (2, 1, 0)
    # X[a,b,c]: sum over the six permutations of W_p * V_p.  NumPy's X.transpose(p)[a,b,c] = X[perm of abc].
    X = W * V
    for p in (p_acb, p_bac, p_bca, p_cab, p_cba):
        X = X + W.transpose(p) * V.transpose(p)
    # reference: Y = V + V.swapaxes(0,1).swapaxes(1,2) + V.swapaxes(0,1).swapaxes(0,2)
    Y = V + V.swapaxes(0, 1).swapaxes(1, 2) + V.swapaxes(0, 1).swapaxes(0, 2)
    Z = V.swapaxes(1, 2) + V.swapaxes(0, 1) + V.swapaxes(0, 2)
    Wc = W + W.transpose(1, 2, 0) + W.transpose(2, 0, 1)       # W_abc + W_bca + W_cab
    Wa = W.transpose(0, 2, 1) + W.transpose(1, 0, 2) + W.transpose(2, 1, 0)   # W_acb + W_bac + W_cba
    denom = triples_denominator(F, blk.no, i, j, k)
    occ = 2.0 - ((i == j) + (i == k) + (j == k))
    term = ((Y - 2.0 * Z) * Wc + (Z - 2.0 * Y) * Wa + 3.0 * X) * occ / denom
    mask = (a >= b) & (b >= c)
    return float(np.sum(term[mask]))


def t_tjl(F, blocks, t1, t2, triplets=None):
    """Lee-Rendell (T) energy, sum over i>=j>=k (pycc/cctriples.py:117-152).
    ``triplets`` restricts the sum to a list of (i,j,k) (used for sampled parity)."""
    blk = blocks if isinstance(blocks, Blocks) else Blocks(blocks)
    if triplets is None:
        triplets = [(i, j, k) for i in range(blk.no) for j in range(i + 1) for k in range(j + 1)]
    et = 0.0
    for (i, j, k) in triplets:
        et += et_triplet(F, blk, t1, t2, i, j, k)
    return et


def t_vikings(F, blocks, t1, t2):
    """All-einsum (T) over all o^3 triplets (pycc/cctriples.py:173-184).
    ERI[v,k,v,v][d,b,c] = <dk|bc> = ovvv[k,d,c,b];  ERI[j,k,o,v] = ooov[j,k];  L[j,k,v,v]."""
    blk = blocks if isinstance(blocks, Blocks) else Blocks(blocks)
    no = blk.no
    Fov = F[:no, no:]
    Loovv = blk.L_oovv()
    X1 = np.zeros_like(t1)
    X2 = np.zeros_like(t2)
    for i in range(no):
        for j in range(no):
            for k in range(no):
                t3 = w_connected(blk, t2, i, j, k) / triples_denominator(F, no, i, j, k)
                u = t3 - t3.swapaxes(0, 2)
                w = 2.0 * t3 - t3.swapaxes(1, 2) - t3.swapaxes(0, 2)
                X1[i] += _ein('abc,bc->a', u, Loovv[j, k])
                X2[i, j] += _ein('abc,c->ab', u, Fov[k])
                X2[i, j] += _ein('abc,dbc->ad', w, blk.ovvv[k].transpose(0, 2, 1))
                X2[i] -= _ein('abc,lc->lab', w, blk.ooov[j, k])
    et = 2.0 * np.sum(t1 * X1)
    et += np.sum((4.0 * t2 - 2.0 * t2.swapaxes(2, 3)) * X2)
    return float(et)


def flops_triples(o, v):
    """Algorithmic flops of the restricted-triplet W build."""
    return (o * (o + 1) * (o + 2) // 6) * (12 * v**4 + 12 * o * v**3)
