"""
TEST INFRASTRUCTURE ONLY: a NumPy interpreter of the C-ABI calls, used by the ``not gpu`` tests
to validate the engine's *host logic* (the contraction plans / descriptors, the DIIS and solver
control flow, the sharding plans) against the oracle without a GPU.

It interprets exactly the structures the CUDA library receives (``pccb_contraction`` with raw
pointers, compound indices, segments, batch strides) by dereferencing the host addresses of
CPU torch tensors.  It is not importable from the product package and the product never
falls back to it: ``pycc_b200.backend.CudaBackend`` is the only back end ``ccwfn`` constructs.
"""
import ctypes

import numpy as np
import torch


def _mem(ptr, n, dtype=np.float64):
    buf = (ctypes.c_char * (n * np.dtype(dtype).itemsize)).from_address(ptr)
    return np.frombuffer(buf, dtype=dtype, count=n)


def _offs(c, extent):
    i = np.arange(extent, dtype=np.int64)
    if c.div <= 0 or c.div >= extent:
        return i * c.s_lo
    return (i // c.div) * c.s_hi + (i % c.div) * c.s_lo


class InterpBackend(object):
    def __init__(self):
        self.device = torch.device("cpu")
        self.n_contract = 0
        self.log = []

    def ensure_workspace(self, nbytes):
        pass

    def bind_stream(self):
        pass

    def synchronize(self):
        pass

    def launch_count(self):
        return self.n_contract

    def empty(self, shape, dtype=torch.float64):
        return torch.empty(shape, dtype=dtype)

    def zeros(self, shape, dtype=torch.float64):
        return torch.zeros(shape, dtype=dtype)

    def to_device(self, x, dtype=torch.float64):
        return torch.as_tensor(np.array(x) if not isinstance(x, torch.Tensor) else x, dtype=dtype).contiguous().clone()

    # ---- pccb_contract ------------------------------------------------------------------------
    def contract(self, d):
        self.n_contract += 1
        npdt = np.float64 if d.dtype == 0 else np.float32
        M, N = d.M, d.N
        cm, cn = _offs(d.cm, M), _offs(d.cn, N)
        for b in range(d.batch):
            acc = np.zeros((M, N), dtype=np.float64)
            for q in range(d.nseg):
                s = d.seg[q]
                K = s.K
                if K == 0:
                    continue
                oa = _offs(s.am, M)[:, None] + _offs(s.ak, K)[None, :] + b * s.a_batch
                ob = _offs(s.bn, N)[:, None] + _offs(s.bk, K)[None, :] + b * s.b_batch
                assert oa.min() >= 0 and ob.min() >= 0
                A = _mem(s.A_dev, int(oa.max()) + 1, npdt)[oa]
                B = _mem(s.B_dev, int(ob.max()) + 1, npdt)[ob]
                acc += s.scale * (A.astype(np.float64) @ B.astype(np.float64).T)
            oc = cm[:, None] + cn[None, :] + b * d.c_batch
            # every output element must be addressed exactly once
            assert np.unique(oc).size == oc.size, "output offsets collide"
            Cmem = _mem(d.C_dev, int(oc.max()) + 1, npdt)
            val = d.alpha * acc
            if d.beta != 0.0:
                val = val + d.beta * Cmem[oc]
            if d.D_dev:
                od = _offs(d.dm, M)[:, None] + _offs(d.dn, N)[None, :] + b * d.d_batch
                val = val + d.gamma * _mem(d.D_dev, int(od.max()) + 1, npdt)[od]
            Cmem[oc] = val.astype(npdt)

    # ---- element-wise entry points ------------------------------------------------------------
    def pair_combine(self, out, X, c0, c1, u, w, c2, no, nv):
        r = c0 * X + c1 * X.transpose(2, 3)
        if u is not None:
            r = r + c2 * u[:, None, :, None] * w[None, :, None, :]
        out.copy_(r)

    def pair_symmetrize(self, r2, no, nv):
        r2.copy_(r2 + r2.permute(1, 0, 3, 2))

    def permute4(self, out, src_view, alpha=1.0, beta=0.0):
        out.copy_(alpha * src_view if beta == 0.0 else alpha * src_view + beta * out)

    def update_amplitudes(self, t1, t2, r1, r2, eps, no, nv, symmetrize):
        eo, ev = eps[:no], eps[no:]
        Dia = eo[:, None] - ev[None, :]
        Dijab = eo[:, None, None, None] + eo[None, :, None, None] - ev[None, None, :, None] - ev[None, None, None, :]
        rr = r2 + r2.permute(1, 0, 3, 2) if symmetrize else r2
        i1, i2 = r1 / Dia, rr / Dijab
        t1 += i1
        t2 += i2
        return float((i1.double() ** 2).sum() + (i2.double() ** 2).sum())

    def update_energy(self, t1, t2, r1, r2, eps, fov, ld_fov, oovv, no, nv, symmetrize, ij_lo, ij_hi, count_t1):
        eo, ev = eps[:no], eps[no:]
        Dia = eo[:, None] - ev[None, :]
        Dijab = eo[:, None, None, None] + eo[None, :, None, None] - ev[None, None, :, None] - ev[None, None, None, :]
        rr = r2 + r2.permute(1, 0, 3, 2) if symmetrize else r2
        i1 = r1 / Dia
        t1 += i1
        sel = torch.zeros(no * no, dtype=torch.bool)
        sel[ij_lo:ij_hi] = True
        sel = sel.reshape(no, no)
        i2 = (rr / Dijab) * sel[:, :, None, None].to(rr.dtype)
        t2 += i2
        tau = t2 + t1[:, None, :, None] * t1[None, :, None, :]
        L = 2.0 * oovv - oovv.transpose(2, 3)
        e2 = float(((tau * L).double() * sel[:, :, None, None]).sum())
        rms2 = float((i2.double() ** 2).sum())
        if count_t1:
            rms2 += float((i1.double() ** 2).sum())
            e2 += float(2.0 * (fov * t1).double().sum())
        return rms2, e2

    def prepare_amplitudes(self, t1, t2, no, nv, vlo, vl, **outs):
        pr = t1[:, None, :, None] * t1[None, :, None, :]
        t2s = 2.0 * t2 - t2.transpose(2, 3)
        tauW = 0.5 * t2 + pr
        val = dict(tau=t2 + pr, tauh=t2 + 0.5 * pr, tauW=tauW, t2s=t2s, t11=pr,
                   t2sP=t2s.permute(0, 2, 1, 3), t2P=t2.permute(0, 2, 1, 3), t2Q=t2.permute(1, 2, 0, 3),
                   tauWPn=-tauW.permute(0, 3, 1, 2)[:, vlo:vlo + vl], t2Rh=0.5 * t2.permute(1, 3, 0, 2)[:, vlo:vlo + vl])
        for k, out in outs.items():
            if out is not None:
                out.copy_(val[k])

    def cc_energy(self, fov, ld_fov, t1, t2, oovv, no, nv):
        tau = t2 + t1[:, None, :, None] * t1[None, :, None, :]
        L = 2.0 * oovv - oovv.transpose(2, 3)
        return float(2.0 * (fov * t1).sum() + (tau * L).sum())

    def dot(self, x, y):
        return float((x.double().reshape(-1) * y.double().reshape(-1)).sum())

    def dot_many(self, x, ys):
        return [self.dot(x, y) for y in ys]

    def axpby(self, out, a, x, b=0.0, y=None):
        out.copy_(a * x if y is None else a * x + b * y)

    def lincomb(self, out, vecs, coefs):
        acc = torch.zeros_like(out, dtype=torch.float64)
        for c, v in zip(coefs, vecs):
            acc += c * v.double()
        out.copy_(acc.to(out.dtype))

    def triples_energy(self, W, ijk_dev, ntrip, t1, t2, oovv, fov, ld_fov, eps, no, nv, et_dev):
        from oracle import triples_oracle as ot   # test-side checker
        F = np.zeros((no + nv, no + nv))
        F[np.arange(no + nv), np.arange(no + nv)] = eps.numpy()
        F[:no, no:] = fov.numpy()
        idx = np.arange(nv)
        a, b, c = idx[:, None, None], idx[None, :, None], idx[None, None, :]
        ijk = ijk_dev.numpy().reshape(-1, 3)[:ntrip]

        class _B(object):
            pass
        blk = _B()
        blk.oovv, blk.no, blk.nv = oovv.numpy(), no, nv
        for t, (i, j, k) in enumerate(ijk):
            Wt = W[t].numpy().astype(np.float64)
            V = (ot.v_disconnected(F, blk, t1.numpy(), t2.numpy(), i, j, k) + Wt) / (1.0 + (a == b) + (a == c) + (b == c))
            X = Wt * V
            for p in ((0, 2, 1), (1, 0, 2), (1, 2, 0), (2, 0, 1), (2, 1, 0)):
                X = X + Wt.transpose(p) * V.transpose(p)
            Y = V + V.transpose(1, 2, 0) + V.transpose(2, 0, 1)
            Z = V.transpose(0, 2, 1) + V.transpose(1, 0, 2) + V.transpose(2, 1, 0)
            Wc = Wt + Wt.transpose(1, 2, 0) + Wt.transpose(2, 0, 1)
            Wa = Wt.transpose(0, 2, 1) + Wt.transpose(1, 0, 2) + Wt.transpose(2, 1, 0)
            den = ot.triples_denominator(F, no, i, j, k)
            occ = 2.0 - ((i == j) + (i == k) + (j == k))
            term = ((Y - 2.0 * Z) * Wc + (Z - 2.0 * Y) * Wa + 3.0 * X) * occ / den
            et_dev[t] = float(term[(a >= b) & (b >= c)].sum())

    def split_lo(self, x, lo):
        hi = (x.view(torch.int32) & -8192).view(torch.float32)
        lo.copy_(x - hi)

    def register_lo_plane(self, x, lo):
        # the interpreter contracts the full FP32 values; the registry is kept so that tests can check the protocol
        self.__dict__.setdefault("lo_registry", {})[x.data_ptr()] = lo

    def unregister_lo_plane(self, x_ptr):
        self.__dict__.setdefault("lo_registry", {}).pop(x_ptr, None)

    def t3_density_combine(self, W, Dd, nk, k0, i, j, eps, no, nv, outs):
        e = eps.double()
        ev = e[no:]
        X = lambda T: 8 * T - 4 * T.permute(1, 0, 2) - 4 * T.permute(0, 2, 1) - 4 * T.permute(2, 1, 0) \
            + 2 * T.permute(2, 0, 1) + 2 * T.permute(1, 2, 0)
        for kk in range(nk):
            d = e[i] + e[j] + e[k0 + kk] - (ev[:, None, None] + ev[None, :, None] + ev[None, None, :])
            M, N = W[kk].double() / d, Dd[kk].double() / d
            val = dict(M=M, A1=M - M.permute(2, 1, 0), A2=2 * M - M.permute(0, 2, 1) - M.permute(2, 1, 0),
                       B1=2 * (M - M.permute(1, 0, 2)),
                       Z3=2 * (M - M.permute(0, 2, 1)) - (M.permute(1, 0, 2) - M.permute(2, 0, 1)),    # permute(2,0,1)[a,b,c] = M[b,c,a]
                       P=2 * X(M) + X(N), Q=X(M) + X(N))
            for k, out in outs.items():
                if out is not None:
                    out[kk].copy_(val[k].to(out.dtype))

    def t3_combine(self, W, nk, k0, i, j, eps, no, nv, A1, A2):
        e = eps.double()
        ev = e[no:]
        for kk in range(nk):
            d = e[i] + e[j] + e[k0 + kk] - (ev[:, None, None] + ev[None, :, None] + ev[None, None, :])
            t3 = W[kk].double() / d
            A1[kk].copy_((t3 - t3.permute(2, 1, 0)).to(A1.dtype))
            A2[kk].copy_((2.0 * t3 - t3.permute(0, 2, 1) - t3.permute(2, 1, 0)).to(A2.dtype))
