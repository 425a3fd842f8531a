"""
CudaBackend: the one execution back end of the engine - every call goes through the C ABI of
libpyccb200.so (include/pyccb200.h) onto the current CUDA stream of PyTorch, which is used
for device-memory ownership only.  Constructing it without the shared object or without a
B200 raises EngineError; nothing here (or anywhere in pycc_b200) computes on the CPU.
"""
import ctypes as C

import torch

from . import _lib
from ._lib import EngineError

_DT = {torch.float64: _lib.F64, torch.float32: _lib.F32}


def dtype_code(t):
    try:
        return _DT[t.dtype if isinstance(t, torch.Tensor) else t]
    except KeyError:
        raise EngineError("unsupported dtype %r (float64 / float32 only)" % (t,))


class CudaBackend(object):
    def __init__(self, device=None, workspace_bytes=256 << 20):
        self.lib = _lib.load()
        if not torch.cuda.is_available() or self.lib.pccb_device_count() <= 0:
            raise EngineError("no CUDA device visible: the pycc_b200 engine has no CPU fallback")
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        if self.device.type != "cuda":
            raise EngineError("pycc_b200 runs on CUDA devices only (got %s)" % self.device)
        index = self.device.index if self.device.index is not None else torch.cuda.current_device()
        self.device = torch.device("cuda", index)
        h = C.c_void_p()
        rc = self.lib.pccb_create(index, C.byref(h))
        if rc != 0:
            raise EngineError("pccb_create failed (%d): %s" % (rc, self.lib.pccb_last_error(None).decode()))
        self.h = h
        self._ws = torch.empty(workspace_bytes, dtype=torch.uint8, device=self.device)
        self._check(self.lib.pccb_set_workspace(self.h, self._ws.data_ptr(), workspace_bytes))
        self._stream = None
        self.bind_stream()

    def ensure_workspace(self, nbytes):
        """Grow the split-K workspace to at least ``nbytes`` (the engine sizes it for the tail-aware split of a
        rank's largest output, pccb_contract in contract.cu).  Launches already queued keep the old buffer alive
        until they finish (stream-ordered free of the caching allocator)."""
        nbytes = int(nbytes)
        if nbytes > self._ws.numel():
            self.synchronize()
            self._ws = torch.empty(nbytes, dtype=torch.uint8, device=self.device)
            self._check(self.lib.pccb_set_workspace(self.h, self._ws.data_ptr(), nbytes))

    # -- plumbing ------------------------------------------------------------------------------
    def _check(self, rc):
        if rc != 0:
            raise EngineError("libpyccb200 error %d: %s" % (rc, self.lib.pccb_last_error(self.h).decode()))

    def bind_stream(self):
        """Launch on PyTorch's current stream for this device (so torch.cuda.Event timing and
        tensor lifetimes see the kernels)."""
        s = torch.cuda.current_stream(self.device).cuda_stream
        if s != self._stream:
            self._check(self.lib.pccb_set_stream(self.h, s))
            self._stream = s

    def close(self):
        if getattr(self, "h", None) is not None and self.h:
            self.lib.pccb_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def empty(self, shape, dtype=torch.float64):
        return torch.empty(shape, dtype=dtype, device=self.device)

    def zeros(self, shape, dtype=torch.float64):
        return torch.zeros(shape, dtype=dtype, device=self.device)

    def to_device(self, x, dtype=torch.float64):
        if isinstance(x, torch.Tensor):
            return x.to(device=self.device, dtype=dtype).contiguous()
        return torch.as_tensor(x, dtype=dtype).to(self.device).contiguous()

    def use_naive_contract(self, flag):
        self._check(self.lib.pccb_set_contract_impl(self.h, _lib.IMPL_NAIVE if flag else _lib.IMPL_TENSOR))

    def launch_count(self):
        n = C.c_int64()
        self._check(self.lib.pccb_launch_count(self.h, C.byref(n)))
        return n.value

    def synchronize(self):
        self._check(self.lib.pccb_synchronize(self.h))

    # -- operator seam ---------------------------------------------------------------------------
    def contract(self, desc):
        self._check(self.lib.pccb_contract(self.h, C.byref(desc)))

    # -- solver seam -----------------------------------------------------------------------------
    def pair_combine(self, out, X, c0, c1, u, w, c2, no, nv):
        self._check(self.lib.pccb_pair_combine(self.h, dtype_code(out), out.data_ptr(), X.data_ptr(), c0, c1,
                                               None if u is None else u.data_ptr(),
                                               None if w is None else w.data_ptr(), c2, no, nv))

    def pair_symmetrize(self, r2, no, nv):
        self._check(self.lib.pccb_pair_symmetrize(self.h, dtype_code(r2), r2.data_ptr(), no, nv))

    def permute4(self, out, src_view, alpha=1.0, beta=0.0):
        """out (contiguous 4-D) = alpha * src_view (any strides) + beta * out"""
        shape = (C.c_int64 * 4)(*out.shape)
        strides = (C.c_int64 * 4)(*src_view.stride())
        self._check(self.lib.pccb_permute4(self.h, dtype_code(out), out.data_ptr(), src_view.data_ptr(), shape,
                                           strides, alpha, beta))

    def update_amplitudes(self, t1, t2, r1, r2, eps, no, nv, symmetrize):
        out = C.c_double()
        self._check(self.lib.pccb_update_amplitudes(self.h, dtype_code(t2), t1.data_ptr(), t2.data_ptr(),
                                                    r1.data_ptr(), r2.data_ptr(), eps.data_ptr(), no, nv,
                                                    1 if symmetrize else 0, C.byref(out)))
        return out.value

    def update_energy(self, t1, t2, r1, r2, eps, fov, ld_fov, oovv, no, nv, symmetrize, ij_lo, ij_hi, count_t1):
        """Fused t += r/D, increment norm and energy of the new amplitudes for the (i,j) blocks
        [ij_lo, ij_hi) of t2 (t1 is always updated in full).  Returns (rms2 part, energy part)."""
        out = (C.c_double * 2)()
        self._check(self.lib.pccb_update_energy(self.h, dtype_code(t2), t1.data_ptr(), t2.data_ptr(), r1.data_ptr(),
                                                r2.data_ptr(), eps.data_ptr(), fov.data_ptr(), ld_fov, oovv.data_ptr(),
                                                no, nv, 1 if symmetrize else 0, ij_lo, ij_hi, 1 if count_t1 else 0, out))
        return out[0], out[1]

    def prepare_amplitudes(self, t1, t2, no, nv, vlo, vl, **outs):
        """One pass over t2 writing the amplitude-dependent operands named in ``outs`` (tau, tauh, tauW, t2s,
        t11, t2sP, t2P, t2Q, tauWPn, t2Rh; see pccb_prepare_amplitudes)."""
        o = _lib.PrepareOutputs()
        for k, t in outs.items():
            if t is not None:
                setattr(o, k, t.data_ptr())
        self._check(self.lib.pccb_prepare_amplitudes(self.h, dtype_code(t2), t1.data_ptr(), t2.data_ptr(), no, nv,
                                                     vlo, vl, C.byref(o)))

    def cc_energy(self, fov, ld_fov, t1, t2, oovv, no, nv):
        out = C.c_double()
        self._check(self.lib.pccb_cc_energy(self.h, dtype_code(t2), fov.data_ptr(), ld_fov, t1.data_ptr(),
                                            t2.data_ptr(), oovv.data_ptr(), no, nv, C.byref(out)))
        return out.value

    def dot(self, x, y):
        out = C.c_double()
        self._check(self.lib.pccb_dot(self.h, dtype_code(x), x.data_ptr(), y.data_ptr(), x.numel(), C.byref(out)))
        return out.value

    def dot_many(self, x, ys):
        """[<x, y> for y in ys] in one pass over x (at most 16 vectors per launch)."""
        out = []
        for k0 in range(0, len(ys), 16):
            grp = ys[k0:k0 + 16]
            m = len(grp)
            ptrs = (C.c_void_p * m)(*[y.data_ptr() for y in grp])
            res = (C.c_double * m)()
            self._check(self.lib.pccb_dot_many(self.h, dtype_code(x), x.data_ptr(), ptrs, m, x.numel(), res))
            out += [res[k] for k in range(m)]
        return out

    def axpby(self, out, a, x, b=0.0, y=None):
        self._check(self.lib.pccb_axpby(self.h, dtype_code(out), out.data_ptr(), a, x.data_ptr(), b,
                                        None if y is None else y.data_ptr(), out.numel()))

    def lincomb(self, out, vecs, coefs):
        n = len(vecs)
        ptrs = (C.c_void_p * n)(*[v.data_ptr() for v in vecs])
        cs = (C.c_double * n)(*[float(c) for c in coefs])
        self._check(self.lib.pccb_lincomb(self.h, dtype_code(out), out.data_ptr(), ptrs, cs, n, out.numel()))

    def triples_energy(self, W, ijk_dev, ntrip, t1, t2, oovv, fov, ld_fov, eps, no, nv, et_dev):
        """W: [>=ntrip, nv, nv, nv]; ijk_dev: int32 device tensor [3*ntrip]; et_dev: float64 device tensor
        [ntrip] that receives the per-triplet energies (asynchronously)."""
        self._check(self.lib.pccb_triples_energy(self.h, dtype_code(W), W.data_ptr(), ijk_dev.data_ptr(), ntrip,
                                                 t1.data_ptr(), t2.data_ptr(), oovv.data_ptr(), fov.data_ptr(),
                                                 ld_fov, eps.data_ptr(), no, nv, et_dev.data_ptr()))

    # -- mixed-precision mode on the tcgen05 tensor cores -------------------------------------------
    def split_lo(self, x, lo):
        """lo = x - trunc_tf32(x) for an FP32 tensor (same layout)."""
        self._check(self.lib.pccb_split_lo(self.h, x.data_ptr(), lo.data_ptr(), x.numel()))

    def register_lo_plane(self, x, lo):
        """From now on FP32 contractions whose operands are views of ``x`` may use ``lo`` as its residual plane."""
        self._check(self.lib.pccb_register_lo_plane(self.h, x.data_ptr(), lo.data_ptr(), x.numel()))

    def unregister_lo_plane(self, x_ptr):
        """Forget the residual plane registered for the tensor whose base address is ``x_ptr``."""
        self._check(self.lib.pccb_unregister_lo_plane(self.h, x_ptr))

    def t3_density_combine(self, W, Dd, nk, k0, i, j, eps, no, nv, outs):
        """outs: dict with any of M, A1, A2, B1, Z3, P, Q (tensors [nk, nv, nv, nv]); see pccb_t3_density_combine."""
        ptrs = (C.c_void_p * 7)(*[(outs[k].data_ptr() if outs.get(k) is not None else None)
                                  for k in ("M", "A1", "A2", "B1", "Z3", "P", "Q")])
        self._check(self.lib.pccb_t3_density_combine(self.h, dtype_code(W), W.data_ptr(), Dd.data_ptr(), nk, k0, i, j,
                                                     eps.data_ptr(), no, nv, ptrs))

    def t3_combine(self, W, nk, k0, i, j, eps, no, nv, A1, A2):
        """A1 / A2 combinations of t3 = W / D for the pair (i, j), k = k0 .. k0+nk-1 (CC3)."""
        self._check(self.lib.pccb_t3_combine(self.h, dtype_code(W), W.data_ptr(), nk, k0, i, j, eps.data_ptr(),
                                             no, nv, A1.data_ptr(), A2.data_ptr()))
