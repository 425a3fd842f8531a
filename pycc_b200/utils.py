"""
utils.py: DIIS extrapolation and the generalized contraction operator, API-compatible with
``pycc/utils.py`` (helper_diis :6-139, cc_contract :141-187) but executing on the B200 engine.
"""
import numpy as np
import torch

from . import _lib
from .dsl import PlanError, TRef, plan


class helper_diis(object):
    """Pulay DIIS on the concatenated (t1, t2) amplitude vector; same bookkeeping as the
    reference (``pycc/utils.py:6-139``):

      * the error vector of an iteration is the difference between the new amplitudes and the
        previous (extrapolated) ones (:26-46);
      * at most ``max_diis`` error vectors are kept, oldest dropped first, and the amplitude
        list stays one entry longer so that ``vals[k+1]`` pairs with ``errs[k]`` (:54-57);
      * B is bordered with -1 / 0, its error block divided by its largest absolute entry (:103-116);
      * the new amplitudes are sum_k c_k vals[k+1] and become the reference point for the next
        error vector (:129-137).

    History vectors live in device memory; the Gram matrix entries are fixed-order FP64 dot
    products (``pccb_dot``), each computed once and cached; the (m+1)x(m+1) solve is done on the
    host in FP64 for both precisions (the reference's SP mode solves in float32, :104-105).

    Several GPUs (``comm``, SURVEY.md 8(e) "DIIS" row): the history is SLICED - rank g keeps only
    elements [g*chunk, (g+1)*chunk) of every amplitude / error vector (1/world of the memory and of
    the axpby / dot / lincomb traffic).  The new Gram entries of an iteration are partial dot products,
    summed by ONE allreduce of <= max_diis doubles; every rank then solves the same small system,
    extrapolates its slice, and one allgather rebuilds the full (t1, t2) on every rank.
    """

    def __init__(self, t1, t2, max_diis, precision='DP', backend=None, comm=None, blocks=None):
        """``blocks`` = (t2_store, block_len, nblock_per_rank) (several GPUs, CCSD engine): slice the history by
        whole (i,j) BLOCKS of t2 instead of flat chunks - rank g keeps the blocks it also updates
        (engine.update_energy) followed by t1 (real on rank 0, zeros elsewhere, so that the summed partial dot
        products count it once) - and gather the extrapolated chunks straight into ``t2_store``, the padded
        buffer t2 is a view of: no staging copy of the amplitude vector."""
        if backend is None:
            raise _lib.EngineError("helper_diis needs the engine back end that owns t1/t2")
        self.be = backend
        self.max_diis = max_diis
        self.precision = precision
        self.n1, self.n2 = t1.numel(), t2.numel()
        self.shape1, self.shape2 = tuple(t1.shape), tuple(t2.shape)
        self.dtype = t2.dtype
        self.diis_size = 0
        self.comm = comm if (comm is not None and comm.world > 1) else None
        self.blocks = blocks if self.comm is not None else None
        n = self.n1 + self.n2
        from .parallel import flat_chunk
        if self.comm is None:
            self.lo, self.hi, self.chunk = 0, n, n
        elif self.blocks is not None:
            store, blen, nper = self.blocks
            self.t2_chunk = blen * nper                        # elements of t2 per rank (padded)
            self.lo = min(self.n2, self.comm.rank * self.t2_chunk)
            self.hi = min(self.n2, self.lo + self.t2_chunk)
            self.chunk = self.t2_chunk + self.n1
            self._full = None
        else:
            w = self.comm.world
            self.lo, self.hi, self.chunk = flat_chunk(n, self.comm.rank, w)
            # gathered extrapolated vector (world * chunk >= n; the tail is padding)
            self._full = self.be.zeros((w * self.chunk,), self.dtype) if max_diis > 0 else None
        # every history vector the bookkeeping can hold at once (max_diis+2 amplitude vectors,
        # max_diis+1 error vectors, the extrapolated one) is allocated now and recycled, so no
        # device allocation happens inside the iteration.  Zero-filled: the padding of the last
        # rank's slice must not contribute to the dot products.
        self._pool = [self.be.zeros((self.chunk,), self.dtype) for _ in range(2 * max_diis + 4)] \
            if max_diis > 0 else []
        self.diis_vals = []
        self.old = None
        self.old = self._pack(t1, t2)
        first = self._alloc()
        first.copy_(self.old)
        self.diis_vals = [first]
        self.diis_errors = []
        self._gram = {}          # (id_a, id_b) -> <e_a, e_b>
        self._ids = []           # identity tags of the error vectors, parallel to diis_errors
        self._next_id = 0

    def _alloc(self):
        return self._pool.pop() if self._pool else self.be.zeros((self.chunk,), self.dtype)

    def _release(self, v):
        if v is not None and not any(v is x for x in self.diis_vals) and v is not self.old:
            self._pool.append(v)

    def _pack(self, t1, t2):
        """This rank's slice [lo, hi) of the concatenated (t1, t2) vector."""
        v = self._alloc()
        lo, hi, n1 = self.lo, self.hi, self.n1
        if self.blocks is not None:      # [ this rank's (i,j) blocks of t2 (zero padding) | t1 on rank 0 ]
            v[:hi - lo].copy_(t2.reshape(-1)[lo:hi])
            if self.comm.rank == 0:
                v[self.t2_chunk:].copy_(t1.reshape(-1))
            return v
        if lo < n1:
            e = min(hi, n1)
            v[:e - lo].copy_(t1.reshape(-1)[lo:e])
        if hi > n1:
            b = max(lo, n1)
            v[b - lo:hi - lo].copy_(t2.reshape(-1)[b - n1:hi - n1])
        return v

    def add_error_vector(self, t1, t2):
        if self.max_diis == 0:
            return        # the history is never read (extrapolate returns at once, utils.py:50-51)
        cur = self._pack(t1, t2)
        self.diis_vals.append(cur)
        err = self._alloc()
        self.be.bind_stream()
        self.be.axpby(err, 1.0, cur, -1.0, self.old)
        self.diis_errors.append(err)
        self._ids.append(self._next_id)
        self._next_id += 1
        prev, self.old = self.old, cur        # cur is never modified afterwards
        self._release(prev)

    def extrapolate(self, t1, t2):
        if self.max_diis == 0:
            return t1, t2
        if len(self.diis_errors) > self.max_diis:
            v0, e0 = self.diis_vals.pop(0), self.diis_errors.pop(0)
            self._pool.append(e0)
            self._release(v0)
            dead = self._ids.pop(0)
            self._gram = {k: v for k, v in self._gram.items() if dead not in k}
        self.diis_size = m = len(self.diis_errors)
        self.be.bind_stream()
        # Gram entries not seen before (in steady state: the m products with the newest error vector)
        missing = [(p, q) for p in range(m) for q in range(p, m) if (self._ids[p], self._ids[q]) not in self._gram]
        # grouped by their second vector: in steady state ALL missing entries pair the newest error vector with the
        # others, i.e. one pccb_dot_many launch and one host synchronisation per DIIS step
        vals_by = {}
        for q in sorted(set(q for (_, q) in missing)):
            ps = [p for (p, qq) in missing if qq == q]
            for p, x in zip(ps, self.be.dot_many(self.diis_errors[q], [self.diis_errors[p] for p in ps])):
                vals_by[(p, q)] = x
        vals = [vals_by[pq] for pq in missing]
        if self.comm is not None and vals:
            vals = self.comm.allreduce_scalars(vals)      # partial dot products of the slices
        for (p, q), x in zip(missing, vals):
            self._gram[(self._ids[p], self._ids[q])] = x
        B = -np.ones((m + 1, m + 1))
        B[-1, -1] = 0.0
        for p in range(m):
            for q in range(p, m):
                B[p, q] = B[q, p] = self._gram[(self._ids[p], self._ids[q])]
        B[:-1, :-1] /= np.abs(B[:-1, :-1]).max()
        resid = np.zeros(m + 1)
        resid[-1] = -1.0
        ci = np.linalg.solve(B, resid)
        # sum_k c_k vals[k+1]; pccb_lincomb takes at most 16 vectors per launch, so a longer history
        # (max_diis > 16, or a late start_diis) is accumulated in groups: the running sum enters the
        # next group as one more vector with coefficient 1
        new = self._alloc()
        G = 16
        self.be.lincomb(new, [self.diis_vals[k + 1] for k in range(min(m, G))], [ci[k] for k in range(min(m, G))])
        k0 = G
        while k0 < m:
            k1 = min(m, k0 + G - 1)
            nxt = self._alloc()
            self.be.lincomb(nxt, [new] + [self.diis_vals[k + 1] for k in range(k0, k1)],
                            [1.0] + [ci[k] for k in range(k0, k1)])
            self._pool.append(new)
            new, k0 = nxt, k1
        prev, self.old = self.old, new
        self._release(prev)
        # the caller's amplitude tensors are updated in place (they are the engine's buffers)
        if self.blocks is not None:
            # chunks of t2 gathered straight into the padded buffer behind t2; t1 lives on rank 0's vector
            # (zeros elsewhere): a sum over the ranks hands it to everybody
            self.comm.allgather_into(self.blocks[0], new[:self.t2_chunk])
            t1.copy_(new[self.t2_chunk:].reshape(self.shape1))
            self.comm.allreduce_(t1)
            return t1, t2
        if self.comm is None:
            full = new
        else:
            self.comm.allgather_into(self._full, new)
            full = self._full
        t1.copy_(full[:self.n1].reshape(self.shape1))
        t2.copy_(full[self.n1:self.n1 + self.n2].reshape(self.shape2))
        return t1, t2


class cc_contract(object):
    """``contract = cc_contract(device)``; ``contract(subscripts, *operands)`` with numpy.einsum
    string semantics (``pycc/utils.py:141-187``).  Two-operand contractions run as one
    ``pccb_contract`` launch on the operands' existing strides; NumPy operands are moved to the
    device first (the reference's GPU mode does the same, :181-183).  Returns a new tensor."""

    def __init__(self, device='GPU', backend=None):
        self.device = device
        self.be = backend
        if backend is not None:
            self.device0 = torch.device('cpu')
            self.device1 = backend.device

    def _backend(self):
        if self.be is None:
            from .backend import CudaBackend
            self.be = CudaBackend()
            self.device0 = torch.device('cpu')
            self.device1 = self.be.device
        return self.be

    def _complex_call(self, spec, ins, out_idx, ops):
        """Complex operands (the real-time code of the reference contracts complex amplitudes through the same
        operator, pycc/rt/rtcc.py): the real and imaginary planes go through the real kernels,
        (Ar + i Ai)(Br + i Bi) = (Ar Br - Ai Bi) + i (Ar Bi + Ai Br); a real operand has no imaginary plane."""
        if len(ops) == 1:
            perm = [ins[0].index(c) for c in out_idx]
            if len(perm) != len(ins[0]):
                raise ValueError("cc_contract: single-operand reductions are not supported: %r" % spec)
            return ops[0].permute(perm).contiguous()
        parts = [(x.real.contiguous(), x.imag.contiguous()) if x.is_complex() else (x, None) for x in ops]
        (ar, ai), (br, bi) = parts
        re = self(spec, ar, br)
        if ai is not None and bi is not None:
            re = re - self(spec, ai, bi)
        im = None
        if bi is not None:
            im = self(spec, ar, bi)
        if ai is not None:
            x = self(spec, ai, br)
            im = x if im is None else im + x
        return torch.complex(re, im)

    def __call__(self, subscripts, *operands):
        be = self._backend()
        spec = subscripts.replace(" ", "")
        if "->" not in spec:
            raise ValueError("cc_contract needs an explicit output ('->') in %r" % subscripts)
        lhs, out_idx = spec.split("->")
        ins = lhs.split(",")
        if len(ins) != len(operands):
            raise ValueError("einsum string %r does not match %d operands" % (subscripts, len(operands)))
        ops = []
        for x in operands:
            if not isinstance(x, torch.Tensor):
                x = torch.as_tensor(np.asarray(x))
            if x.device != be.device:
                x = x.to(be.device)
            ops.append(x)
        if any(x.is_complex() for x in ops) and len(ops) <= 2:
            return self._complex_call(spec, ins, out_idx, ops)
        dtype = next((x.dtype for x in ops if x.is_complex()), ops[0].dtype)
        if dtype not in (torch.float64, torch.float32, torch.complex128, torch.complex64):
            raise ValueError("cc_contract: float64 / float32 (or complex128 / complex64) operands only (got %s)" % dtype)
        if not dtype.is_complex:
            ops = [x if x.dtype == dtype else x.to(dtype) for x in ops]
        dcode = _lib.F64 if dtype in (torch.float64, torch.complex128) else _lib.F32
        if len(ops) == 1:
            perm = [ins[0].index(c) for c in out_idx]
            if len(perm) != len(ins[0]):
                raise ValueError("cc_contract: single-operand reductions are not supported: %r" % subscripts)
            return ops[0].permute(perm).contiguous()
        if len(ops) != 2:
            # chain left to right, keeping every index still needed by the output or a later operand
            cur, cur_idx = ops[0], ins[0]
            for n in range(1, len(ops)):
                later = set(out_idx).union(*[set(s) for s in ins[n + 1:]]) if n + 1 < len(ops) else set(out_idx)
                keep = "".join(c for c in dict.fromkeys(cur_idx + ins[n]) if c in later)
                cur = self("%s,%s->%s" % (cur_idx, ins[n], keep), cur, ops[n])
                cur_idx = keep
            return cur
        ext = {}
        for idx, x in zip(ins, ops):
            if len(idx) != x.dim():
                raise ValueError("operand has %d dimensions but subscripts %r" % (x.dim(), idx))
            for c, n in zip(idx, x.shape):
                if ext.setdefault(c, n) != n:
                    raise ValueError("size mismatch for index %r in %r" % (c, subscripts))
        for c in out_idx:
            if c not in ext:
                raise ValueError("output index %r does not appear in the operands (%r)" % (c, subscripts))
        # indices summed over a single operand are reduced first (rare; not on the CCSD path)
        for n in (0, 1):
            other = ins[1 - n]
            lone = [c for c in ins[n] if c not in other and c not in out_idx]
            if lone:
                ops[n] = ops[n].sum(dim=[ins[n].index(c) for c in lone])
                ins[n] = "".join(c for c in ins[n] if c not in lone)
        out = be.empty(tuple(ext[c] for c in out_idx), dtype)
        if out_idx == "":
            out = be.empty((1,), dtype)
            T = {"C": TRef(out.data_ptr(), (), (), out.element_size(), "C", out), "A": ops[0], "B": ops[1]}
        else:
            T = {"C": out, "A": ops[0], "B": ops[1]}
        be.bind_stream()
        try:
            launches = plan("C[%s] = A[%s]*B[%s]" % (out_idx, ins[0], ins[1]), T, dcode)
        except PlanError:
            # an index group that does not fit the descriptor's two stride levels (e.g. 'iajb,ijab->'): contract
            # contiguous copies in the canonical orders [batch, M, K] x [batch, N, K] -> [batch, M, N] instead
            Bt = [c for c in ins[0] if c in ins[1] and c in out_idx]
            K = [c for c in ins[0] if c in ins[1] and c not in out_idx]
            M = [c for c in ins[0] if c not in ins[1]]
            N = [c for c in ins[1] if c not in ins[0]]
            a_idx, b_idx, c_idx = "".join(Bt + M + K), "".join(Bt + N + K), "".join(Bt + M + N)
            a = ops[0].permute([ins[0].index(c) for c in a_idx]).contiguous()
            b = ops[1].permute([ins[1].index(c) for c in b_idx]).contiguous()
            if (a_idx, b_idx, c_idx) == (ins[0], ins[1], out_idx):
                raise
            res = self("%s,%s->%s" % (a_idx, b_idx, c_idx), a, b)
            return res if c_idx == out_idx else res.permute([c_idx.index(c) for c in out_idx]).contiguous()
        for L in launches:
            be.contract(L.desc)
        return out.reshape(()) if out_idx == "" else out
