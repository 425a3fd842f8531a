"""
Planner: einsum-style statements  ->  pccb_contraction descriptors (include/pyccb200.h).

The reference expresses every term of the CCSD residual as ``contract('imae,mbej->ijab', X, Y)``
(pycc/ccwfn.py:413-622) and lets opt_einsum transpose/reshape operands into a GEMM.  Here a
statement such as

    r2[ijab] += 0.5 * tau[ijef] * vvvv[abef] + t2a[imae] * Wmbej[mbej]

is planned ONCE into strided-GEMM descriptors: the einsum index groups become the M / N / K /
batch dimensions of ``pccb_contract``, each described by a two-level compound index over the
operands' *existing* strides, so no operand or result is ever permuted in memory.  Terms with
the same M/N structure are K-concatenated into one launch (up to 4 segments).

Pure host logic (no device access): ``tests/test_plan_cpu.py`` interprets the emitted
descriptors with NumPy and checks them against the reference's results.
"""
import itertools
import re

from . import _lib

FLAT = 0  # pccb_cidx.div <= 0 means "flat index"


class PlanError(ValueError):
    pass


class TRef(object):
    """A strided view of device (or, in host-logic tests, host) memory: pointer, shape, strides in
    elements.  ``keep`` pins the owning tensor."""
    __slots__ = ("ptr", "shape", "strides", "itemsize", "name", "keep")

    def __init__(self, ptr, shape, strides, itemsize, name=None, keep=None):
        self.ptr, self.shape, self.strides = int(ptr), tuple(int(s) for s in shape), tuple(int(s) for s in strides)
        self.itemsize, self.name, self.keep = int(itemsize), name, keep

    @staticmethod
    def of(t, name=None):
        """From a torch tensor (any device, any strides)."""
        return TRef(t.data_ptr(), t.shape, t.stride(), t.element_size(), name, t)


_FACTOR = re.compile(r"\s*([A-Za-z_][A-Za-z_0-9]*)\s*\[\s*([A-Za-z]*)\s*\]\s*")
_NUMBER = re.compile(r"\s*([0-9]*\.?[0-9]+(?:[eE][-+]?[0-9]+)?)\s*")


def parse(stmt):
    """'out[idx] (=|+=|-=) [c *] X[idx] [* Y[idx]] (+|-) ...'  ->  (out_name, out_idx, accumulate, terms)
    with terms = [(coef, [(name, idx), ...]), ...] (one or two factors per term)."""
    m = re.match(r"^\s*([A-Za-z_][A-Za-z_0-9]*)\s*\[\s*([A-Za-z]*)\s*\]\s*(\+=|-=|=)(.*)$", stmt, re.S)
    if not m:
        raise PlanError("cannot parse statement: %r" % stmt)
    out_name, out_idx, op, rhs = m.group(1), m.group(2), m.group(3), m.group(4)
    terms = []
    pos, sign = 0, 1.0
    rhs = rhs.strip()
    if rhs.startswith("-"):
        sign, rhs = -1.0, rhs[1:]
    elif rhs.startswith("+"):
        rhs = rhs[1:]
    while True:
        coef = sign
        mnum = _NUMBER.match(rhs, pos)
        if mnum:
            coef *= float(mnum.group(1))
            pos = mnum.end()
            if pos < len(rhs) and rhs[pos] == "*":
                pos += 1
            else:
                raise PlanError("expected '*' after coefficient in %r" % stmt)
        factors = []
        while True:
            mf = _FACTOR.match(rhs, pos)
            if not mf:
                raise PlanError("expected tensor factor at ...%r in %r" % (rhs[pos:pos + 20], stmt))
            factors.append((mf.group(1), mf.group(2)))
            pos = mf.end()
            if pos < len(rhs) and rhs[pos] == "*":
                pos += 1
                continue
            break
        if len(factors) > 2:
            raise PlanError("at most two tensor factors per term: %r" % stmt)
        terms.append((coef, factors))
        if pos >= len(rhs):
            break
        if rhs[pos] == "+":
            sign = 1.0
        elif rhs[pos] == "-":
            sign = -1.0
        else:
            raise PlanError("unexpected %r in %r" % (rhs[pos], stmt))
        pos += 1
    if op == "-=":
        terms = [(-c, f) for (c, f) in terms]
    return out_name, out_idx, op != "=", terms


def _view(ref, idx, ext):
    if len(idx) != len(ref.shape):
        raise PlanError("%s has rank %d but indices %r" % (ref.name, len(ref.shape), idx))
    if len(set(idx)) != len(idx):
        raise PlanError("repeated index in %s[%s]" % (ref.name, idx))
    v = {}
    for letter, n, s in zip(idx, ref.shape, ref.strides):
        if letter in ext and ext[letter] != n:
            raise PlanError("extent mismatch for index %r (%d vs %d) in %s" % (letter, ext[letter], n, ref.name))
        ext[letter] = n
        v[letter] = s
    return v


def _flatten(group, views, ext, prefer_inner, prefer_flat=False):
    """Order the letters of one index group so that it is a <=2-level compound index in every
    view.  Returns (total_extent, inner_extent_or_FLAT, [(s_hi, s_lo) per view]).
    ``prefer_flat``: among otherwise equal orders take the one whose level break is a real break in the
    fewest views, i.e. that leaves the most operands plain matrices (the FP32 tcgen05 kernel takes
    plain-matrix operands only; without it ties fall to the alphabetical order)."""
    letters = sorted(x for x in group if ext[x] > 1)
    total = 1
    for x in group:
        total *= ext[x]
    if not letters:
        return total, FLAT, [(0, 0) for _ in views]
    best = None
    for perm in itertools.permutations(letters):
        breaks, nbroken = [], 0
        for q in range(len(perm) - 1):
            x, y = perm[q], perm[q + 1]
            bad = sum(1 for v in views if v.get(x, 0) != v.get(y, 0) * ext[y])
            if bad:
                breaks.append(q)
                nbroken += bad
        if len(breaks) > 1:
            continue
        rank = prefer_inner.index(perm[-1]) if perm[-1] in prefer_inner else len(prefer_inner)
        score = (len(breaks), rank, nbroken if prefer_flat else 0, perm)
        if best is None or score < best[0]:
            best = (score, perm, breaks)
    if best is None:
        raise PlanError("index group %r needs more than two levels" % (group,))
    _, perm, breaks = best
    if not breaks:
        return total, FLAT, [(0, v.get(perm[-1], 0)) for v in views]
    q = breaks[0]
    inner = 1
    for x in perm[q + 1:]:
        inner *= ext[x]
    return total, inner, [(v.get(perm[q], 0), v.get(perm[-1], 0)) for v in views]


def _cidx(div, s):
    return _lib.CIdx(div, s[0], s[1])


class Launch(object):
    """One pccb_contract call: the ctypes descriptor plus the tensors it borrows."""

    def __init__(self, desc, keep, text):
        self.desc, self.keep, self.text = desc, keep, text


def _roles(fx, fy, out_set):
    """Index sets for A=fx, B=fy."""
    sa, sb = set(fx[1]), set(fy[1])
    batch = sa & sb & out_set
    M = (sa & out_set) - sb
    N = (sb & out_set) - sa
    K = (sa & sb) - out_set
    stray = (sa | sb) - out_set - K
    return M, N, K, batch, stray


def plan(stmt, tensors, dtype=_lib.F64, prefer_flat=None):
    """Plan one statement into a list of Launch objects (usually one).
    ``tensors`` maps names to TRef (or torch tensors)."""
    out_name, out_idx, accumulate, terms = parse(stmt)
    if prefer_flat is None:
        prefer_flat = dtype == _lib.F32

    def ref(name):
        if name not in tensors:
            raise PlanError("unknown tensor %r in %r" % (name, stmt))
        t = tensors[name]
        return t if isinstance(t, TRef) else TRef.of(t, name)

    out = ref(out_name)
    out_set = set(out_idx)
    products = [(c, f) for (c, f) in terms if len(f) == 2]
    addends = [(c, f[0]) for (c, f) in terms if len(f) == 1]

    # ---- group product terms into launches with identical (M, N, batch) structure -------------
    groups = []   # each: dict(M,N,batch, segs=[(coef, A, B, K)])
    smallest = min(zip(out.strides, out_idx))[1] if out_idx else None
    for coef, (fx, fy) in products:
        placed = False
        cands = []
        for (fa, fb) in ((fx, fy), (fy, fx)):
            M, N, K, batch, stray = _roles(fa, fb, out_set)
            if stray or (M | N | batch) != out_set:
                continue
            cands.append((fa, fb, frozenset(M), frozenset(N), frozenset(K), frozenset(batch)))
        if not cands:
            raise PlanError("term %s*%s is not a pairwise contraction onto %s[%s]" % (fx, fy, out_name, out_idx))
        for g in groups:
            for (fa, fb, M, N, K, batch) in cands:
                if M == g["M"] and N == g["N"] and batch == g["batch"] and len(g["segs"]) < _lib.MAX_SEGMENTS:
                    g["segs"].append((coef, fa, fb, K))
                    placed = True
                    break
            if placed:
                break
        if not placed:
            # first term of a new launch: let N carry the output's fastest index (coalesced stores)
            # tile waste first (the kernel's CTA tile is (16*MI) x 128, MI in {1,2,4,5,6,8}: a
            # 150-wide index belongs on M as 2 x 80, not on N as 2 x 128), then coalesced stores
            letters = {}
            for (nm, idx) in (fx, fy, (out_name, out_idx)):
                for letter, n in zip(idx, ref(nm).shape):
                    letters[letter] = n

            def padded(c):
                Mx = Nx = 1
                for x in c[2]:
                    Mx *= letters[x]
                for x in c[3]:
                    Nx *= letters[x]
                pm = min(-(-Mx // bm) * bm for bm in (128, 96, 80, 64, 32, 16))
                return pm * (-(-Nx // 128) * 128)

            best = min(padded(c) for c in cands)
            cands.sort(key=lambda c: (0 if padded(c) <= 1.03 * best else 1, 0 if smallest in c[3] else 1))
            fa, fb, M, N, K, batch = cands[0]
            groups.append(dict(M=M, N=N, batch=batch, segs=[(coef, fa, fb, K)]))

    launches = []
    first = True
    pending_addends = list(addends)

    def emit(g):
        nonlocal first
        ext = {}
        keep = [out]
        cv = _view(out, out_idx, ext)
        d = _lib.Contraction()
        d.dtype = dtype
        dref, dv, gamma = None, None, 0.0
        if pending_addends:
            gamma, (dname, didx) = pending_addends.pop(0)
            dref = ref(dname)
            if set(didx) != out_set:
                raise PlanError("addend %s[%s] does not match output indices %r" % (dname, didx, out_idx))
            dv = _view(dref, didx, ext)
            keep.append(dref)
        segs = []
        if g is not None:
            for (coef, fa, fb, K) in g["segs"]:
                ra, rb = ref(fa[0]), ref(fb[0])
                segs.append((coef, ra, _view(ra, fa[1], ext), rb, _view(rb, fb[1], ext), K))
                keep += [ra, rb]
            Mset, Nset, Bset = g["M"], g["N"], g["batch"]
        else:
            # epilogue-only launch: N takes the output's fastest index, M the rest
            Nset = frozenset([smallest]) if smallest is not None else frozenset()
            Mset, Bset = frozenset(out_set - Nset), frozenset()
        mviews = [s[2] for s in segs] + [cv] + ([dv] if dv else [])
        nviews = [s[4] for s in segs] + [cv] + ([dv] if dv else [])
        pref_m = [x for s in segs for x in Mset if s[2][x] == 1] + [x for x in Mset if cv[x] == 1]
        pref_n = [x for s in segs for x in Nset if s[4][x] == 1] + [x for x in Nset if cv[x] == 1]
        bviews = [s[2] for s in segs] + [s[4] for s in segs] + [cv] + ([dv] if dv else [])

        def flatten_all(Ms, Ns, Bs):
            Mt, md, ms_ = _flatten(Ms, mviews, ext, pref_m, prefer_flat)
            Nt, nd, ns_ = _flatten(Ns, nviews, ext, pref_n, prefer_flat)
            Bt, bd, bs_ = _flatten(Bs, bviews, ext, [])
            if bd != FLAT:
                raise PlanError("batch indices %r do not flatten to one stride" % (sorted(Bs),))
            return Mt, md, ms_, Nt, nd, ns_, Bt, bs_

        # An index group that needs three levels gives its slowest output index to the batch
        # dimension (operands that do not carry it get batch stride 0).
        trials = [(Mset, Nset, Bset)]
        for x in sorted(Mset, key=lambda l: -cv[l]):
            trials.append((Mset - {x}, Nset, Bset | {x}))
        for x in sorted(Nset, key=lambda l: -cv[l]):
            trials.append((Mset, Nset - {x}, Bset | {x}))
        for x in sorted(Mset, key=lambda l: -cv[l]):
            for y in sorted(Nset, key=lambda l: -cv[l]):
                trials.append((Mset - {x}, Nset - {y}, Bset | {x, y}))
        err = None
        for (Ms, Ns, Bs) in trials:
            try:
                Mtot, mdiv, ms, Ntot, ndiv, nsr, Btot, bs = flatten_all(Ms, Ns, Bs)
                break
            except PlanError as e:
                err = err or e
        else:
            raise PlanError("%s  [while planning %r]" % (err, stmt))
        d.M, d.N, d.batch = Mtot, Ntot, Btot
        d.nseg = len(segs)
        nsg = len(segs)
        for q, (coef, ra, va, rb, vb, K) in enumerate(segs):
            sg = d.seg[q]
            pref_k = [x for x in K if va[x] == 1] + [x for x in K if vb[x] == 1]
            Ktot, kdiv, ks = _flatten(K, [va, vb], ext, pref_k)
            sg.A_dev, sg.B_dev = ra.ptr, rb.ptr
            sg.am, sg.ak = _cidx(mdiv, ms[q]), _cidx(kdiv, ks[0])
            sg.bn, sg.bk = _cidx(ndiv, nsr[q]), _cidx(kdiv, ks[1])
            sg.K = Ktot
            sg.a_batch, sg.b_batch = bs[q][1], bs[nsg + q][1]
            sg.scale = coef
        d.C_dev = out.ptr
        d.cm, d.cn, d.c_batch = _cidx(mdiv, ms[nsg]), _cidx(ndiv, nsr[nsg]), bs[2 * nsg][1]
        if dref is not None:
            d.D_dev = dref.ptr
            d.dm, d.dn, d.d_batch = _cidx(mdiv, ms[nsg + 1]), _cidx(ndiv, nsr[nsg + 1]), bs[2 * nsg + 1][1]
            d.gamma = gamma
        d.alpha = 1.0
        d.beta = 1.0 if (accumulate or not first) else 0.0
        first = False
        launches.append(Launch(d, keep, stmt))

    for g in groups:
        emit(g)
    while pending_addends:
        emit(None)
    if not launches:
        raise PlanError("empty statement %r" % stmt)
    return launches
