"""
ORACLE (test infrastructure, not product code): CPU/NumPy restatement of the reference's
RHF-CCSD amplitude iteration, working from the six symmetry-distinct Dirac ERI blocks instead
of the reference's full ``n^4`` ``ERI`` and ``L`` arrays.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import this package.  The engine (``pycc_b200``) never does.

Parity status: PINNED.  ``tests/test_oracle_golden.py`` checks every function here against
``tests/golden/*.npz``, which were produced by running the reference's own, unmodified
``pycc/ccwfn.py`` / ``pycc/cctriples.py`` / ``pycc/utils.py`` on identical synthetic inputs
(generator: ``tests/golden/make_golden.py``).

Blocks (all C-contiguous, see ``pycc_b200/synthetic.py``):
    oooo[m,n,i,j] = <mn|ij>   ooov[m,n,i,e] = <mn|ie>   oovv[i,j,a,b] = <ij|ab>
    ovov[m,b,j,e] = <mb|je>   ovvv[m,b,e,f] = <mb|ef>   vvvv[a,b,e,f] = <ab|ef>
Every other slice the reference takes out of ``ERI``/``L`` follows from the 8-fold symmetry of
real orbitals, <pq|rs> = <qp|sr> = <rs|pq> = <ps|rq> (= (pr|qs) in Mulliken order):
    ERI[o,o,v,o][m,n,e,j] = ooov[n,m,j,e]       ERI[o,v,v,o][m,b,e,j] = oovv[m,j,e,b]
    ERI[v,v,v,o][a,b,e,j] = ovvv[j,e,b,a]       ERI[o,v,o,o][m,b,i,j] = ooov[i,j,m,b]
    L = 2<pq|rs> - <pq|sr>                      (pycc/hamiltonian.py:39)
"""
import numpy as np


def _ein(spec, *ops):
    return np.einsum(spec, *ops, optimize=True)


class Blocks(object):
    """The reference's ERI/L slices expressed through the six stored blocks."""

    def __init__(self, blocks):
        self.oooo = blocks["oooo"]
        self.ooov = blocks["ooov"]
        self.oovv = blocks["oovv"]
        self.ovov = blocks["ovov"]
        self.ovvv = blocks["ovvv"]
        self.vvvv = blocks.get("vvvv")
        self.no = self.oovv.shape[0]
        self.nv = self.oovv.shape[2]
        # bench.py, shapes whose <ab|ef> block does not fit the host: ``vvvv`` may be a callable
        # (lo, hi) -> rows a in [lo, hi); the ladder then runs chunk by chunk (same einsum per chunk) and
        # the seconds spent manufacturing the rows are kept apart in ``gen_seconds`` (they are set-up)
        self.vvvv_chunk = blocks.get("vvvv_chunk", 16)
        self.gen_seconds = 0.0

    # ERI slices in the reference's index order
    def oovo(self):  # <mn|ej>
        return self.ooov.transpose(1, 0, 3, 2)

    def ovvo(self):  # <mb|ej>
        return self.oovv.transpose(0, 3, 2, 1)

    def vvvo(self):  # <ab|ej>
        return self.ovvv.transpose(3, 2, 1, 0)

    def ovoo(self):  # <mb|ij>
        return self.ooov.transpose(2, 3, 0, 1)

    # spin-adapted combinations
    def L_oovv(self):
        return 2.0 * self.oovv - self.oovv.transpose(0, 1, 3, 2)

    def L_ovvv(self):
        return 2.0 * self.ovvv - self.ovvv.transpose(0, 1, 3, 2)

    def L_ooov(self):
        return 2.0 * self.ooov - self.ooov.transpose(1, 0, 2, 3)

    def L_ovvo(self):  # L[n,a,f,i] = 2<na|fi> - <na|if>
        return 2.0 * self.ovvo() - self.ovov.transpose(0, 1, 3, 2)

    def L_oovo(self):  # L[n,m,e,i] = 2<nm|ei> - <nm|ie>
        return 2.0 * self.oovo() - self.ooov.transpose(0, 1, 3, 2)


def denominators(F, no):
    """Dia, Dijab from the Fock diagonal (pycc/ccwfn.py:164-167)."""
    eps = np.diag(F)
    eo, ev = eps[:no], eps[no:]
    Dia = eo[:, None] - ev[None, :]
    Dijab = eo[:, None, None, None] + eo[None, :, None, None] - ev[None, None, :, None] - ev[None, None, None, :]
    return Dia, Dijab


def tau(t1, t2, f1=1.0, f2=1.0):
    """build_tau (pycc/ccwfn.py:408-410)."""
    return f1 * t2 + f2 * t1[:, None, :, None] * t1[None, :, None, :]


def build_Fae(F, blk, t1, t2):
    """pycc/ccwfn.py:421-428 (CCSD branch)."""
    no = blk.no
    Fov, Fvv = F[:no, no:], F[no:, no:]
    out = Fvv - 0.5 * (t1.T @ Fov)            # [a,e] = -1/2 sum_m t_ma F_me
    out = out + _ein('mf,mafe->ae', t1, blk.L_ovvv())
    out = out - _ein('mnaf,mnef->ae', tau(t1, t2, 1.0, 0.5), blk.L_oovv())
    return out


def build_Fmi(F, blk, t1, t2):
    """pycc/ccwfn.py:441-447."""
    no = blk.no
    Foo, Fov = F[:no, :no], F[:no, no:]
    out = Foo + 0.5 * (Fov @ t1.T)            # [m,i] = 1/2 sum_e F_me t_ie
    out = out + _ein('ne,mnie->mi', t1, blk.L_ooov())
    out = out + _ein('inef,mnef->mi', tau(t1, t2, 1.0, 0.5), blk.L_oovv())
    return out


def build_Fme(F, blk, t1):
    """pycc/ccwfn.py:456-460."""
    no = blk.no
    return F[:no, no:] + _ein('nf,mnef->me', t1, blk.L_oovv())


def build_Wmnij(blk, t1, t2):
    """pycc/ccwfn.py:473-482."""
    W = blk.oooo + _ein('je,mnie->mnij', t1, blk.ooov)
    W = W + _ein('ie,mnej->mnij', t1, blk.oovo())
    W = W + _ein('ijef,mnef->mnij', tau(t1, t2), blk.oovv)
    return W


def build_Wmbej(blk, t1, t2):
    """pycc/ccwfn.py:498-505."""
    W = blk.ovvo() + _ein('jf,mbef->mbej', t1, blk.ovvv)
    W = W - _ein('nb,mnej->mbej', t1, blk.oovo())
    W = W - _ein('jnfb,mnef->mbej', tau(t1, t2, 0.5, 1.0), blk.oovv)
    W = W + 0.5 * _ein('njfb,mnef->mbej', t2, blk.L_oovv())
    return W


def build_Wmbje(blk, t1, t2):
    """pycc/ccwfn.py:520-526."""
    W = -1.0 * blk.ovov - _ein('jf,mbfe->mbje', t1, blk.ovvv)
    W = W + _ein('nb,mnje->mbje', t1, blk.ooov)
    W = W + _ein('jnfb,mnfe->mbje', tau(t1, t2, 0.5, 1.0), blk.oovv)
    return W


def build_Zmbij(blk, t1, t2):
    """pycc/ccwfn.py:537."""
    return _ein('mbef,ijef->mbij', blk.ovvv, tau(t1, t2))


def r_T1(F, blk, t1, t2, Fae, Fme, Fmi):
    """pycc/ccwfn.py:548-557."""
    no = blk.no
    t2s = 2.0 * t2 - t2.transpose(0, 1, 3, 2)
    r = F[:no, no:] + t1 @ Fae.T
    r = r - Fmi.T @ t1
    r = r + _ein('imae,me->ia', t2s, Fme)
    r = r + _ein('nf,nafi->ia', t1, blk.L_ovvo())
    r = r + _ein('mief,maef->ia', t2s, blk.ovvv)
    r = r - _ein('mnae,nmei->ia', t2, blk.L_oovo())
    return r


def ladder(blk, tt):
    """sum_ef tau[ijef] <ab|ef> (pycc/ccwfn.py:607), one einsum - or, with provided rows, one per chunk."""
    if blk.vvvv is None:          # bench.py's bounded CPU sample times the ladder separately, on a slice
        return 0.0
    if not callable(blk.vvvv):
        return _ein('ijef,abef->ijab', tt, blk.vvvv)
    import time
    out = np.empty_like(tt)
    for lo in range(0, blk.nv, blk.vvvv_chunk):
        hi = min(blk.nv, lo + blk.vvvv_chunk)
        t0 = time.perf_counter()
        rows = blk.vvvv(lo, hi)
        blk.gen_seconds += time.perf_counter() - t0
        out[:, :, lo:hi, :] = _ein('ijef,abef->ijab', tt, rows)
        del rows
    return out


def r_T2(F, blk, t1, t2, Fae, Fme, Fmi, Wmnij, Wmbej, Wmbje, Zmbij):
    """pycc/ccwfn.py:596-621 (CCSD branch, including the final pair symmetrisation)."""
    tt = tau(t1, t2)
    r = 0.5 * blk.oovv
    r = r + _ein('ijae,be->ijab', t2, Fae)
    r = r - 0.5 * _ein('ijae,be->ijab', t2, _ein('mb,me->be', t1, Fme))
    r = r - _ein('imab,mj->ijab', t2, Fmi)
    r = r - 0.5 * _ein('imab,jm->ijab', t2, _ein('je,me->jm', t1, Fme))
    r = r + 0.5 * _ein('mnab,mnij->ijab', tt, Wmnij)
    r = r + 0.5 * ladder(blk, tt)
    r = r - _ein('ma,mbij->ijab', t1, Zmbij)
    r = r + _ein('imae,mbej->ijab', t2 - t2.transpose(0, 1, 3, 2), Wmbej)
    r = r + _ein('imae,mbej->ijab', t2, Wmbej + Wmbje.transpose(0, 1, 3, 2))
    r = r + _ein('mjae,mbie->ijab', t2, Wmbje)
    tmp = t1[:, None, :, None] * t1[None, :, None, :]      # tmp[i,m,e,a] = t_ie t_ma
    r = r - _ein('imea,mbej->ijab', tmp, blk.ovvo())
    r = r - _ein('imeb,maje->ijab', tmp, blk.ovov)
    r = r + _ein('ie,abej->ijab', t1, blk.vvvo())
    r = r - _ein('ma,mbij->ijab', t1, blk.ovoo())
    return r + r.transpose(1, 0, 3, 2)


def residuals(F, blk, t1, t2, want_intermediates=False):
    """ccwfn.residuals, CCSD/CCSD(T) branch (pycc/ccwfn.py:357-366)."""
    Fae = build_Fae(F, blk, t1, t2)
    Fmi = build_Fmi(F, blk, t1, t2)
    Fme = build_Fme(F, blk, t1)
    Wmnij = build_Wmnij(blk, t1, t2)
    Wmbej = build_Wmbej(blk, t1, t2)
    Wmbje = build_Wmbje(blk, t1, t2)
    Zmbij = build_Zmbij(blk, t1, t2)
    r1 = r_T1(F, blk, t1, t2, Fae, Fme, Fmi)
    r2 = r_T2(F, blk, t1, t2, Fae, Fme, Fmi, Wmnij, Wmbej, Wmbje, Zmbij)
    if want_intermediates:
        return r1, r2, dict(Fae=Fae, Fmi=Fmi, Fme=Fme, Wmnij=Wmnij, Wmbej=Wmbej, Wmbje=Wmbje, Zmbij=Zmbij)
    return r1, r2


def cc_energy(F, blk, t1, t2):
    """pycc/ccwfn.py:717-718."""
    no = blk.no
    return 2.0 * np.sum(F[:no, no:] * t1) + np.sum(tau(t1, t2) * blk.L_oovv())


class Diis(object):
    """helper_diis (pycc/utils.py:6-139), NumPy branch; same bookkeeping, same B-matrix scaling."""

    def __init__(self, t1, t2, max_diis):
        self.old = np.concatenate((t1.ravel(), t2.ravel()))
        self.vals = [self.old.copy()]
        self.errs = []
        self.max_diis = max_diis
        self.shape1, self.shape2 = t1.shape, t2.shape

    def add_error_vector(self, t1, t2):
        cur = np.concatenate((t1.ravel(), t2.ravel()))
        self.vals.append(cur)
        self.errs.append(cur - self.old)
        self.old = cur.copy()

    def extrapolate(self, t1, t2):
        if self.max_diis == 0:
            return t1, t2
        if len(self.errs) > self.max_diis:
            del self.vals[0]
            del self.errs[0]
        m = len(self.errs)
        B = -np.ones((m + 1, m + 1))
        B[-1, -1] = 0.0
        for p in range(m):
            for q in range(p, m):
                B[p, q] = B[q, p] = np.dot(self.errs[p], self.errs[q])
        B[:-1, :-1] /= np.abs(B[:-1, :-1]).max()
        rhs = np.zeros(m + 1)
        rhs[-1] = -1.0
        c = np.linalg.solve(B, rhs)
        new = np.zeros_like(self.old)
        for k in range(m):
            new += c[k] * self.vals[k + 1]
        self.old = new.copy()
        n1 = int(np.prod(self.shape1))
        return new[:n1].reshape(self.shape1), new[n1:].reshape(self.shape2)


def solve_cc(F, blocks, e_conv=1e-7, r_conv=1e-7, maxiter=100, max_diis=8, start_diis=1, log=None):
    """ccwfn.solve_cc, NumPy branch without local filtering (pycc/ccwfn.py:265-329).
    Returns (ecc, t1, t2, niter) or (None, t1, t2, maxiter) when not converged."""
    blk = blocks if isinstance(blocks, Blocks) else Blocks(blocks)
    no, nv = blk.no, blk.nv
    Dia, Dijab = denominators(F, no)
    t1 = np.zeros((no, nv))
    t2 = blk.oovv / Dijab
    ecc = cc_energy(F, blk, t1, t2)
    if log is not None:
        log.append((0, ecc, -ecc, None))
    diis = Diis(t1, t2, max_diis)
    for niter in range(1, maxiter + 1):
        ecc_last = ecc
        r1, r2 = residuals(F, blk, t1, t2)
        inc1, inc2 = r1 / Dia, r2 / Dijab
        t1 = t1 + inc1
        t2 = t2 + inc2
        rms = np.sqrt(np.sum(inc1 * inc1) + np.sum(inc2 * inc2))
        ecc = cc_energy(F, blk, t1, t2)
        ediff = ecc - ecc_last
        if log is not None:
            log.append((niter, ecc, ediff, rms))
        if abs(ediff) < e_conv and abs(rms) < r_conv:
            return ecc, t1, t2, niter
        diis.add_error_vector(t1, t2)
        if niter >= start_diis:
            t1, t2 = diis.extrapolate(t1, t2)
    return None, t1, t2, maxiter


def flops_ccsd_iteration(o, v):
    """Algorithmic flop count of one reference residuals() call."""
    return 2 * (o**2 * v**4 + 9 * o**3 * v**3 + 2 * o**4 * v**2 + 7 * o**2 * v**3 + 8 * o**3 * v**2
                + o * v**3 + 2 * o**4 * v + 12 * o**2 * v**2 + o**3 * v + 3 * o * v**2 + 3 * o**2 * v)
