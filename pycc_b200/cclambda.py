"""
cclambda.py: Lambda-amplitude solver, API-compatible with ``pycc/cclambda.py`` (class cclambda
:16-581) for the models CCD, CC2, CCSD, CCSD(T) and CC3, executing on the B200 engine
(``hbar_engine.LambdaEngine``).

    hbar = pycc_b200.cchbar(cc)
    lam  = pycc_b200.cclambda(cc, hbar)
    lecc = lam.solve_lambda(e_conv, r_conv)          # same signature, prints and return value
"""
import time

import numpy as np
import torch

from .utils import helper_diis


class cclambda(object):
    def __init__(self, ccwfn, hbar):
        from .hbar_engine import LambdaEngine
        self.ccwfn = ccwfn
        self.hbar = hbar
        self.contract = ccwfn.contract
        self.engine = LambdaEngine(hbar.engine)
        if ccwfn.model == 'CCSD(T)':
            # the reference reads ccwfn.S1 / ccwfn.S2 (cclambda.py:474-475, 519-520): they must exist
            # (solve_cc(..., make_t3_density=True) or ccwfn.t3_density())
            if not hasattr(ccwfn, "S1") or not hasattr(ccwfn, "S2"):
                raise Exception("CCSD(T) Lambda needs the (T) source terms: construct ccwfn with make_t3_density=True "
                                "or call ccwfn.t3_density() first.")
            self.engine.t3_source = (ccwfn.S1, ccwfn.S2)
        if ccwfn.model == 'CC3':
            # triples contributions Y1 / Y2 of the loops in solve_lambda (cclambda.py:118-139, 160-235)
            from .cc3_lambda_engine import CC3LambdaEngine
            self.engine.cc3 = CC3LambdaEngine(self.engine)
        self.engine.guess()                       # l1 = 2 t1, l2 = 2 (2 t2 - t2^T)  (cclambda.py:57-58)
        self.l1, self.l2 = self.engine.l1, self.engine.l2

    def solve_lambda(self, e_conv=1e-7, r_conv=1e-7, maxiter=100, max_diis=8, start_diis=1):
        """Solve the Lambda equations (pycc/cclambda.py:60-274).  Returns the pseudoenergy, or None
        when not converged in ``maxiter`` iterations."""
        lambda_tstart = time.time()
        self._ensure_hbar()
        eng = self.engine
        lecc = eng.pseudoenergy()
        print("\nLCC Iter %3d: LCC PseudoE = %.15f  dE = % .5E" % (0, lecc, -lecc))
        diis = helper_diis(self.l1, self.l2, max_diis, self.ccwfn.precision, backend=self.ccwfn.backend,
                           comm=self.ccwfn.comm)
        self.iteration_log = []
        for niter in range(1, maxiter + 1):
            lecc_last = lecc
            lecc, rms = self._iteration()
            ediff = lecc - lecc_last
            print("LCC Iter %3d: LCC PseudoE = %.15f  dE = % .5E  rms = % .5E" % (niter, lecc, ediff, rms))
            self.iteration_log.append((niter, lecc, ediff, rms))
            if (abs(ediff) < e_conv) and abs(rms) < r_conv:
                print("\nLambda-CC has converged in %.3f seconds.\n" % (time.time() - lambda_tstart))
                return lecc
            diis.add_error_vector(self.l1, self.l2)
            if niter >= start_diis:
                self.l1, self.l2 = diis.extrapolate(self.l1, self.l2)

    def _iteration(self):
        """Residuals, l += r/D (the r2 + r2^T of :542 folded into the update), increment norm and
        pseudoenergy (cclambda.py:141-151, 244-254).  Returns (lecc, rms)."""
        self._ensure_hbar()
        eng = self.engine
        eng.residuals(symmetrize=False)
        rms = np.sqrt(eng.update(symmetrize_on_the_fly=True))
        return eng.pseudoenergy(), rms

    def _load(self, l1, l2):
        eng = self.engine
        if l1 is not eng.l1:
            eng.l1.copy_(torch.as_tensor(l1).to(eng.l1.device, eng.l1.dtype))
        if l2 is not eng.l2:
            eng.l2.copy_(torch.as_tensor(l2).to(eng.l2.device, eng.l2.dtype))

    def residuals(self, F, t1, t2, l1, l2):
        """L1 and L2 residuals for the given Fock operator and amplitudes; HBAR is rebuilt from
        (F, t1, t2) first (pycc/cclambda.py:281-323).  Returns new tensors."""
        from .ccwfn import _is_complex
        if any(_is_complex(x) for x in (F, t1, t2, l1, l2)):
            # rtcc.f (pycc/rt/rtcc.py:132): complex amplitudes, HBAR rebuilt from them
            return self.ccwfn._complex().lambda_residuals(F, t1, t2, l1, l2,
                                                          real_time=bool(getattr(self.ccwfn, "real_time", False)))
        w, eng, le = self.ccwfn, self.ccwfn.engine, self.engine
        cc3 = getattr(le, "cc3", None)
        if cc3 is not None:
            # the CC3 branch (cclambda.py:323-440) repeats the solve_lambda loops for THESE (F, t1, t2): the T-only
            # intermediates are rebuilt from them, with the t3_pert_ijk term when the wave function is a real-time one
            cc3.ready = False
            cc3.pert = bool(getattr(w, "real_time", False))
        own_t = t1 is eng.t1 and t2 is eng.t2
        own_l = l1 is le.l1 and l2 is le.l2
        self.hbar.engine.snapshot()
        if not own_t:                    # the objects' amplitudes are put back afterwards (the HBAR
            kt1, kt2 = eng.t1.clone(), eng.t2.clone()     # blocks keep the values of this call)
        if not own_l:
            kl1, kl2 = le.l1.clone(), le.l2.clone()
        w._load(F, t1, t2)
        self.hbar.engine.build()
        self._load(l1, l2)
        r1, r2 = le.residuals(symmetrize=True)
        r1, r2 = r1.clone(), r2.clone()
        if cc3 is not None:
            cc3.ready, cc3.pert = False, False     # the next solve_lambda rebuilds them from the object's own amplitudes
        if not own_t:
            eng.t1.copy_(kt1)
            eng.t2.copy_(kt2)
        if not own_l:
            le.l1.copy_(kl1)
            le.l2.copy_(kl2)
        eng.set_fock(eng.F0)
        # the HBAR blocks now hold the values of THIS call; the reference's hbar.H* are never touched
        # by residuals(): rebuild them from the object's own (F0, t1, t2) before the next use
        foreign_F = F is not None and F is not eng.F0 and F is not eng.F
        if not own_t or foreign_F:
            self._hbar_dirty = True
        return r1, r2

    def _ensure_hbar(self):
        """HBAR of the object's own amplitudes (rebuilt after a residuals() call on other inputs)."""
        if getattr(self, "_hbar_dirty", False):
            eng = self.ccwfn.engine
            eng.set_fock(eng.F0)
            self.hbar.engine.snapshot()
            self.hbar.engine.build()
            self._hbar_dirty = False

    def build_Goo(self, t2, l2):
        self.ccwfn._load(None, self.ccwfn.engine.t1, t2)
        self._load(self.engine.l1, l2)
        self.engine.run("G")
        return self.engine.Goo.clone()

    def build_Gvv(self, t2, l2):
        self.ccwfn._load(None, self.ccwfn.engine.t1, t2)
        self._load(self.engine.l1, l2)
        self.engine.run("G")
        return self.engine.Gvv.clone()

    def _set_hbar(self, **blocks):
        h = self.hbar.engine
        for k, x in blocks.items():
            if x is None:
                continue
            dst = getattr(h, k)
            if x is not dst:
                dst.copy_(torch.as_tensor(x).to(dst.device, dst.dtype))

    def r_L1(self, o, v, l1, l2, Hov, Hvv, Hoo, Hovvo, Hovov, Hvvvo, Hovoo, Hvovv, Hooov, Gvv, Goo):
        self._load(l1, l2)
        self._set_hbar(Hov=Hov, Hvv=Hvv, Hoo=Hoo, Hovvo=Hovvo, Hovov=Hovov, Hvvvo=Hvvvo, Hovoo=Hovoo,
                       Hvovv=Hvovv, Hooov=Hooov)
        e = self.engine
        e.Gvv.copy_(torch.as_tensor(Gvv).to(e.Gvv.device, e.Gvv.dtype))
        e.Goo.copy_(torch.as_tensor(Goo).to(e.Goo.device, e.Goo.dtype))
        e.run("r_L1")
        return e.r1.clone()

    def r_L2(self, o, v, l1, l2, L, Hov, Hvv, Hoo, Hoooo, Hvvvv, Hovvo, Hovov, Hvvvo, Hovoo, Hvovv, Hooov, Gvv, Goo):
        self._load(l1, l2)
        self._set_hbar(Hov=Hov, Hvv=Hvv, Hoo=Hoo, Hoooo=Hoooo, Hvvvv=Hvvvv, Hovvo=Hovvo, Hovov=Hovov,
                       Hvvvo=Hvvvo, Hovoo=Hovoo, Hvovv=Hvovv, Hooov=Hooov)
        e = self.engine
        e.Gvv.copy_(torch.as_tensor(Gvv).to(e.Gvv.device, e.Gvv.dtype))
        e.Goo.copy_(torch.as_tensor(Goo).to(e.Goo.device, e.Goo.dtype))
        e.run("r_L2")
        e.be.pair_symmetrize(e.r2, e.eng.no, e.eng.nv)
        return e.r2.clone()

    def pseudoenergy(self, o, v, ERI, l2):
        self._load(self.engine.l1, l2)
        return self.engine.pseudoenergy()
