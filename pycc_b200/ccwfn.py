"""
ccwfn.py: CC T-amplitude solver, API-compatible with ``pycc/ccwfn.py`` for the RHF-CCSD and
CCSD(T) models, executing on the B200 engine (``pycc_b200.engine`` over libpyccb200.so).

    wfn = ccwfn(scf_wfn, model='CCSD(T)')            # psi4 wavefunction (needs psi4), or
    wfn = ccwfn.from_arrays(F, ERI, no)              # full Dirac ERI (n,n,n,n), or
    wfn = ccwfn.from_arrays(F, blocks, no)           # dict of the six blocks (see hamiltonian.py)
    ecc = wfn.solve_cc(e_conv=1e-7, r_conv=1e-7, maxiter=100, max_diis=8, start_diis=1)

Same constructor kwargs, method names, argument orders, printed iteration lines and return
conventions as the reference (``pycc/ccwfn.py:62-329``); arrays are ``torch.Tensor`` on the
device (``.cpu().numpy()`` gives the reference's ndarrays).  There is no CPU path: without the
CUDA library or a B200 the constructor raises.
"""
import time

import numpy as np
import torch

from . import _lib
from .engine import CCSDEngine
from .hamiltonian import Hamiltonian, BLOCKS
from .utils import helper_diis, cc_contract


def _is_complex(x):
    if x is None:
        return False
    return x.is_complex() if isinstance(x, torch.Tensor) else np.iscomplexobj(x)


class ccwfn(object):
    """An RHF-CC wave function and energy object (see module docstring).

    Attributes mirror the reference (``pycc/ccwfn.py:23-52``): ref, eref, nfzc, no, nv, nmo, nact,
    H, o, v, Dia, Dijab, t1, t2, ecc, model, local, precision, device, contract.
    """

    valid_cc_models = ['CCD', 'CC2', 'CCSD', 'CCSD(T)', 'CC3']
    engine_models = ['CCD', 'CC2', 'CCSD', 'CCSD(T)', 'CC3']

    def __init__(self, scf_wfn, **kwargs):
        try:
            import psi4  # noqa: F401
        except ImportError:
            raise Exception("ccwfn(scf_wfn) needs psi4 for the MO integrals (pycc/hamiltonian.py:32-38); "
                            "use ccwfn.from_arrays(F, ERI_or_blocks, no) instead.")
        import psi4
        time_init = time.time()
        ref = scf_wfn
        nfzc = ref.frzcpi()[0]
        no = ref.doccpi()[0] - nfzc
        nmo = ref.nmo()
        nv = nmo - no - nfzc
        C = ref.Ca_subset("AO", "ACTIVE")
        npC = np.asarray(C)
        # Hamiltonian arrays exactly as pycc/hamiltonian.py:32-38 builds them
        F = npC.T @ np.asarray(ref.Fa()) @ npC
        mints = psi4.core.MintsHelper(ref.basisset())
        ERI = np.asarray(mints.mo_eri(C, C, C, C)).swapaxes(1, 2)
        self._setup(F, ERI, no, nv, kwargs, eref=ref.energy(), nfzc=nfzc, ref=ref, C=C, time_init=time_init)

    @classmethod
    def from_arrays(cls, F, ERI_or_blocks, no, eref=0.0, **kwargs):
        """psi4-free constructor: ``F`` (n,n) Fock matrix, ``ERI_or_blocks`` either the full Dirac
        array ``<pq|rs>`` (n,n,n,n) or a dict with the six blocks oooo, ooov, oovv, ovov, ovvv, vvvv
        (NumPy or torch, host or device), ``no`` active occupied orbitals."""
        self = object.__new__(cls)
        n = F.shape[0]
        self._setup(F, ERI_or_blocks, int(no), int(n - no), kwargs, eref=eref, nfzc=0, ref=None, C=None,
                    time_init=time.time())
        return self

    @classmethod
    def from_factor(cls, F, B, no, eref=0.0, **kwargs):
        """psi4-free constructor from a density-fitting / Cholesky factor B[Q,n,n] in the MO basis,
        (pq|rs) = sum_Q B[Q,p,q] B[Q,r,s]: every rank forms the six blocks - its own rows of <ab|ef> / <mb|ef> only -
        on the device (hamiltonian.blocks_from_factor); no n^4 array exists anywhere."""
        self = object.__new__(cls)
        n = F.shape[0]
        self._setup(F, ("factor", B), int(no), int(n - no), kwargs, eref=eref, nfzc=0, ref=None, C=None,
                    time_init=time.time())
        return self

    @classmethod
    def from_ao(cls, Fao, ao_eri, C, no, eref=0.0, **kwargs):
        """psi4-free constructor from AO-basis integrals: Fock matrix Fao[mu,nu], two-electron integrals in
        chemists' order ao_eri[mu,nu,la,si] = (mu nu|la si), MO coefficients C[mu,p] (pycc/hamiltonian.py:32-38
        without MintsHelper: hamiltonian.mo_transform, four pccb_contract launches)."""
        self = object.__new__(cls)
        nmo = C.shape[1]
        self._setup(None, ("ao", Fao, ao_eri, C), int(no), int(nmo - no), kwargs, eref=eref, nfzc=0, ref=None, C=C,
                    time_init=time.time())
        return self

    # -----------------------------------------------------------------------------------------
    def _setup(self, F, ERI, no, nv, kwargs, eref, nfzc, ref, C, time_init):
        model = kwargs.pop('model', 'CCSD').upper()
        if model not in self.valid_cc_models:
            raise Exception("%s is not an allowed CC model." % (model))
        if model not in self.engine_models:
            raise Exception("%s is not available in the B200 engine." % (model))
        self.model = model
        self.need_singles = ['CCSD', 'CCSD(T)', 'CC2', 'CC3']
        self.need_t1_transform = ['CC3']
        self.make_t3_density = kwargs.pop('make_t3_density', False)
        self.real_time = kwargs.pop('real_time', False)

        valid_local_models = [None, 'PNO', 'PAO', 'CPNO++', 'PNO++']
        local = kwargs.pop('local', None)
        if local not in valid_local_models:
            raise Exception("%s is not an allowed local-CC model." % (local))
        if local is not None:
            raise Exception("local-correlation models are outside the B200 engine's scope (local=None only).")
        self.local = local
        self.local_cutoff = kwargs.pop('local_cutoff', 1e-5)
        valid_local_MOs = ['PIPEK_MEZEY', 'BOYS']
        local_MOs = kwargs.pop('local_mos', 'PIPEK_MEZEY')
        if local_MOs not in valid_local_MOs:
            raise Exception("%s is not an allowed MO localization method." % (local_MOs))
        self.local_MOs = local_MOs
        it2_opt = kwargs.pop('it2_opt', True)
        if it2_opt not in [True, False]:
            raise Exception("%s is not an allowed initial t2 amplitudes." % (it2_opt))
        self.it2_opt = it2_opt
        filt = kwargs.pop('filter', False)
        if filt not in [True, False]:
            raise Exception("%s is not an allowed local filter." % (filt))
        self.filter = filt

        valid_precision = ['SP', 'DP']
        precision = kwargs.pop('precision', 'DP')
        if precision.upper() not in valid_precision:
            raise Exception('%s is not an allowed precision arithmetic.' % (precision))
        self.precision = precision.upper()
        # 'SP' = the engine's mixed-precision mode: FP32 storage and contractions (3xTF32 on the tensor
        # cores), FP64 accumulation of the energy, the increment norm and the DIIS algebra.  (The
        # reference's SP mode keeps everything in float32: pycc/ccwfn.py:189-196, utils.py:104-105.)
        valid_device = ['CPU', 'GPU']
        device = kwargs.pop('device', 'GPU')
        if device.upper() not in valid_device:
            raise Exception("%s is not an allowed device." % (device))
        # the engine has one execution path: the B200.  device='CPU' is accepted for call-site
        # compatibility with the reference (its default) but does not select a CPU implementation.
        self.device = 'GPU'

        # engine-only kwarg: group=None (one GPU), or 'world' / a torch.distributed process group
        # (one process per GPU; the virtual index and the (T) triplets are sharded over its ranks)
        from .parallel import Comm
        self.comm = Comm(kwargs.pop('group', None))
        backend = kwargs.pop('_backend', None)
        if backend is None:
            from .backend import CudaBackend
            backend = CudaBackend()
        self.backend = backend
        self.device0 = torch.device('cpu')
        self.device1 = backend.device

        self.ref = ref
        self.eref = eref
        self.nfzc = nfzc
        self.no, self.nv = no, nv
        self.nmo = no + nv + nfzc
        self.nact = no + nv
        print("NMO = %d; NACT = %d; NO = %d; NV = %d" % (self.nmo, self.nact, self.no, self.nv))
        self.o = slice(0, no)
        self.v = slice(no, no + nv)
        self.C = C

        dtype = torch.float32 if self.precision == 'SP' else torch.float64
        from .parallel import virtual_slice
        vslice = virtual_slice(nv, self.comm.rank, self.comm.world)
        # engine-only kwarg: keep every row of <mb|ef> on each rank (needed by HBAR / Lambda with group=...)
        full_ovvv = bool(kwargs.pop('full_ovvv', False))
        keep_ovvv = model in ('CCSD(T)', 'CC3') or self.comm.world == 1 or full_ovvv
        if isinstance(ERI, tuple) and ERI[0] == "factor":
            self.H = Hamiltonian.from_factor(F, ERI[1], no, backend, dtype, vslice, keep_ovvv)
        elif isinstance(ERI, tuple) and ERI[0] == "ao":
            self.H = Hamiltonian.from_ao(ERI[1], ERI[2], ERI[3], no, backend, dtype, vslice, keep_ovvv)
        elif isinstance(ERI, dict):
            self.H = Hamiltonian.from_blocks(F, ERI, no, backend.device, dtype, vslice, keep_ovvv)
        else:
            if tuple(ERI.shape) != (no + nv,) * 4:
                raise Exception("ERI must be the full (n,n,n,n) Dirac array or a dict of blocks %s" % (BLOCKS,))
            self.H = Hamiltonian.from_full(F, ERI, no, backend.device, dtype, vslice, keep_ovvv)
        self.engine = CCSDEngine(backend, self.H.F, self.H.blocks, no, nv, dtype, comm=self.comm,
                                 keep_full_ovvv=keep_ovvv, model=model)
        self.H.F = self.engine.F0
        self._triples = None

        # first-order amplitudes (pycc/ccwfn.py:170-175): t1 = 0, t2 = <ij|ab>/Dijab
        self.t1, self.t2 = self.engine.t1, self.engine.t2
        self.t1.zero_()
        self.t2.zero_()
        zero1 = backend.zeros((no, nv), dtype)
        backend.bind_stream()
        backend.update_amplitudes(self.t1, self.t2, zero1, self.H.blocks["oovv"], self.engine.eps, no, nv, False)
        self.contract = cc_contract(device=self.device, backend=backend)
        self.ecc = None
        print("CCWFN object initialized in %.3f seconds." % (time.time() - time_init))

    # denominators (pycc/ccwfn.py:164-167), materialised on demand for API compatibility; the
    # engine's kernels recompute them from the Fock diagonal instead of reading N2-sized arrays
    @property
    def Dia(self):
        eps = self.engine.eps
        return eps[:self.no].reshape(-1, 1) - eps[self.no:]

    @property
    def Dijab(self):
        eps = self.engine.eps
        eo, ev = eps[:self.no], eps[self.no:]
        return eo.reshape(-1, 1, 1, 1) + eo.reshape(-1, 1, 1) - ev.reshape(-1, 1) - ev

    # -----------------------------------------------------------------------------------------
    def solve_cc(self, e_conv=1e-7, r_conv=1e-7, maxiter=100, max_diis=8, start_diis=1):
        """Solve the CC amplitude equations (pycc/ccwfn.py:234-329).  Returns the correlation
        energy (plus E(T) for model 'CCSD(T)'), or None when not converged in ``maxiter``."""
        ccsd_tstart = time.time()
        eng = self.engine
        eng.set_fock(eng.F0)
        ecc = eng.energy()
        print("CC Iter %3d: CC Ecorr = %.15f  dE = % .5E  MP2" % (0, ecc, -ecc))
        diis = self.make_diis(max_diis)
        self.diis = diis
        self.iteration_log = []
        for niter in range(1, maxiter + 1):
            ecc_last = ecc
            ecc, rms = self._iteration(diis, niter, start_diis, extrapolate=False)
            ediff = ecc - ecc_last
            print("CC Iter %3d: CC Ecorr = %.15f  dE = % .5E  rms = % .5E" % (niter, ecc, ediff, rms))
            self.iteration_log.append((niter, ecc, ediff, rms))
            if (abs(ediff) < e_conv) and abs(rms) < r_conv:
                eng.sync_amplitudes()
                print("\nCCWFN converged in %.3f seconds.\n" % (time.time() - ccsd_tstart))
                print("E(REF)  = %20.15f" % self.eref)
                if self.model == 'CCSD(T)':
                    print("E(CCSD) = %20.15f" % ecc)
                    if self.make_t3_density is True:
                        et = self.t3_density()
                    else:
                        from .cctriples import t_tjl
                        et = t_tjl(self)
                    print("E(T)    = %20.15f" % et)
                    ecc = ecc + et
                else:
                    print("E(%s) = %20.15f" % (self.model, ecc))
                self.ecc = ecc
                print("E(TOT)  = %20.15f" % (ecc + self.eref))
                return ecc
            self._diis_step(diis, niter, start_diis)
        eng.sync_amplitudes()

    def t3_density(self):
        """(T) contributions to the Lambda equations and the densities (pycc/ccwfn.py:721-855): sets S1, S2, Dvv, Doo,
        Dov, Goovv, Gooov, Gvvvo (device tensors in the reference's index order) and returns E(T)."""
        if getattr(self, "_t3d", None) is None:
            from .t3density_engine import T3DensityEngine
            self._t3d = T3DensityEngine(self.engine)
        self.engine.sync_amplitudes()
        d = self._t3d
        et = d.run()
        self.S1, self.S2 = d.S1, d.S2
        self.Dvv, self.Doo, self.Dov = d.Dvv, d.Doo, d.Dov
        self.Goovv, self.Gooov, self.Gvvvo = d.Goovv, d.Gooov, d.Gvvvo
        return et

    def make_diis(self, max_diis=8):
        """helper_diis over this object's amplitudes; with several GPUs its history is sliced by the same
        (i,j) blocks of t2 that each rank updates."""
        eng = self.engine
        blocks = (eng._t2_store, self.nv * self.nv, eng.chunk_pairs) if self.comm.world > 1 else None
        return helper_diis(self.t1, self.t2, max_diis, self.precision, backend=self.backend, comm=self.comm,
                           blocks=blocks)

    def _iteration(self, diis, niter, start_diis, extrapolate=True):
        """One pass of the solve_cc loop body (pycc/ccwfn.py:274-296, 327-329): residuals, t += r/D
        with the pair symmetrisation of r2 (:621) folded into the update, the increment norm, the
        energy and (``extrapolate``) the DIIS step.  Returns (ecc, rms).  This is the unit bench.py
        times as one step."""
        eng = self.engine
        ev = getattr(self, "phase_events", None)     # bench.py: list receiving 4 CUDA events per step
        mark = (lambda: None) if ev is None else (lambda: (ev.append(torch.cuda.Event(enable_timing=True)), ev[-1].record()))
        mark()
        eng.sync_amplitudes()
        eng.residuals(symmetrize=False)
        mark()
        rms2, ecc = eng.update_energy()          # one fused pass; split over the ranks by (i,j) blocks
        rms = np.sqrt(rms2)
        mark()
        if extrapolate:
            self._diis_step(diis, niter, start_diis)
        mark()
        return ecc, rms

    def _diis_step(self, diis, niter, start_diis):
        """DIIS push + extrapolation.  After a split update each rank's history slice is exactly the blocks it
        updated, and the allgather of the extrapolated chunks completes t2; without an extrapolation the
        updated chunks themselves are gathered."""
        eng = self.engine
        if diis.blocks is None:
            eng.sync_amplitudes()        # a flat-chunk history packs slices of the COMPLETE vector
        diis.add_error_vector(self.t1, self.t2)
        if niter >= start_diis and diis.max_diis > 0:
            self.t1, self.t2 = diis.extrapolate(self.t1, self.t2)
            if diis.blocks is not None:
                eng.t2_stale = False     # the allgather of the extrapolated chunks completed t2
        eng.sync_amplitudes()

    def _load(self, F, t1, t2):
        eng = self.engine
        if _is_complex(F) or _is_complex(t1) or _is_complex(t2):
            raise Exception("complex amplitudes / Fock matrices go through ccwfn.residuals() only "
                            "(two real planes); the builders of the B200 engine are real.")
        if F is None:
            F = eng.F0                   # the Hamiltonian's own Fock matrix
        if F is not eng.F:
            eng.set_fock(F)
        if t1 is not eng.t1:
            eng.t1.copy_(torch.as_tensor(t1).to(eng.t1.device, eng.dtype))
        if t2 is not eng.t2:
            eng.t2.copy_(torch.as_tensor(t2).to(eng.t2.device, eng.dtype))

    class _Pure(object):
        """Context manager that makes a public builder pure, like the reference's: the object's own
        amplitudes and Fock matrix are put back when the call used other ones (the engine's
        intermediates keep the values of the call)."""

        def __init__(self, w, t1, t2):
            self.eng = w.engine
            e = self.eng
            self.k1 = None if (t1 is None or t1 is e.t1) else e.t1.clone()
            self.k2 = None if (t2 is None or t2 is e.t2) else e.t2.clone()

        def __enter__(self):
            return self

        def __exit__(self, *exc):
            e = self.eng
            if self.k1 is not None:
                e.t1.copy_(self.k1)
            if self.k2 is not None:
                e.t2.copy_(self.k2)
            e.set_fock(e.F0)
            return False

    def residuals(self, F, t1, t2, real_time=False):
        """r1, r2 = <mu|HBAR|0> for the given Fock operator and amplitudes (pycc/ccwfn.py:331-406).
        Pure, like the reference's: the object's amplitudes and ``H.F`` are unchanged afterwards
        (the engine's intermediates hold the values of this call).  Returns new tensors; complex
        inputs give complex results."""
        if _is_complex(F) or _is_complex(t1) or _is_complex(t2):
            # complex amplitudes (real-time propagation, pycc/rt/rtcc.py:125): two real planes
            return self._complex().residuals(F, t1, t2, real_time=bool(real_time))
        eng = self.engine
        # CC3 with real_time=True: the reference adds the t3_pert_ijk term built from V = F - H.F
        # (pycc/ccwfn.py:378-386); complex CC3 amplitudes are outside the engine (_complex() raises)
        pert = bool(real_time) and self.model == 'CC3'
        if pert and eng.cc3 is None:
            from .cc3_engine import CC3Engine
            eng.cc3 = CC3Engine(eng)
        with self._Pure(self, t1, t2):   # pure: the object's amplitudes and H.F are put back afterwards
            self._load(F, t1, t2)
            if pert:
                eng.cc3.pert = True
            try:
                r1, r2 = eng.residuals(symmetrize=True)
            finally:
                if pert:
                    eng.cc3.pert = False
            r1, r2 = r1.clone(), r2.clone()
        return r1, r2

    # ---- host <-> device transfer of the amplitudes, sharded over the ranks -------------------------
    # (engine-only API)  With several GPUs the amplitude vector (t1, t2) lives in host memory ONCE: rank g
    # holds elements [lo, hi) of the flattened concatenation.  Loading copies each shard over its own
    # PCIe link and rebuilds the replicated (t1, t2) with one allgather over NVLink; storing writes back
    # only the rank's shard.  On one GPU the shard is the whole vector.
    def amplitude_shard(self):
        """(lo, hi): this rank's range of the flattened (t1, t2) vector."""
        from .parallel import flat_chunk
        lo, hi, _ = flat_chunk(self.t1.numel() + self.t2.numel(), self.comm.rank, self.comm.world)
        return lo, hi

    def _shard_buffers(self):
        if getattr(self, "_shard", None) is None:
            from .parallel import flat_chunk
            n = self.t1.numel() + self.t2.numel()
            lo, hi, chunk = flat_chunk(n, self.comm.rank, self.comm.world)
            dev = self.backend.zeros((chunk,), self.engine.dtype)
            full = self.backend.zeros((chunk * self.comm.world,), self.engine.dtype) if self.comm.world > 1 else dev
            self._shard = (lo, hi, dev, full)
        return self._shard

    def load_amplitude_shard(self, host_shard):
        """(t1, t2) on every rank from the ranks' host shards (``host_shard``: this rank's elements
        [lo, hi), ideally pinned).  Asynchronous on the current stream."""
        lo, hi, dev, full = self._shard_buffers()
        n1 = self.t1.numel()
        dev[:hi - lo].copy_(host_shard, non_blocking=True)
        self.comm.allgather_into(full, dev)
        self.t1.copy_(full[:n1].reshape(self.t1.shape))
        self.t2.copy_(full[n1:n1 + self.t2.numel()].reshape(self.t2.shape))

    def store_amplitude_shard(self, host_shard):
        """This rank's elements [lo, hi) of the current (t1, t2) into ``host_shard``.  Asynchronous."""
        lo, hi, dev, _ = self._shard_buffers()
        n1 = self.t1.numel()
        if lo < n1:
            e = min(hi, n1)
            dev[:e - lo].copy_(self.t1.reshape(-1)[lo:e])
        if hi > n1:
            b = max(lo, n1)
            dev[b - lo:hi - lo].copy_(self.t2.reshape(-1)[b - n1:hi - n1])
        host_shard.copy_(dev[:hi - lo], non_blocking=True)

    def _complex(self):
        if getattr(self, "_cplx", None) is None:
            from .complex_engine import ComplexResiduals
            self._cplx = ComplexResiduals(self.engine)
        return self._cplx

    def build_tau(self, t1, t2, fact1=1.0, fact2=1.0):
        out = self.backend.empty(tuple(t2.shape), self.engine.dtype)
        t1d = torch.as_tensor(t1).to(out.device, out.dtype).contiguous()
        t2d = torch.as_tensor(t2).to(out.device, out.dtype).contiguous()
        self.backend.bind_stream()
        self.backend.pair_combine(out, t2d, fact1, 0.0, t1d, t1d, fact2, self.no, self.nv)
        return out

    # Builders keep the reference's signatures (pycc/ccwfn.py:413-537).  The integral arguments
    # (ERI, L) are accepted for call-site compatibility; the engine uses the blocks of self.H.
    def _builder(self, name, F, t1, t2):
        """False when the model has no such intermediate (the reference's builders return None then:
        CCD has no Fme / Zmbij, CC2 no Wmbej / Wmbje; pycc/ccwfn.py:453-454, 494-496, 516-518, 532-533)."""
        if name not in self.engine.plans or name not in self.engine.builder_order() or \
                (self.model == 'CCD' and name in ('Fme', 'Zmbij')):
            return False
        with self._Pure(self, t1, t2):
            self._load(F, t1, t2)
            self.engine.run_builder(name)
        return True

    def build_Fae(self, o, v, F, L, t1, t2):
        if not self._builder("Fae", F, t1, t2):
            return None
        return self.engine.Fae.clone()

    def build_Fmi(self, o, v, F, L, t1, t2):
        if not self._builder("Fmi", F, t1, t2):
            return None
        return self.engine.Fmi.clone()

    def build_Fme(self, o, v, F, L, t1):
        if not self._builder("Fme", F, t1, self.engine.t2):
            return None
        return self.engine.Fme.clone()

    def build_Wmnij(self, o, v, ERI, t1, t2):
        if not self._builder("Wmnij", None, t1, t2):
            return None
        return self.engine.Wmnij.clone()

    def build_Wmbej(self, o, v, ERI, L, t1, t2):
        if not self._builder("Wmbej", None, t1, t2):
            return None
        return self.engine.Wmbej_view().contiguous()

    def build_Wmbje(self, o, v, ERI, t1, t2):
        if not self._builder("Wmbje", None, t1, t2):
            return None
        return self.engine.Wmbje_view().contiguous()

    def build_Zmbij(self, o, v, ERI, t1, t2):
        if not self._builder("Zmbij", None, t1, t2):
            return None
        return self.engine.Zmbij_view().contiguous()

    def r_T1(self, o, v, F, ERI, L, t1, t2, Fae, Fme, Fmi):
        eng = self.engine
        with self._Pure(self, t1, t2):
            self._load(F, t1, t2)
            eng.be.bind_stream()
            eng.prepare()
            eng.Fae.copy_(torch.as_tensor(Fae).to(eng.Fae.device, eng.dtype))
            eng.Fme.copy_(torch.as_tensor(Fme).to(eng.Fme.device, eng.dtype))
            eng.Fmi.copy_(torch.as_tensor(Fmi).to(eng.Fmi.device, eng.dtype))
            eng._run("r_T1")
            out = eng.r1.clone()
        return out

    def r_T2(self, o, v, F, ERI, L, t1, t2, Fae, Fme, Fmi, Wmnij, Wmbej, Wmbje, Zmbij):
        eng = self.engine
        with self._Pure(self, t1, t2):
            self._load(F, t1, t2)
            eng.be.bind_stream()
            eng.prepare()
            dev = lambda x: torch.as_tensor(x).to(eng.t2.device, eng.dtype)
            eng.Fae.copy_(dev(Fae))
            eng.Fme.copy_(dev(Fme))
            eng.Fmi.copy_(dev(Fmi))
            eng.Wmnij.copy_(dev(Wmnij))
            eng.Wmbej_view().copy_(dev(Wmbej))
            eng.Wmbje_view().copy_(dev(Wmbje))
            eng.Zmbij_view().copy_(dev(Zmbij))
            eng._run("r_T2")
            eng.be.pair_symmetrize(eng.r2, self.no, self.nv)
            out = eng.r2.clone()
        return out

    def cc_energy(self, o, v, F, L, t1, t2):
        with self._Pure(self, t1, t2):
            self._load(F, t1, t2)
            e = self.engine.energy()
        return e
