"""
cchbar.py: the CC similarity-transformed Hamiltonian, API-compatible with ``pycc/cchbar.py``
(class cchbar :13-375) but built by the B200 engine (``hbar_engine.HbarEngine``).

    hbar = pycc_b200.cchbar(cc)          # cc: a pycc_b200.ccwfn holding (converged) amplitudes

Attributes Hov Hvv Hoo Hoooo Hvvvv Hvovv Hooov Hovvo Hovov Hvvvo Hovoo are device tensors in the
reference's index order (Hvvvv is a permuted view of its transposed storage).
"""
import time

import torch


class cchbar(object):
    def __init__(self, ccwfn):
        time_init = time.time()
        if ccwfn.model not in ('CCD', 'CC2', 'CCSD', 'CCSD(T)', 'CC3'):
            raise Exception("HBAR in the B200 engine: models CCD, CC2, CCSD, CCSD(T) and CC3 (got %s)." % (ccwfn.model))
        from .hbar_engine import HbarEngine
        self.ccwfn = ccwfn
        self.contract = ccwfn.contract
        self.o, self.v = ccwfn.o, ccwfn.v
        self.no, self.nv = ccwfn.no, ccwfn.nv
        self.engine = HbarEngine(ccwfn.engine)
        self.engine.snapshot()
        self.engine.build()
        e = self.engine
        self.Hov, self.Hvv, self.Hoo = e.Hov, e.Hvv, e.Hoo
        self.Hoooo, self.Hvvvv, self.Hvovv, self.Hooov = e.Hoooo, e.Hvvvv, e.Hvovv, e.Hooov
        self.Hovvo, self.Hovov, self.Hvvvo, self.Hovoo = e.Hovvo, e.Hovov, e.Hvvvo, e.Hovoo
        ccwfn.backend.synchronize()
        print("\nHBAR constructed in %.3f seconds." % (time.time() - time_init))

    # Builders with the reference's signatures (pycc/cchbar.py:90-375).  The integral and HBAR
    # arguments are accepted for call-site compatibility: the engine uses the blocks of ccwfn.H and
    # rebuilds what a block depends on.  Each returns a new tensor.
    def _build(self, name, F, t1, t2, attr=None):
        w = self.ccwfn
        w._load(F, t1 if t1 is not None else w.engine.t1, t2 if t2 is not None else w.engine.t2)
        self.engine.build(only=name)
        x = getattr(self.engine, attr or name)
        return x.contiguous() if not x.is_contiguous() else x.clone()

    def build_Hov(self, o, v, F, L, t1):
        return self._build("Hov", F, t1, None)

    def build_Hvv(self, o, v, F, L, t1, t2):
        return self._build("Hvv", F, t1, t2)

    def build_Hoo(self, o, v, F, L, t1, t2):
        return self._build("Hoo", F, t1, t2)

    def build_Hoooo(self, o, v, ERI, t1, t2):
        return self._build("Hoooo", None, t1, t2)

    def build_Hvvvv(self, o, v, ERI, t1, t2):
        return self._build("Hvvvv", None, t1, t2)

    def build_Hvovv(self, o, v, ERI, t1):
        return self._build("Hvovv", None, t1, None)

    def build_Hooov(self, o, v, ERI, t1):
        return self._build("Hooov", None, t1, None)

    def build_Hovvo(self, o, v, ERI, L, t1, t2):
        return self._build("Hovvo", None, t1, t2)

    def build_Hovov(self, o, v, ERI, t1, t2):
        return self._build("Hovov", None, t1, t2)

    def build_Hvvvo(self, o, v, ERI, L, Hov, Hvvvv, t1, t2):
        return self._build("Hvvvo", None, t1, t2)

    def build_Hovoo(self, o, v, ERI, L, Hov, Hoooo, t1, t2):
        return self._build("Hovoo", None, t1, t2)
