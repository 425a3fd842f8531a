"""
Import the UNMODIFIED reference (jattakumi/pycc) from /root/reference for golden-vector generation.

Only usable in the build container (the GPU box has no /root/reference): nothing under
``tests/`` that runs with ``-m gpu``, ``bench.py`` or ``smoke()`` imports this module.

psi4 and opt_einsum are not installed offline, so two stub modules are injected
(recipe: SURVEY.md section 8(c)):
  * ``psi4``        -> empty module (only used by ``ccwfn.__init__``, which is bypassed)
  * ``opt_einsum``  -> ``contract = numpy.einsum(..., optimize=True)`` (same pairwise-tensordot
                       strategy; opt_einsum itself contains no arithmetic)
The ``ccwfn`` object is created with ``object.__new__`` and given exactly the attributes
``pycc/ccwfn.py:76-200`` would set from a psi4 wavefunction.
"""
import os
import sys
import types

import numpy as np

REFERENCE_ROOT = os.environ.get("PYCC_REFERENCE_ROOT", "/root/reference")


def reference_available():
    return os.path.isdir(os.path.join(REFERENCE_ROOT, "pycc"))


def _install_stubs():
    if "psi4" not in sys.modules:
        psi4 = types.ModuleType("psi4")
        psi4.core = types.ModuleType("psi4.core")
        sys.modules["psi4"] = psi4
        sys.modules["psi4.core"] = psi4.core
    if "opt_einsum" not in sys.modules:
        oe = types.ModuleType("opt_einsum")

        def contract(subscripts, *operands, **kwargs):
            return np.einsum(subscripts, *operands, optimize=True)

        oe.contract = contract
        sys.modules["opt_einsum"] = oe


def import_reference():
    """Returns the reference's (ccwfn class, cctriples module, utils module)."""
    if not reference_available():
        raise RuntimeError("reference tree not present at %s" % REFERENCE_ROOT)
    _install_stubs()
    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)
    from pycc.ccwfn import ccwfn          # noqa: E402
    import pycc.cctriples as cctriples    # noqa: E402
    import pycc.utils as utils            # noqa: E402
    return ccwfn, cctriples, utils


class _H(object):
    pass


def reference_ccwfn_from_arrays(F, ERI, no, model="CCSD", precision="DP"):
    """A reference ``ccwfn`` whose Hamiltonian is (F, ERI) instead of a psi4 wavefunction."""
    ccwfn, _, utils = import_reference()
    n = F.shape[0]
    nv = n - no
    w = object.__new__(ccwfn)
    w.model = model.upper()
    w.need_singles = ['CCSD', 'CCSD(T)', 'CC2', 'CC3']
    w.need_t1_transform = ['CC3']
    w.make_t3_density = False
    w.real_time = False
    w.local = None
    w.filter = False
    w.ref = None
    w.eref = 0.0
    w.nfzc = 0
    w.no, w.nv, w.nmo, w.nact = no, nv, n, n
    w.o = slice(0, no)
    w.v = slice(no, n)
    H = _H()
    H.F = np.array(F, dtype=np.float64)
    H.ERI = np.array(ERI, dtype=np.float64)
    H.L = 2.0 * H.ERI - H.ERI.swapaxes(2, 3)
    w.H = H
    o, v = w.o, w.v
    eps_occ = np.diag(H.F)[o]
    eps_vir = np.diag(H.F)[v]
    w.Dia = eps_occ.reshape(-1, 1) - eps_vir
    w.Dijab = eps_occ.reshape(-1, 1, 1, 1) + eps_occ.reshape(-1, 1, 1) - eps_vir.reshape(-1, 1) - eps_vir
    w.t1 = np.zeros((no, nv))
    w.t2 = H.ERI[o, o, v, v] / w.Dijab
    w.precision = precision.upper()
    w.device = 'CPU'
    if w.precision == 'SP':
        H.F = np.float32(H.F)
        w.t1 = np.float32(w.t1)
        w.t2 = np.float32(w.t2)
        w.Dia = np.float32(w.Dia)
        w.Dijab = np.float32(w.Dijab)
        H.ERI = np.float32(H.ERI)
        H.L = np.float32(H.L)
    w.contract = utils.cc_contract(device='CPU')
    return w
