"""
pycc_b200: B200-native engine for PyCC's RHF-CCSD amplitude iteration and (T) correction,
plus the HBAR / Lambda step that follows it.
Public names mirror ``pycc/__init__.py:9-18`` for the path it covers.
"""
from .ccwfn import ccwfn
from .cctriples import t_tjl, t_vikings, t3c_ijk, t3d_ijk
from .cchbar import cchbar
from .cclambda import cclambda
from .utils import helper_diis, cc_contract
from .hamiltonian import Hamiltonian

__all__ = ["ccwfn", "cchbar", "cclambda", "t_tjl", "t_vikings", "t3c_ijk", "t3d_ijk", "helper_diis", "cc_contract", "Hamiltonian"]
