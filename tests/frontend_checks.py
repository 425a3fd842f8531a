"""Shared checks of the psi4-free integral front end (pycc_b200.hamiltonian.from_factor / from_ao;
reference: pycc/hamiltonian.py:26-39) for the CPU (interpreter back end) and GPU (C ABI) tests."""
import contextlib
import io

import numpy as np


def check_factor_blocks(backend):
    from pycc_b200.hamiltonian import Hamiltonian
    from pycc_b200.synthetic import SyntheticSystem
    s = SyntheticSystem(5, 12, seed=2)
    H = Hamiltonian.from_factor(s.F, s.B, 5, backend)
    for k, ref in s.blocks.items():
        got = H.blocks[k].cpu().numpy()
        assert got.shape == ref.shape and np.abs(ref).max() > 1e-3
        assert np.abs(got - ref).max() < 1e-13, k
    # a rank's share: rows [4, 10) of <ab|ef> and of the b index of <mb|ef>
    H2 = Hamiltonian.from_factor(s.F, s.B, 5, backend, vslice=(4, 10), keep_full_ovvv=False)
    assert np.abs(H2.blocks["vvvv"].cpu().numpy() - s.blocks["vvvv"][4:10]).max() < 1e-13
    assert np.abs(H2.blocks["ovvv"].cpu().numpy() - s.blocks["ovvv"][:, 4:10]).max() < 1e-13
    # the provider views the reference slices out of ERI / L
    o, v = H.o, H.v
    full = s.full_eri()
    L = 2.0 * full - full.swapaxes(2, 3)
    n = 17
    oo, vv = slice(0, 5), slice(5, n)
    assert np.abs(H.ERI[o, v, v, o].cpu().numpy() - full[oo, vv, vv, oo]).max() < 1e-13
    assert np.abs(H.L[o, v, v, v].cpu().numpy() - L[oo, vv, vv, vv]).max() < 1e-13


def check_ao_transform(backend):
    """AO -> MO quarter transformations against numpy on a synthetic 8-fold-symmetric AO tensor, then a CCSD solve
    from the transformed Hamiltonian against the oracle on the same MO integrals."""
    import pycc_b200
    from pycc_b200.hamiltonian import mo_transform
    from oracle import ccsd_oracle as oc
    rng = np.random.default_rng(0)
    nao, nmo, no = 11, 10, 3
    G = rng.standard_normal((3 * nao, nao, nao))
    G = 0.06 * (G + G.transpose(0, 2, 1))
    ao = np.einsum('Qmn,Qls->mnls', G, G)                                  # chemists' (mu nu|la si)
    C, _ = np.linalg.qr(rng.standard_normal((nao, nao)))
    C = C[:, :nmo]
    eps = np.concatenate((np.sort(-2.0 + rng.random(no)), np.sort(0.5 + 2.0 * rng.random(nmo - no))))
    Fa = C @ np.diag(eps) @ C.T                                            # so that C^T Fa C = diag(eps)
    F, ERI = mo_transform(backend, Fa, ao, C)
    mo = np.einsum('mp,nq,lr,st,mnls->pqrt', C, C, C, C, ao, optimize=True)
    ref = mo.swapaxes(1, 2)                                                # hamiltonian.py:38
    assert np.abs(ref).max() > 1e-3
    assert np.abs(ERI.cpu().numpy() - ref).max() < 1e-13
    assert np.abs(F.cpu().numpy() - C.T @ Fa @ C).max() < 1e-13
    with contextlib.redirect_stdout(io.StringIO()):
        w = pycc_b200.ccwfn.from_ao(Fa, ao, C, no, _backend=backend)
        e = w.solve_cc(e_conv=1e-11, r_conv=1e-11)
    from pycc_b200.hamiltonian import blocks_from_full
    eo, t1, t2, niter = oc.solve_cc(C.T @ Fa @ C, {k: np.ascontiguousarray(x) for k, x in blocks_from_full(ref, no).items()},
                                    1e-11, 1e-11)
    assert abs(e - eo) < 1e-10 and len(w.iteration_log) == niter
    assert np.abs(w.t2.cpu().numpy() - t2).max() < 1e-9
