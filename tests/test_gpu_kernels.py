"""GPU parity tests of the C-ABI kernels (run with -m gpu on a B200).  Every call goes through
libpyccb200.so; the comparisons are against the reference-generated golden vectors, the CPU
oracle, and (for the raw contraction kernel) torch.einsum in FP64 on the same device."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def be():
    from pycc_b200.backend import CudaBackend
    return CudaBackend()


def _plan_run(be, stmt, tensors):
    from pycc_b200.dsl import plan
    for L in plan(stmt, tensors):
        be.contract(L.desc)


CASES = [
    # (statement, shapes) - exercises K-major / MN-major tiles, 8- and 16-byte cp.async, compound
    # indices, multi-segment K, batch promotion, split-K, edges
    ("C[ijab] = 0.5*A[ijef]*B[abef]", dict(A=(4, 4, 10, 10), B=(10, 10, 10, 10), C=(4, 4, 10, 10))),
    ("C[ijab] = 0.5*A[ijef]*B[abef]", dict(A=(3, 3, 7, 7), B=(7, 7, 7, 7), C=(3, 3, 7, 7))),
    ("C[ijab] = A[imae]*B[jbme] + A2[imae]*B2[jbme]",
     dict(A=(5, 5, 12, 12), A2=(5, 5, 12, 12), B=(5, 12, 5, 12), B2=(5, 12, 5, 12), C=(5, 5, 12, 12))),
    ("C[ijab] = A[mjae]*B[ibme]", dict(A=(6, 6, 9, 9), B=(6, 9, 6, 9), C=(6, 6, 9, 9))),
    ("C[ae] = A[mnaf]*B[mnef]", dict(A=(8, 8, 40, 40), B=(8, 8, 40, 40), C=(40, 40))),          # split-K
    ("C[mi] = A[mnef]*B[inef]", dict(A=(8, 8, 30, 30), B=(8, 8, 30, 30), C=(8, 8))),              # split-K
    ("C[jbme] = A[jnfb]*B[mnfe]", dict(A=(4, 4, 14, 14), B=(4, 4, 14, 14), C=(4, 14, 4, 14))),    # MN-major both
    ("C[jbme] = A[jf]*B[mbef]", dict(A=(4, 14), B=(4, 14, 14, 14), C=(4, 14, 4, 14))),            # batch promotion
    ("C[ijab] = A[ie]*B[jabe]", dict(A=(4, 13), B=(4, 13, 13, 13), C=(4, 4, 13, 13))),
    ("C[mn] = A[mk]*B[nk]", dict(A=(300, 500), B=(260, 500), C=(300, 260))),
    ("C[mn] = A[km]*B[kn]", dict(A=(333, 129), B=(333, 257), C=(129, 257))),                      # odd, MN-major, vec 1
    ("C[mn] = A[km]*B[nk]", dict(A=(64, 130), B=(258, 64), C=(130, 258))),
    ("C[abc] = 0.5*A[eab]*B[ce] + 2.0*A2[eba]*B2[ce] - 3.0*A3[mc]*B3[mab]",
     dict(A=(20, 22, 22), B=(22, 20), A2=(20, 22, 22), B2=(22, 20), A3=(6, 22), B3=(6, 22, 22), C=(22, 22, 22))),
    ("C[ijab] = 0.5*A[ijef]*B[abef]", dict(A=(9, 9, 24, 24), B=(24, 24, 24, 24), C=(9, 9, 24, 24))),   # MI=6 tiles
    ("C[ijab] = A[ijef]*B[abef]", dict(A=(20, 20, 16, 16), B=(16, 16, 16, 16), C=(20, 20, 16, 16))),    # MI=5 tiles
    ("C[abc] = A[eab]*B[ce] + A2[eba]*B2[ce] - A3[mc]*B3[mab]",
     dict(A=(20, 20, 20), B=(20, 20), A2=(20, 20, 20), B2=(20, 20), A3=(6, 20), B3=(6, 20, 20), C=(20, 20, 20))),
    # short K, >= 2 tiles per SM: the kernel with epilogue warps (MI = 5, 6, 4; outputs contiguous along
    # M, along N and compound; the (T) W build at v = 150 in its three output orders)
    ("C[abc] = A[eab]*B[ce] + A2[eba]*B2[ce] - A3[mc]*B3[mab]",
     dict(A=(150, 150, 150), B=(150, 150), A2=(150, 150, 150), B2=(150, 150), A3=(7, 150), B3=(7, 150, 150),
          C=(150, 150, 150))),
    ("C[abc] = A[eac]*B[be] - A3[mb]*B3[mac]",
     dict(A=(150, 150, 150), B=(150, 150), A3=(7, 150), B3=(7, 150, 150), C=(150, 150, 150))),
    ("C[abc] = A[ebc]*B[ae] - A3[ma]*B3[mbc]",
     dict(A=(150, 150, 150), B=(150, 150), A3=(7, 150), B3=(7, 150, 150), C=(150, 150, 150))),
    ("C[mn] = A[mk]*B[nk]", dict(A=(90, 200), B=(38011, 200), C=(90, 38011))),
    ("C[nm] = A[km]*B[kn]", dict(A=(137, 60), B=(137, 38002), C=(38002, 60))),
    ("C[nm] = A[mk]*B[nk]", dict(A=(157, 300), B=(24000, 300), C=(24000, 157))),
    # plain K-contiguous matrices with M = 256 / 1024: the 128-row TMA-staged kernel, contract_f64_tma_kernel<8>
    ("C[mn] = A[mk]*B[nk]", dict(A=(256, 520), B=(390, 520), C=(256, 390))),
    ("C[mn] = A[mk]*B[nk] + A2[ml]*B2[nl]", dict(A=(1024, 200), B=(2100, 200), A2=(1024, 72), B2=(2100, 72), C=(1024, 2100))),
    # M = 900 (= o^2 at o = 30): the 48-row tile, TMA-staged kernel <3> and the cp.async kernel <.,3> (row-contiguous operands)
    ("C[mn] = A[mk]*B[nk]", dict(A=(900, 300), B=(520, 300), C=(900, 520))),
    ("C[mn] = A[km]*B[kn]", dict(A=(2000, 900), B=(2000, 260), C=(900, 260))),
    # batched plain matrices (batch index on A, B and C; on A and C only): the TMA kernel with 3-D tensor maps,
    # ragged K (70 = 4 k-tiles + 6: partial last k-tile, zero k-steps skipped), two segments
    ("C[kmn] = A[kmp]*B[knp]", dict(A=(5, 300, 70), B=(5, 260, 70), C=(5, 300, 260))),
    ("C[kmn] = A[kmp]*B[knp] - 0.5*A2[kmq]*B2[nq]",
     dict(A=(3, 320, 150), B=(3, 130, 150), A2=(3, 320, 30), B2=(130, 30), C=(3, 320, 130))),
    # short K, M = v^2 with v = 250 rows (128-row tiles by padding): rerouted to the kernel with epilogue warps
    ("C[abc] = A[eab]*B[ce] - A3[mc]*B3[mab]",
     dict(A=(250, 250, 250), B=(250, 250), A3=(9, 250), B3=(9, 250, 250), C=(250, 250, 250))),
    # a few waves of long tiles (6 x 52 = 312 tiles = 2.1 waves, K = 4200 with a ragged tail): tail-aware split-K
    # (one rank's ladder out of eight) on the TMA-staged kernel, plain and row-contiguous operands
    ("C[mn] = A[mk]*B[nk]", dict(A=(480, 4200), B=(6650, 4200), C=(480, 6650))),
    ("C[mn] = A[km]*B[kn]", dict(A=(4200, 480), B=(4200, 6650), C=(480, 6650))),
]


@pytest.mark.parametrize("naive", [False, True])
@pytest.mark.parametrize("case", range(len(CASES)))
def test_contract_vs_einsum(be, case, naive):
    from pycc_b200.dsl import parse
    stmt, shapes = CASES[case]
    g = torch.Generator(device="cpu").manual_seed(case)
    T = {k: torch.randn(s, generator=g, dtype=torch.float64).to(be.device) for k, s in shapes.items()}
    out_name, out_idx, acc, terms = parse(stmt)
    ref = torch.zeros_like(T[out_name])
    for coef, fac in terms:
        spec = ",".join(f[1] for f in fac) + "->" + out_idx
        ref += coef * torch.einsum(spec, *[T[f[0]] for f in fac])
    be.use_naive_contract(naive)
    try:
        T[out_name].fill_(float("nan"))
        _plan_run(be, stmt, T)
        # accumulate form: run "+=" on top of a known C
        stmt2 = stmt.replace(" = ", " += ", 1)
        C0 = torch.randn(shapes[out_name], generator=g, dtype=torch.float64).to(be.device)
        got1 = T[out_name].clone()
        T[out_name].copy_(C0)
        _plan_run(be, stmt2, T)
        got2 = T[out_name].clone()
    finally:
        be.use_naive_contract(False)
    scale = ref.abs().max().item()
    assert (got1 - ref).abs().max().item() < 1e-12 * max(scale, 1.0) * 50
    assert (got2 - (ref + C0)).abs().max().item() < 1e-12 * max(scale, 1.0) * 50


F32_EXTRA = [
    # shapes that allow 16-byte (4-float) and 8-byte (2-float) cp.async chunks in the FP32 kernel
    ("C[ijab] = 0.5*A[ijef]*B[abef]", dict(A=(6, 6, 12, 12), B=(12, 12, 12, 12), C=(6, 6, 12, 12))),
    ("C[mn] = A[mk]*B[nk]", dict(A=(200, 512), B=(136, 512), C=(200, 136))),
    ("C[mn] = A[km]*B[kn]", dict(A=(320, 132), B=(320, 264), C=(132, 264))),
    ("C[ijab] = A[imae]*B[jbme] + A2[imae]*B2[jbme]",
     dict(A=(4, 4, 16, 16), A2=(4, 4, 16, 16), B=(4, 16, 4, 16), B2=(4, 16, 4, 16), C=(4, 4, 16, 16))),
]


@pytest.mark.parametrize("naive", [False, True])
@pytest.mark.parametrize("case", range(len(CASES) + len(F32_EXTRA)))
def test_contract_f32_vs_einsum(be, case, naive):
    """FP32 operands (precision='SP'): 3xTF32 tensor-core kernel and the FP32 naive kernel against an
    FP64 einsum of the same FP32 inputs.  Tolerance: FP32 rounding of a length-K dot product."""
    from pycc_b200.dsl import parse, plan
    from pycc_b200 import _lib
    stmt, shapes = (CASES + F32_EXTRA)[case]
    g = torch.Generator(device="cpu").manual_seed(100 + case)
    T = {k: torch.randn(s, generator=g, dtype=torch.float32).to(be.device) for k, s in shapes.items()}
    out_name, out_idx, acc, terms = parse(stmt)
    ref = torch.zeros(shapes[out_name], dtype=torch.float64, device=be.device)
    mag = torch.zeros_like(ref)
    for coef, fac in terms:
        spec = ",".join(f[1] for f in fac) + "->" + out_idx
        ref += coef * torch.einsum(spec, *[T[f[0]].double() for f in fac])
        mag += abs(coef) * torch.einsum(spec, *[T[f[0]].double().abs() for f in fac])
    be.use_naive_contract(naive)
    try:
        T[out_name].fill_(float("nan"))
        for L in plan(stmt, T, _lib.F32):
            be.contract(L.desc)
        got1 = T[out_name].double().clone()
        C0 = torch.randn(shapes[out_name], generator=g, dtype=torch.float32).to(be.device)
        T[out_name].copy_(C0)
        for L in plan(stmt.replace(" = ", " += ", 1), T, _lib.F32):
            be.contract(L.desc)
        got2 = T[out_name].double().clone()
    finally:
        be.use_naive_contract(False)
    # |error| <= c * eps32 * sum|a||b| (3xTF32 drops only the lo*lo term, ~2^-22 relative)
    tol = 4e-6 * mag + 1e-6
    assert ((got1 - ref).abs() <= tol).all()
    assert ((got2 - (ref + C0.double())).abs() <= tol + 1e-6 * C0.double().abs()).all()


def test_elementwise_vs_torch(be):
    no, nv = 5, 37
    g = torch.Generator(device="cpu").manual_seed(3)
    R = lambda *s: torch.randn(s, generator=g, dtype=torch.float64).to(be.device)
    X, t1 = R(no, no, nv, nv), R(no, nv)
    out = torch.empty_like(X)
    be.pair_combine(out, X, 2.0, -1.0, t1, t1, 0.5, no, nv)
    ref = 2.0 * X - X.transpose(2, 3) + 0.5 * t1[:, None, :, None] * t1[None, :, None, :]
    assert (out - ref).abs().max().item() < 1e-14
    r2 = X.clone()
    be.pair_symmetrize(r2, no, nv)
    assert (r2 - (X + X.permute(1, 0, 3, 2))).abs().max().item() < 1e-14
    # permute4
    P = torch.empty(nv, no, nv, no, dtype=torch.float64, device=be.device)
    be.permute4(P, X.permute(2, 0, 3, 1), 2.0, 0.0)
    assert (P - 2.0 * X.permute(2, 0, 3, 1)).abs().max().item() == 0.0
    # update + norm
    eps = torch.cat((torch.sort(-2 + torch.rand(no, generator=g, dtype=torch.float64))[0],
                     torch.sort(0.5 + torch.rand(nv, generator=g, dtype=torch.float64))[0])).to(be.device)
    t2, r1 = R(no, no, nv, nv), R(no, nv)
    eo, ev = eps[:no], eps[no:]
    Dia = eo[:, None] - ev[None, :]
    Dijab = eo[:, None, None, None] + eo[None, :, None, None] - ev[None, None, :, None] - ev[None, None, None, :]
    for symm in (False, True):
        a1, a2 = t1.clone(), t2.clone()
        rr = X + X.permute(1, 0, 3, 2) if symm else X
        rms2 = be.update_amplitudes(a1, a2, r1, X, eps, no, nv, symm)
        assert (a1 - (t1 + r1 / Dia)).abs().max().item() < 1e-13
        assert (a2 - (t2 + rr / Dijab)).abs().max().item() < 1e-13
        ref2 = ((r1 / Dia) ** 2).sum().item() + ((rr / Dijab) ** 2).sum().item()
        assert abs(rms2 - ref2) < 1e-11 * ref2
    # energy
    F = R(no + nv, no + nv)
    e = be.cc_energy(F[:no, no:], F.stride(0), t1, t2, X, no, nv)
    tau = t2 + t1[:, None, :, None] * t1[None, :, None, :]
    eref = (2.0 * (F[:no, no:] * t1).sum() + (tau * (2.0 * X - X.transpose(2, 3))).sum()).item()
    assert abs(e - eref) < 1e-11 * abs(eref)
    # DIIS algebra
    x, y = R(100003), R(100003)
    assert abs(be.dot(x, y) - torch.dot(x, y).item()) < 1e-10
    o2 = torch.empty_like(x)
    be.axpby(o2, 2.0, x, -3.0, y)
    assert (o2 - (2 * x - 3 * y)).abs().max().item() < 1e-14
    be.lincomb(o2, [x, y, x], [0.5, 0.25, -1.0])
    assert (o2 - (-0.5 * x + 0.25 * y)).abs().max().item() < 1e-14


@pytest.mark.parametrize("naive", [False, True])
@pytest.mark.parametrize("tag", ["sym", "gen"])
def test_residuals_golden(be, golden, tag, naive):
    from pycc_b200.engine import CCSDEngine
    g = golden
    eng = CCSDEngine(be, g.F, g.blocks, g.no, g.nv)
    eng.set_amplitudes(g[tag + "_t1"], g[tag + "_t2"])
    be.use_naive_contract(naive)
    try:
        r1, r2 = eng.residuals()
        got = dict(Fae=eng.Fae, Fmi=eng.Fmi, Fme=eng.Fme, Wmnij=eng.Wmnij, Wmbej=eng.Wmbej_view(),
                   Wmbje=eng.Wmbje_view(), Zmbij=eng.Zmbij_view(), r1=r1, r2=r2)
        for k, t in got.items():
            assert np.abs(t.cpu().numpy() - g[tag + "_" + k]).max() < 1e-12, k
        assert abs(eng.energy() - float(g[tag + "_energy"])) < 1e-12
    finally:
        be.use_naive_contract(False)


TC05_CASES = [
    ("C[ijab] = 0.5*A[ijef]*B[abef]", dict(A=(12, 12, 16, 16), B=(16, 16, 16, 16), C=(12, 12, 16, 16))),
    ("C[ijab] = 0.5*A[ijef]*B[abef]", dict(A=(13, 13, 18, 18), B=(18, 18, 18, 18), C=(13, 13, 18, 18))),   # ragged M, N, K
    ("C[mn] = A[mk]*B[nk]", dict(A=(300, 1000), B=(520, 1000), C=(300, 520))),
    ("C[nm] = A[mk]*B[nk]", dict(A=(260, 512), B=(300, 512), C=(300, 260))),                                # strided output
    ("C[mn] = A[mk]*B[nk] + A2[mk]*B2[nk]", dict(A=(256, 260), B=(256, 260), A2=(256, 260), B2=(256, 260), C=(256, 256))),
    # the epilogue that leaves through shared memory with lane = column (n contiguous in C): D read along n as well,
    ("C[iajb] = D[ijab] + A[iak]*B[jbk]", dict(A=(12, 12, 260), B=(20, 16, 260), D=(12, 20, 12, 16), C=(12, 12, 20, 16))),
    # compound m and n of the output (r2[ijaB] of the ring terms) with D read along m (thread = row mapping),
    ("C[ijab] = 2.0*D[jbia] + A[iak]*B[jbk]", dict(A=(11, 13, 300), B=(19, 17, 300), D=(19, 17, 11, 13), C=(11, 19, 13, 17))),
    # and a short M (t1[ie] <ja|be>: M = o rows in a 128-row tile)
    ("C[mn] = A[mk]*B[nk]", dict(A=(40, 300), B=(1000, 300), C=(40, 1000))),
]


def test_contract_f32_tcgen05_k_panels(be, monkeypatch):
    """The K-panel loop of the tcgen05 path (one segment, N >= 8 M, A planes larger than the panel budget:
    the SP ladder at (40,300)) with the budget lowered to 1 MB so that a small shape runs it: three panels
    of 512 + 512 + 376 k, the last one ragged against the 32-float k-block, beta = 1 from the second on.
    (M = 256, N = 2304: both whole tiles, so the planner keeps m on M and the launch stays tcgen05-eligible.)"""
    from pycc_b200.dsl import plan
    from pycc_b200 import _lib
    monkeypatch.setenv("PCCB_TC05_PANEL_MB", "1")
    g = torch.Generator(device="cpu").manual_seed(77)
    M, N, K = 256, 2304, 1400
    A = torch.randn(M, K, generator=g, dtype=torch.float32).to(be.device)
    B = torch.randn(N, K, generator=g, dtype=torch.float32).to(be.device)
    C0 = torch.randn(M, N, generator=g, dtype=torch.float32).to(be.device)
    los = []
    for t in (A, B):
        lo = torch.empty_like(t)
        be.split_lo(t, lo)
        be.register_lo_plane(t, lo)
        los.append(lo)
    ref = A.double() @ B.double().T
    mag = A.double().abs() @ B.double().abs().T
    tol = 4e-6 * mag + 1e-6
    try:
        for stmt, base in (("C[mn] = A[mk]*B[nk]", None), ("C[mn] += 0.5*A[mk]*B[nk]", C0)):
            C = torch.full((M, N), float("nan"), dtype=torch.float32, device=be.device) if base is None else base.clone()
            n0 = be.launch_count()
            for L in plan(stmt, dict(A=A, B=B, C=C), _lib.F32):
                be.contract(L.desc)
            assert be.launch_count() - n0 == 3          # three K panels, tcgen05 kernel each
            want = ref if base is None else base.double() + 0.5 * ref
            assert ((C.double() - want).abs() <= tol).all()
    finally:
        be._check(be.lib.pccb_clear_lo_planes(be.h))


@pytest.mark.parametrize("case", range(len(TC05_CASES)))
def test_contract_f32_tcgen05_vs_einsum(be, case):
    """FP32 plain-matrix contractions whose operands have registered residual planes run on the tcgen05 /
    TMEM kernel (3xTF32 in tensor memory); compared with an FP64 einsum of the same FP32 inputs and with
    the mma.sync 3xTF32 kernel (same accuracy model)."""
    from pycc_b200.dsl import parse, plan
    from pycc_b200 import _lib
    stmt, shapes = TC05_CASES[case]
    g = torch.Generator(device="cpu").manual_seed(300 + case)
    T = {k: torch.randn(s, generator=g, dtype=torch.float32).to(be.device) for k, s in shapes.items()}
    out_name, out_idx, acc, terms = parse(stmt)
    ref = torch.zeros(shapes[out_name], dtype=torch.float64, device=be.device)
    mag = torch.zeros_like(ref)
    for coef, fac in terms:
        spec = ",".join(f[1] for f in fac) + "->" + out_idx
        ref += coef * torch.einsum(spec, *[T[f[0]].double() for f in fac])
        mag += abs(coef) * torch.einsum(spec, *[T[f[0]].double().abs() for f in fac])
    # the mma.sync kernel first (no planes registered for these tensors yet)
    T[out_name].fill_(float("nan"))
    for L in plan(stmt, T, _lib.F32):
        be.contract(L.desc)
    got_sync = T[out_name].double().clone()
    los = {}
    for k, t in T.items():
        if k != out_name:
            los[k] = torch.empty_like(t)
            be.split_lo(t, los[k])
            be.register_lo_plane(t, los[k])
            hi = (t.view(torch.int32) & -8192).view(torch.float32)
            assert torch.equal(los[k], t - hi)
    n0 = be.launch_count()
    T[out_name].fill_(float("nan"))
    for L in plan(stmt, T, _lib.F32):
        be.contract(L.desc)
    got1 = T[out_name].double().clone()
    C0 = torch.randn(shapes[out_name], generator=g, dtype=torch.float32).to(be.device)
    T[out_name].copy_(C0)
    for L in plan(stmt.replace(" = ", " += ", 1), T, _lib.F32):
        be.contract(L.desc)
    got2 = T[out_name].double().clone()
    be._check(be.lib.pccb_clear_lo_planes(be.h))
    tol = 4e-6 * mag + 1e-6
    assert ((got1 - ref).abs() <= tol).all()
    assert ((got2 - (ref + C0.double())).abs() <= tol).all()
    assert ((got1 - got_sync).abs() <= tol).all()
    # one launch each (two when the number of m-tiles is odd: the paired-CTA kernel + the last tile column): no split-K
    # reduction, i.e. not the mma.sync path
    assert be.launch_count() - n0 in (2, 4)


@pytest.mark.parametrize("dtype", [torch.float64, torch.float32])
def test_permute4_forms(be, dtype):
    """pccb_permute4 picks one of three kernels: rows (the input's contiguous index stays last), 64 x 64 tiles (it becomes
    output index 0, 1 or 2) and the generic gather.  Every permutation of a ragged 4-index tensor, a narrowed (strided)
    source, and the accumulate form; exact (a permutation times a power of two)."""
    import itertools
    g = torch.Generator(device="cpu").manual_seed(21)
    X = torch.randn(5, 70, 9, 67, generator=g, dtype=torch.float64).to(dtype).to(be.device)
    for perm in itertools.permutations(range(4)):
        src = X.permute(*perm)
        out = torch.full(tuple(src.shape), float("nan"), dtype=dtype, device=be.device)
        be.permute4(out, src, 2.0, 0.0)
        assert torch.equal(out, 2.0 * src), perm
    for perm, dim in (((0, 3, 1, 2), 1), ((0, 1, 2, 3), 3), ((2, 3, 0, 1), 2)):
        src = X.permute(*perm).narrow(dim, 1, X.permute(*perm).shape[dim] - 2)
        out = torch.randn(tuple(src.shape), generator=g, dtype=torch.float64).to(dtype).to(be.device)
        want = 0.5 * src + 2.0 * out
        be.permute4(out, src, 0.5, 2.0)
        assert torch.equal(out, want), (perm, dim)


@pytest.mark.parametrize("dtype", [torch.float64, torch.float32])
def test_dot_many_vs_dot(be, dtype):
    """pccb_dot_many (one pass, one synchronisation: the new DIIS Gram row) against pccb_dot entry by entry and
    an FP64 torch dot; odd length (scalar tail) and an unaligned view (VEC = 1 path)."""
    g = torch.Generator(device="cpu").manual_seed(5)
    n = 1_000_003
    x = torch.randn(n + 1, generator=g, dtype=torch.float64).to(dtype).to(be.device)
    ys = [torch.randn(n + 1, generator=g, dtype=torch.float64).to(dtype).to(be.device) for _ in range(9)]
    for off in (0, 1):
        xv, yv = x[off:off + n], [y[off:off + n] for y in ys]
        got = be.dot_many(xv, yv)
        for k, y in enumerate(yv):
            ref = float((xv.double() * y.double()).sum())
            assert abs(got[k] - ref) <= 1e-9 * n ** 0.5
            assert abs(got[k] - be.dot(xv, y)) <= 1e-9 * n ** 0.5
