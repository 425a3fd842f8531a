"""Achieved HBM bandwidth of the memory-bound kernels (elementwise.cu) through the C ABI, at sizes beyond L2:

    python tools/bench_elementwise.py [o v] > profiles/rNN_elementwise.txt      (default 40 300, FP64: t2 = 1.15 GB)

Bytes are ALGORITHMIC (each tensor the kernel must read or write counted once); time = CUDA events around 5
back-to-back launches after a warm-up, every tensor larger than the 126 MB L2.  Peak = MEASURED_PEAKS.json hbm_gbs."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from pycc_b200.backend import CudaBackend

o, v = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (40, 300)
try:
    peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    src = "MEASURED_PEAKS.json"
except Exception:
    peak, src = 6536.0, "fallback (earlier MEASURED_PEAKS.json of this pool)"
be = CudaBackend()
dev, dt, w = be.device, torch.float64, 8
g = torch.Generator(device="cpu").manual_seed(0)
N2, N1 = o * o * v * v, o * v
R = lambda *s: torch.randn(*s, dtype=dt, device=dev) * 0.01
t1, t2, r1, r2, oovv = R(o, v), R(o, o, v, v), R(o, v), R(o, o, v, v), R(o, o, v, v)
eps = torch.cat([-2.0 + torch.rand(o, dtype=dt, device=dev), 0.5 + torch.rand(v, dtype=dt, device=dev)])
F = torch.diag(eps).contiguous()
fov = F[:o, o:]
outs = {k: torch.empty(o, o, v, v, dtype=dt, device=dev) for k in ("tau", "tauh", "tauW", "t2s")}
perm = {k: torch.empty(o, v, o, v, dtype=dt, device=dev) for k in ("t2sP", "t2P", "t2Q", "tauWPn", "t2Rh")}
x, y, z = R(N2), R(N2), R(N2)
vecs = [R(N2) for _ in range(8)]


def timed(fn, reps=5):
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps * 1e3          # us


rows = []


def row(name, ref, nbytes, fn):
    us = timed(fn)
    rows.append((name, ref, nbytes, us, nbytes / us * 1e-3))


be.bind_stream()
row("update_energy_kernel (update + norm + energy, symmetrising read)", "ccwfn.py:287-294,621,712-719", 5 * w * N2,
    lambda: be.update_energy(t1, t2, r1, r2, eps, fov, F.stride(0), oovv, o, v, True, 0, o * o, True))
row("update_energy_kernel (r2 already symmetrised: multi-GPU path)", "ccwfn.py:287-294,712-719", 4 * w * N2,
    lambda: be.update_energy(t1, t2, r1, r2, eps, fov, F.stride(0), oovv, o, v, False, 0, o * o, True))
row("update_energy_kernel without energy (pccb_update_amplitudes)", "ccwfn.py:287-294,621", 4 * w * N2,
    lambda: be.update_amplitudes(t1, t2, r1, r2, eps, o, v, True))
row("energy_kernel (pccb_cc_energy)", "ccwfn.py:712-719", 2 * w * N2, lambda: be.cc_energy(fov, F.stride(0), t1, t2, oovv, o, v))
row("prepare_kernel: tau, tauh, tauW, t2s (1 read + 4 writes)", "ccwfn.py:408-410,554", 5 * w * N2,
    lambda: be.prepare_amplitudes(t1, t2, o, v, 0, v, **outs))
row("prepare_kernel: + t2sP t2P t2Q tauWPn t2Rh (1 read + 9 writes)", "ccwfn.py:408-410,498-505,609-611", 10 * w * N2,
    lambda: be.prepare_amplitudes(t1, t2, o, v, 0, v, **outs, **perm))
row("pair_combine_kernel (2 t2 - t2^T)", "ccwfn.py:554", 2 * w * N2, lambda: be.pair_combine(outs["t2s"], t2, 2.0, -1.0, None, None, 0.0, o, v))
row("pair_symmetrize_kernel (r2 += r2^T in place)", "ccwfn.py:621", 2 * w * N2, lambda: be.pair_symmetrize(r2, o, v))
row("dot_vec_kernel (DIIS Gram entry)", "utils.py:108-112", 2 * w * N2, lambda: be.dot(x, y))
row("axpby_vec_kernel (DIIS error vector)", "utils.py:26-46", 3 * w * N2, lambda: be.axpby(z, 1.0, x, -1.0, y))
row("lincomb_vec_kernel, 8 vectors (DIIS extrapolation)", "utils.py:129-133", 9 * w * N2, lambda: be.lincomb(z, vecs, [0.125] * 8))
row("permute4_rows_kernel [i,a,m,e] <- [i,m,a,e] (set-up only)", "set-up", 2 * w * N2, lambda: be.permute4(perm["t2P"], t2.permute(0, 2, 1, 3)))
tauT = torch.empty(v, v, o, o, dtype=t2.dtype, device=t2.device)
row("permute4_tiles_kernel [a,b,m,n] <- [m,n,a,b] (tauT, FP32 mode)", "layout", 2 * w * N2, lambda: be.permute4(tauT, t2.permute(2, 3, 0, 1)))

print("# element-wise kernels of libpyccb200.so at o=%d v=%d FP64 (o^2 v^2 tensor = %.2f GB), B200, CUDA events" % (o, v, w * N2 / 1e9))
print("# peak = %.1f GB/s (%s: copy bandwidth, read + write bytes)" % (peak, src))
print("%-68s %-32s %10s %10s %9s %6s" % ("kernel", "reference lines", "bytes", "us", "GB/s", "frac"))
for name, ref, nb, us, gbs in rows:
    print("%-68s %-32s %10.3e %10.1f %9.1f %6.3f" % (name, ref, nb, us, gbs, gbs / peak))
