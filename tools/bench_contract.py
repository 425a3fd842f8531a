"""Quick timing of single pccb_contract launches at CCSD shapes (CUDA events, L2 not flushed:
operands >> L2).  Usage: python tools/bench_contract.py [o v]"""
import sys, os, json, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from pycc_b200.backend import CudaBackend
from pycc_b200.dsl import plan

o, v = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (20, 150)
only = sys.argv[3].split(",") if len(sys.argv) > 3 else None
be = CudaBackend()
dev = be.device
F32 = os.environ.get("PCCB_BENCH_DTYPE", "f64") == "f32"
DT = torch.float32 if F32 else torch.float64
from pycc_b200 import _lib
R = lambda *s: torch.randn(s, dtype=DT, device=dev)
T = dict(tau=R(o, o, v, v), vvvv=R(v, v, v, v), r2=R(o, o, v, v), t2=R(o, o, v, v), Wr=R(o, v, o, v),
         ovvv=R(o, v, v, v), Zr=R(o, o, o, v), Wmnij=R(o, o, o, o), Loovv=R(o, o, v, v), Fae=R(v, v), t1=R(o, v))
tests = [("ladder", "r2[ijab] += 0.5*tau[ijef]*vvvv[abef]", 2.0 * o * o * v ** 4),
         ("ring", "r2[ijab] += t2[imae]*Wr[jbme]", 2.0 * o ** 3 * v ** 3),
         ("ring2", "r2[ijab] += t2[mjae]*Wr[ibme]", 2.0 * o ** 3 * v ** 3),
         ("Wmbej", "Wr[jbme] = tau[jnfb]*Loovv[mnef]", 2.0 * o ** 3 * v ** 3),
         ("Zmbij", "Zr[mijb] = ovvv[mbef]*tau[ijef]", 2.0 * o ** 3 * v ** 3),
         ("Wmnij", "Wmnij[mnij] = Loovv[mnef]*tau[ijef]", 2.0 * o ** 4 * v ** 2),
         ("Fae", "Fae[ae] = tau[mnaf]*Loovv[mnef]", 2.0 * o ** 2 * v ** 3),
         ("t1ovvv", "r2[ijab] += t1[ie]*ovvv[jabe]", 2.0 * o ** 2 * v ** 3)]
res = {}
for name, stmt, fl in tests:
    if only and name not in only: continue
    Ls = plan(stmt, T, _lib.F32 if F32 else _lib.F64)
    for _ in range(2):
        for L in Ls: be.contract(L.desc)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n = 5
    e0.record()
    for _ in range(n):
        for L in Ls: be.contract(L.desc)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / n
    res[name] = dict(ms=ms, tflops=fl / ms * 1e-9)
    print(name, "%.3f ms  %.2f TFLOP/s" % (ms, fl / ms * 1e-9), flush=True)
# cuBLAS comparison for the ladder shape
A = T["tau"].reshape(o * o, v * v); B = T["vvvv"].reshape(v * v, v * v)
for _ in range(2): torch.matmul(A, B.t())
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(5): torch.matmul(A, B.t())
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 5
print("cublas ladder %.3f ms %.2f TFLOP/s" % (ms, 2.0 * o * o * v ** 4 / ms * 1e-9))
res["cublas_ladder"] = dict(ms=ms, tflops=2.0 * o * o * v ** 4 / ms * 1e-9)
json.dump(res, open("gpurun_out/bench_contract_%d_%d.json" % (o, v), "w"), indent=1)
