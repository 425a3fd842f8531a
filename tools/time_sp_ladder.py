"""The two big tcgen05 3xTF32 products of the mixed-precision iteration, timed alone with CUDA events:
    python tools/time_sp_ladder.py [o v]       ladder M = o^2, N = v^2, K = v^2 ; ring M = N = K = o v
(env knobs: PCCB_TC05_PREFETCH, PCCB_TC05_PAIRS, PCCB_TC05_PANEL_MB)"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from pycc_b200.backend import CudaBackend
from pycc_b200.dsl import plan
from pycc_b200 import _lib
o, v = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (40, 300)
be = CudaBackend()
dev = be.device
def run(M, N, K, name):
    A = torch.randn(M, K, dtype=torch.float32, device=dev); B = torch.randn(N, K, dtype=torch.float32, device=dev)
    C = torch.zeros(M, N, dtype=torch.float32, device=dev)
    los = []
    for t in (A, B):
        lo = torch.empty_like(t); be.split_lo(t, lo); be.register_lo_plane(t, lo); los.append(lo)
    Ls = plan("C[mn] = A[mk]*B[nk]", dict(A=A, B=B, C=C), _lib.F32)
    for L in Ls: be.contract(L.desc)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(3):
        for L in Ls: be.contract(L.desc)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 3
    # spot check against FP64 on a few rows
    idx = torch.tensor([0, M // 2, M - 1], device=dev)
    ref = A[idx].double() @ B.double().T
    err = ((C[idx].double() - ref).abs().max() / ref.abs().max()).item()
    print("%s M=%d N=%d K=%d: %.3f ms  %.1f TFLOP/s FP32-accurate  rel.err %.1e" % (name, M, N, K, ms, 2.0 * M * N * K / ms * 1e-9, err))
    be._check(be.lib.pccb_clear_lo_planes(be.h))
    del A, B, C, los
    torch.cuda.empty_cache()
run(o * o, v * v, v * v, "ladder")
run(o * v, o * v, o * v, "ring  ")
