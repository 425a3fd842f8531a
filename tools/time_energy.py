"""(T) energy kernel alone on the W batch of one pair, repeated back to back (W stays in L2 when it fits):
    python tools/time_energy.py o v i,j [reps]"""
import contextlib, io, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import pycc_b200
import pycc_b200.cctriples as cct
from pycc_b200.synthetic import make_factor, calibrated_scale, device_blocks

o, v = int(sys.argv[1]), int(sys.argv[2])
pairs = [tuple(int(x) for x in a.split(",")) for a in sys.argv[3:] if "," in a]
reps = 20
B, eps, P = make_factor(o, v, 0)
B = B * calibrated_scale(B, eps, o, -0.10)
blocks = device_blocks(B, o, torch.device("cuda", 0))
with contextlib.redirect_stdout(io.StringIO()):
    w = pycc_b200.ccwfn.from_arrays(np.diag(eps), blocks, o, model="CCSD(T)")
d = w.make_diis(8)
for n in range(1, 3):
    w._iteration(d, n, 1)
tri = cct._engine(w)
eng, be = tri.eng, tri.be
tri.refresh()
fov = eng.F[:o, o:]
for (i, j) in pairs:
    n = j + 1
    trip = [(i, j, k) for k in range(n)]
    ijk = torch.tensor(trip, dtype=torch.int32).reshape(-1).to(be.device)
    et = be.zeros((n,), torch.float64)
    tri._launch_pair(i, j)
    args = (tri.W, ijk, n, eng.t1, eng.t2, eng.blk["oovv"], fov, eng.F.stride(0), eng.eps, o, v, et)
    be.triples_energy(*args)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        be.triples_energy(*args)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    gb = n * v**3 * tri.W.element_size() / 1e9
    print("pair (%d,%d): %d triplets, W %.3f GB: %.4f ms per call = %.2f TB/s, E=%.15e" % (i, j, n, gb, ms, gb / ms, et.sum().item()))
