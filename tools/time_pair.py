"""(T) W build + energy kernel of a few occupied pairs, timed with CUDA events:
    python tools/time_pair.py o v [i,j ...]        (env knobs: PCCB_TMAE, PCCB_EPI_MI, PCCB_T_PLAIN, PCCB_WSE_MI)
Prints ms per pair, TFLOP/s of the restricted-triplet flop count, and checks the pair energies against a second run."""
import contextlib, io, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import pycc_b200
import pycc_b200.cctriples as cct
from pycc_b200.synthetic import make_factor, calibrated_scale, device_blocks

o, v = int(sys.argv[1]), int(sys.argv[2])
pairs = [tuple(int(x) for x in a.split(",")) for a in sys.argv[3:]] or [(o - 1, o - 1), (o - 1, o // 2)]
B, eps, P = make_factor(o, v, 0)
B = B * calibrated_scale(B, eps, o, -0.10)
blocks = device_blocks(B, o, torch.device("cuda", 0))
with contextlib.redirect_stdout(io.StringIO()):
    w = pycc_b200.ccwfn.from_arrays(np.diag(eps), blocks, o, model="CCSD(T)")
d = w.make_diis(8)
for n in range(1, 4):
    w._iteration(d, n, 1)
tri = cct._engine(w)
e_first, _, _ = tri.energy(pairs)
for (i, j) in pairs:
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(3):
        et, per, trip = tri.energy([(i, j)])
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 3
    fl = (j + 1) * (12.0 * v**4 + 12.0 * o * v**3)
    print("pair (%d,%d): %d triplets %.3f ms  %.2f TFLOP/s  E=%.15e" % (i, j, j + 1, ms, fl / ms * 1e-9, et))
print("sum over pairs, first run: %.15e" % e_first)
