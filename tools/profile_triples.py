"""(T) on a few occupied pairs of a synthetic system (for ncu launch lists / captures):
python tools/profile_triples.py o v npairs"""
import sys, os, io, contextlib, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import numpy as np
import pycc_b200
from pycc_b200.synthetic import make_factor, calibrated_scale, device_blocks
import pycc_b200.cctriples as cct

o, v, npairs = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
B, eps, P = make_factor(o, v, 0)
B = B * calibrated_scale(B, eps, o, -0.10)
dev = torch.device("cuda", 0)
blocks = device_blocks(B, o, dev)
with contextlib.redirect_stdout(io.StringIO()):
    w = pycc_b200.ccwfn.from_arrays(np.diag(eps), blocks, o, model="CCSD(T)")
tri = cct._engine(w)
pairs = tri.pairs()[-npairs:]
tri.energy(pairs[:1])
torch.cuda.synchronize()
t0 = time.perf_counter()
et, per, trip = tri.energy(pairs)
torch.cuda.synchronize()
dt = time.perf_counter() - t0
fl = len(trip) * (12 * v**4 + 12 * o * v**3)
print("pairs %d triplets %d: %.4f s, %.2f TFLOP/s, E=%.12e" % (len(pairs), len(trip), dt, fl / dt / 1e12, et))
