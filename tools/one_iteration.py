"""One CCSD iteration (and optionally one (T) occupied pair) of the shipped code between cudaProfilerStart/Stop,
for an ncu launch list of exactly that region:

    ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv \\
        --log-file gpurun_out/launches.csv python tools/one_iteration.py 30 250 [--sp] [--pair i,j] [--world W]

``--world W`` runs rank 0's share of a W-rank job on one GPU (collectives are no-ops), as tools/profile_rank.py does.
Summarise with tools/launch_shares.py.
"""
import argparse
import contextlib
import io
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

ap = argparse.ArgumentParser()
ap.add_argument("o", type=int)
ap.add_argument("v", type=int)
ap.add_argument("--sp", action="store_true")
ap.add_argument("--pair", default=None, help="also run the (T) W build + energy kernel of this occupied pair i,j")
ap.add_argument("--world", type=int, default=1)
ap.add_argument("--only-pair", action="store_true", help="profile only the (T) pair, not the iteration")
args = ap.parse_args()

import pycc_b200
import pycc_b200.parallel as par
from pycc_b200.synthetic import make_factor, calibrated_scale, device_blocks

if args.world > 1:
    class FakeComm(par.Comm):
        def __init__(self, group=None):
            self.group, self.rank, self.world, self.bytes_reduced = None, 0, args.world, 0

        def allreduce_(self, *t):
            return

        def reduce_scatter_(self, *a):
            return

        def allreduce_scalar(self, x):
            return float(x)

        def allreduce_scalars(self, xs):
            return [float(x) for x in xs]

        def allgather_into(self, full, chunk, in_place=False):
            return

        def barrier(self):
            return
    par.Comm = FakeComm

o, v = args.o, args.v
B, eps, P = make_factor(o, v, 0)
B = B * calibrated_scale(B, eps, o, -0.10)
dev = torch.device("cuda", 0)
dt = torch.float32 if args.sp else torch.float64
want_t = args.pair is not None
blocks = device_blocks(B, o, dev, vslice=par.virtual_slice(v, 0, args.world), keep_full_ovvv=want_t or args.world == 1, dtype=dt)
with contextlib.redirect_stdout(io.StringIO()):
    w = pycc_b200.ccwfn.from_arrays(np.diag(eps), blocks, o, model="CCSD(T)" if want_t else "CCSD",
                                    precision="SP" if args.sp else "DP")
diis = w.make_diis(8)
n = 0
for _ in range(3):
    n += 1
    w._iteration(diis, n, 1)
tri = None
if want_t:
    import pycc_b200.cctriples as cct
    tri = cct._engine(w)
    i, j = [int(x) for x in args.pair.split(",")]
    tri.energy([(i, j)])
torch.cuda.synchronize()
torch.cuda.profiler.start()
if not args.only_pair:
    n += 1
    w._iteration(diis, n, 1)
if want_t:
    tri.energy([(i, j)])
torch.cuda.synchronize()
torch.cuda.profiler.stop()
print("done: o=%d v=%d world=%d sp=%s pair=%s" % (o, v, args.world, args.sp, args.pair))
