"""What ONE rank of a W-rank CCSD job does, measured on a single GPU (no communication):
python tools/profile_rank.py o v world [rank]

The rank's virtual slice, its rows of <ab|ef> / <mb|ef> and its plans are exactly those of the
sharded run (pycc_b200.parallel.virtual_slice); the allreduce is a no-op.  Prints
  * the iteration time (CUDA events over 10 iterations, as bench.py times it),
  * the host time to enqueue one iteration (how far ahead of the GPU the launching thread can get),
  * the sum of the individual launch times of the residual (each timed alone),
so that launch-bound, quantisation-bound and redundant (unsharded) shares can be told apart.
"""
import sys, os, io, contextlib, json, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import numpy as np
import pycc_b200
import pycc_b200.parallel as par
from pycc_b200.synthetic import make_factor, calibrated_scale, device_blocks
from pycc_b200.utils import helper_diis

o, v, world = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
rank = int(sys.argv[4]) if len(sys.argv) > 4 else 0


class FakeComm(par.Comm):
    def __init__(self, group=None):
        self.group = None
        self.rank, self.world = rank, world
        self.bytes_reduced = 0

    def allreduce_(self, *tensors):
        return

    def allreduce_scalar(self, x):
        return float(x)

    def barrier(self):
        return


par.Comm = FakeComm
B, eps, P = make_factor(o, v, 0)
B = B * calibrated_scale(B, eps, o, -0.10)
dev = torch.device("cuda", 0)
vs = par.virtual_slice(v, rank, world)
blocks = device_blocks(B, o, dev, vslice=vs, keep_full_ovvv=False)
with contextlib.redirect_stdout(io.StringIO()):
    w = pycc_b200.ccwfn.from_arrays(np.diag(eps), blocks, o, model="CCSD")
eng, be = w.engine, w.backend
diis = helper_diis(w.t1, w.t2, 8, "DP", backend=be)
n = 0
for _ in range(3):
    n += 1
    w._iteration(diis, n, 1)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10):
    n += 1
    w._iteration(diis, n, 1)
e1.record()
torch.cuda.synchronize()
it_ms = e0.elapsed_time(e1) / 10

# residual only (no host synchronisation inside)
e0.record()
for _ in range(10):
    eng.residuals(symmetrize=False)
e1.record()
torch.cuda.synchronize()
res_ms = e0.elapsed_time(e1) / 10
t0 = time.perf_counter()
for _ in range(10):
    eng.residuals(symmetrize=False)
host_ms = (time.perf_counter() - t0) * 100
torch.cuda.synchronize()

rows, tot = [], 0.0
for name, launches in eng.plans.items():
    for L in launches:
        d = L.desc
        fl = sum(2.0 * d.M * d.N * d.batch * d.seg[q].K for q in range(d.nseg))
        be.contract(d)
        e0.record()
        for _ in range(3):
            be.contract(d)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 3
        tot += ms
        rows.append(dict(builder=name, stmt=L.text, M=d.M, N=d.N, batch=d.batch,
                         K=[d.seg[q].K for q in range(d.nseg)], ms=ms, tflops=fl / ms * 1e-9))
        print("%-6s M=%-6d N=%-6d b=%-4d K=%-22s %8.3f ms %6.2f TF/s  %s" %
              (name, d.M, d.N, d.batch, rows[-1]["K"], ms, rows[-1]["tflops"], L.text[:70]), flush=True)
nl = sum(len(x) for x in eng.plans.values())
print("rank %d of %d, slice %s: iteration %.3f ms | residuals() %.3f ms on the GPU, %.3f ms of host time "
      "to enqueue (%d contractions) | sum of contraction launches timed alone %.3f ms"
      % (rank, world, vs, it_ms, res_ms, host_ms, nl, tot))
os.makedirs("gpurun_out", exist_ok=True)
json.dump(dict(o=o, v=v, world=world, rank=rank, iteration_ms=it_ms, residuals_ms=res_ms, host_enqueue_ms=host_ms,
               sum_launch_ms=tot, launches=rows), open("gpurun_out/rank_%d_%d_w%d.json" % (o, v, world), "w"), indent=1)
