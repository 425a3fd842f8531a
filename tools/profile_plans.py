"""Time every planned pccb_contract launch of one CCSD residual individually (CUDA events,
3 repetitions after a warm-up): python tools/profile_plans.py o v  -> table + gpurun_out JSON."""
import sys, os, io, contextlib, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import numpy as np
import pycc_b200
from pycc_b200.synthetic import make_factor, calibrated_scale, device_blocks

o, v = int(sys.argv[1]), int(sys.argv[2])
B, eps, P = make_factor(o, v, 0)
B = B * calibrated_scale(B, eps, o, -0.10)
dev = torch.device("cuda", 0)
blocks = device_blocks(B, o, dev)
with contextlib.redirect_stdout(io.StringIO()):
    w = pycc_b200.ccwfn.from_arrays(np.diag(eps), blocks, o, model="CCSD")
eng, be = w.engine, w.backend
eng.residuals()
torch.cuda.synchronize()
rows, tot = [], 0.0
for name, launches in eng.plans.items():
    for L in launches:
        d = L.desc
        fl = sum(2.0 * d.M * d.N * d.batch * d.seg[q].K for q in range(d.nseg))
        be.contract(d)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(3):
            be.contract(d)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 3
        tot += ms
        rows.append(dict(builder=name, stmt=L.text, M=d.M, N=d.N, batch=d.batch, nseg=d.nseg,
                         K=[d.seg[q].K for q in range(d.nseg)], ms=ms, tflops=fl / ms * 1e-9))
        print("%-6s M=%-6d N=%-6d b=%-4d K=%-22s %8.3f ms %6.2f TF/s  %s" %
              (name, d.M, d.N, d.batch, rows[-1]["K"], ms, rows[-1]["tflops"], L.text[:70]), flush=True)
print("sum of launches: %.3f ms" % tot)
os.makedirs("gpurun_out", exist_ok=True)
json.dump(rows, open("gpurun_out/plans_%d_%d.json" % (o, v), "w"), indent=1)
