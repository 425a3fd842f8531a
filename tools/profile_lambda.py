"""HBAR build and Lambda iteration at (o, v): whole-step times (CUDA events) and every planned
pccb_contract launch timed individually.  python tools/profile_lambda.py o v  -> table + JSON."""
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
    for it in range(3):
        w._iteration(None, it + 1, 1, extrapolate=False)
    hb = pycc_b200.cchbar(w)
    lam = pycc_b200.cclambda(w, hb)
be = w.backend


def timed(fn, reps):
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


res = dict(o=o, v=v)
res["hbar_ms"] = timed(hb.engine.build, 3)
res["hbar_flops"] = hb.engine.flops_issued()
res["lambda_iter_ms"] = timed(lam._iteration, 5)
res["lambda_flops"] = lam.engine.flops_issued()
print("HBAR build %.3f ms (%.2f TF/s issued)   Lambda iteration %.3f ms (%.2f TF/s issued)" %
      (res["hbar_ms"], res["hbar_flops"] / res["hbar_ms"] * 1e-9, res["lambda_iter_ms"],
       res["lambda_flops"] / res["lambda_iter_ms"] * 1e-9), flush=True)
rows = []
for tag, plans in (("hbar", hb.engine.plans), ("lambda", lam.engine.plans)):
    tot = 0.0
    for name, launches in plans.items():
        for L in launches:
            d = L.desc
            fl = sum(2.0 * d.M * d.N * d.batch * d.seg[q].K for q in range(d.nseg))
            ms = timed(lambda: be.contract(d), 3)
            tot += ms
            rows.append(dict(part=tag, builder=name, stmt=L.text, M=d.M, N=d.N, batch=d.batch, nseg=d.nseg,
                             K=[d.seg[q].K for q in range(d.nseg)], ms=ms, tflops=fl / ms * 1e-9))
            print("%-6s M=%-7d N=%-7d b=%-4d K=%-18s %8.3f ms %6.2f TF/s  %s" %
                  (name, d.M, d.N, d.batch, rows[-1]["K"], ms, rows[-1]["tflops"], L.text[:64]), flush=True)
    print("%s: sum of launches %.3f ms" % (tag, tot))
res["launches"] = rows
os.makedirs("gpurun_out", exist_ok=True)
json.dump(res, open("gpurun_out/lambda_%d_%d.json" % (o, v), "w"), indent=1)
