"""Wall-clock of the round-2 widened rows on one GPU (synthetic factorised integrals):
    python tools/time_lambda_triples.py [o v]
  * ccwfn.t3_density()            (pycc/ccwfn.py:721-855: all o^3 triplets, densities + Lambda sources)
  * one CC3 Lambda iteration      (pycc/cclambda.py:160-235: t3 and l3 of all o^3 triplets + 8 consumers each)
  * the one-time CC3 Lambda setup (pycc/cclambda.py:118-139)
Prints one JSON line."""
import contextlib, io, json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import pycc_b200
from pycc_b200.synthetic import make_factor, calibrated_scale, device_blocks

o, v = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (12, 80)
B, eps, P = make_factor(o, v, 0)
B = B * calibrated_scale(B, eps, o, -0.10)
dev = torch.device("cuda", 0)
out = dict(o=o, v=v)


def timed(fn, reps=1):
    torch.cuda.synchronize()
    t0 = time.time()
    for _ in range(reps):
        r = fn()
    torch.cuda.synchronize()
    return (time.time() - t0) / reps, r


with contextlib.redirect_stdout(io.StringIO()):
    w = pycc_b200.ccwfn.from_arrays(np.diag(eps), device_blocks(B, o, dev), o, model="CCSD(T)")
    ecc = w.solve_cc(e_conv=1e-10, r_conv=1e-10)
    w.t3_density()                                   # plans
    out["t3_density_s"], et = timed(w.t3_density)
    out["E_T_vikings"] = et
    out["E_T_tjl"] = pycc_b200.t_tjl(w)
    out["t3_density_tflops"] = o ** 3 * (12.0 * v ** 4 + 12.0 * o * v ** 3 + 2.0 * 12 * v ** 4) / out["t3_density_s"] * 1e-12
    del w
    torch.cuda.empty_cache()
    w = pycc_b200.ccwfn.from_arrays(np.diag(eps), device_blocks(B, o, dev), o, model="CC3")
    d = w.make_diis(8)
    out["cc3_iteration_s"], _ = timed(lambda: [w._iteration(d, n, 1) for n in (1, 2, 3)])
    out["cc3_iteration_s"] /= 3
    hb = pycc_b200.cchbar(w)
    lam = pycc_b200.cclambda(w, hb)
    out["cc3_lambda_setup_s"], _ = timed(lam.engine.cc3.setup)
    lam._iteration()
    out["cc3_lambda_iteration_s"], _ = timed(lam._iteration, 2)
print(json.dumps(out))
