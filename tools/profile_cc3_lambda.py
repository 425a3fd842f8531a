import contextlib, io, os, sys
sys.path.insert(0, "/root/repo")
import numpy as np, torch
import pycc_b200
from pycc_b200.synthetic import make_factor, calibrated_scale, device_blocks
o, v = 16, 120
B, eps, P = make_factor(o, v, 0)
B = B * calibrated_scale(B, eps, o, -0.10)
dev = torch.device("cuda", 0)
with contextlib.redirect_stdout(io.StringIO()):
    w = pycc_b200.ccwfn.from_arrays(np.diag(eps), device_blocks(B, o, dev), o, model="CC3")
    d = w.make_diis(8)
    w._iteration(d, 1, 1)
    lam = pycc_b200.cclambda(w, pycc_b200.cchbar(w))
    lam._iteration()
    torch.cuda.synchronize()
    torch.cuda.cudart().cudaProfilerStart()
    lam._iteration()
    torch.cuda.synchronize()
    torch.cuda.cudart().cudaProfilerStop()
