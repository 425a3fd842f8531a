"""One CC3 amplitude iteration between cudaProfilerStart/Stop (for an ncu launch list): python tools/profile_cc3.py [o v]"""
import contextlib, io, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import pycc_b200
from pycc_b200.synthetic import make_factor, calibrated_scale, device_blocks
o, v = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (16, 120)
B, eps, P = make_factor(o, v, 0)
B = B * calibrated_scale(B, eps, o, -0.10)
with contextlib.redirect_stdout(io.StringIO()):
    w = pycc_b200.ccwfn.from_arrays(np.diag(eps), device_blocks(B, o, torch.device("cuda", 0)), o, model="CC3")
    d = w.make_diis(8)
    w._iteration(d, 1, 1)
    w._iteration(d, 2, 1)
    torch.cuda.synchronize()
    torch.cuda.cudart().cudaProfilerStart()
    w._iteration(d, 3, 1)
    torch.cuda.synchronize()
    torch.cuda.cudart().cudaProfilerStop()
