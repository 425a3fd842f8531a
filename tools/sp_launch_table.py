"""Every planned launch of one mixed-precision CCSD iteration, timed alone (CUDA events, 2nd of 2 runs), sorted by time:
    python tools/sp_launch_table.py [o v] [--dp]
Columns: ms, TFLOP/s (2 M N K per segment), builder, M N batch [K per segment], statement text."""
import contextlib, io, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import pycc_b200
from pycc_b200.synthetic import make_factor, calibrated_scale, device_blocks
args = [a for a in sys.argv[1:] if not a.startswith("--")]
o, v = (int(args[0]), int(args[1])) if len(args) > 1 else (40, 300)
dp = "--dp" in sys.argv
B, eps, P = make_factor(o, v, 0)
B = B * calibrated_scale(B, eps, o, -0.10)
dev = torch.device("cuda", 0)
dt = torch.float64 if dp else torch.float32
with contextlib.redirect_stdout(io.StringIO()):
    w = pycc_b200.ccwfn.from_arrays(np.diag(eps), device_blocks(B, o, dev, dtype=dt), o, model="CCSD", precision="DP" if dp else "SP")
d = w.make_diis(8)
for n in (1, 2):
    w._iteration(d, n, 1)
eng = w.engine
if getattr(eng, "sp_tc", False):       # the engine registers these two planes only while r_T2 runs; here launches run alone
    eng.be.register_lo_plane(eng.t1, eng._lo["t1"])
    eng.be.register_lo_plane(eng.t2, eng._lo["t2"])
VARIANTS = [("mma.sync", {"PCCB_TC05_MIN_K": "2000000000"}), ("tc05 row-epi", {"PCCB_TC05_EPI": "0"}), ("tc05", {})]
rows = []
plans = dict(eng.plans)
plans.update(getattr(eng, "pre_plans", {}) or {})


def timed(ds, env):
    for k, x in env.items():
        os.environ[k] = x
    try:
        for rep in range(2):
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); eng.be.contract(ds); e1.record(); torch.cuda.synchronize()
    finally:
        for k in env:
            del os.environ[k]
    return e0.elapsed_time(e1)


for name, launches in plans.items():
    for L in launches:
        ds = L.desc
        fl = sum(2.0 * ds.M * ds.N * ds.batch * ds.seg[q].K for q in range(ds.nseg))
        vs = VARIANTS if not dp else VARIANTS[-1:]
        ms = [timed(ds, env) for _, env in vs]
        if not dp and ms[0] > 1.3 * ms[-1]:        # a tcgen05 launch: once more in the opposite order (clocks drift under load)
            again = [timed(ds, env) for _, env in reversed(vs)][::-1]
            ms = [min(a, b) for a, b in zip(ms, again)]
        rows.append((ms[-1], fl / ms[-1] * 1e-9 if ms[-1] > 0 else 0.0, name, ds.M, ds.N, ds.batch,
                     [ds.seg[q].K for q in range(ds.nseg)], getattr(L, "text", ""), ms))
rows.sort(key=lambda r: -r[0])
tot = sum(r[0] for r in rows)
print("# %d launches, %.1f ms summed, o=%d v=%d %s" % (len(rows), tot, o, v, "FP64" if dp else "FP32 contraction mode"))
print("# ms per variant: " + " | ".join(n for n, _ in VARIANTS))
for r in rows[:40]:
    print("%9.3f ms %7.1f TF  %-6s M=%-7d N=%-7d b=%-3d K=%s  %s" % r[:8] + "   [" + " | ".join("%.3f" % x for x in r[8]) + "]")
