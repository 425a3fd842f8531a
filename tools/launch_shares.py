"""Share table of an ncu launch list (gpu__time_duration.sum per launch):

    python tools/launch_shares.py gpurun_out/launches.csv [title] > profiles/rNN_iteration_shares_*.txt

Groups by kernel name (template arguments kept), prints launches, total / mean time and share; then the raw
launch order with grid sizes.  ncu times are cold-cache and serialised: the SHARES are what is comparable with the
CUDA-event time of the step."""
import collections
import csv
import re
import sys

rows = [r for r in csv.reader(open(sys.argv[1])) if r]
hi = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
hdr = rows[hi]
ix = {h: i for i, h in enumerate(hdr)}
recs = []
for r in rows[hi + 1:]:
    if len(r) < len(hdr) or r[ix["Metric Name"]] != "gpu__time_duration.sum":
        continue
    val = float(r[ix["Metric Value"]].replace(",", ""))
    unit = r[ix["Metric Unit"]]
    us = val * {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6, "nsecond": 1e-3, "usecond": 1.0, "msecond": 1e3, "second": 1e6}.get(unit, 1e-3)
    name = re.sub(r"\(anonymous namespace\)::", "", r[ix["Kernel Name"]])
    name = re.sub(r"\(.*$", "", name)
    recs.append((name, us, r[ix["Grid Size"]] if "Grid Size" in ix else "", r[ix["Block Size"]] if "Block Size" in ix else ""))
tot = sum(x[1] for x in recs)
print("# %s" % (sys.argv[2] if len(sys.argv) > 2 else sys.argv[1]))
print("# %d launches, %.3f ms summed (ncu gpu__time_duration, --clock-control none)" % (len(recs), tot / 1e3))
agg = collections.OrderedDict()
for name, us, g, b in recs:
    a = agg.setdefault(name, [0, 0.0])
    a[0] += 1
    a[1] += us
print("%-70s %5s %12s %10s %7s" % ("kernel", "n", "total us", "mean us", "share"))
for name, (n, us) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print("%-70s %5d %12.1f %10.1f %6.2f%%" % (name[:70], n, us, us / n, 100.0 * us / tot))
print("\n# launch order")
for q, (name, us, g, b) in enumerate(recs):
    print("%4d %-62s %10.1f us  grid %s block %s" % (q, name[:62], us, g, b))
