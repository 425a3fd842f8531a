"""Opcode histogram per kernel of the in-tree libpyccb200.so (no GPU needed):

    python tools/sass_histogram.py [path/to/lib.so] > profiles/rNN_sass_histogram.txt

Runs ``cuobjdump -sass`` and counts, per kernel, the mnemonics that show which hardware path a kernel uses:
DMMA (FP64 tensor), HMMA (mma.sync tf32/f16), UTCHMMA / UTCBAR / LDTM / STTM (tcgen05 + tensor memory),
UTMALDG / UTMASTG (TMA tensor copies), UBLKCP (bulk copies), LDGSTS (cp.async), SYNCS (mbarrier), LDG/STG width.
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "pycc_b200", "libpyccb200.so")
WATCH = ("DMMA", "HMMA", "UTCHMMA", "UTCQMMA", "UTCBAR", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UBLKCP", "LDGSTS",
         "SYNCS", "LDG.E.128", "LDG.E.64", "STG.E.128", "STG.E.64", "LDS.64", "LDS.128", "STS.64", "STS.128", "BAR",
         "USETMAXREG", "DFMA", "FFMA")

sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True, check=True).stdout
names = subprocess.run(["c++filt"], input="\n".join(re.findall(r"Function : (\S+)", sass)), capture_output=True,
                       text=True).stdout.splitlines()
kern, cur = collections.OrderedDict(), None
it = iter(names)
for line in sass.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = next(it)
        cur = re.sub(r"\(anonymous namespace\)::", "", cur)
        cur = re.sub(r"\(.*$", "", cur)                  # drop the parameter list
        kern[cur] = collections.Counter()
        continue
    m = re.match(r"\s+/\*[0-9a-f]{4,6}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
    if m and cur is not None:
        op = m.group(1)
        kern[cur]["_total"] += 1
        for w in WATCH:
            if op == w or op.startswith(w + ".") or (w.count(".") and op.startswith(w)):
                kern[cur][w] += 1

print("# cuobjdump -sass %s : instructions per kernel (static counts), sm_100a" % os.path.relpath(lib, ROOT))
print("# columns: total | " + " ".join(WATCH))
for k, c in kern.items():
    hot = " ".join("%s=%d" % (w, c[w]) for w in WATCH if c[w])
    print("%-78s %6d | %s" % (k[:78], c["_total"], hot))
