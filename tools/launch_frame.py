#!/usr/bin/env python
"""Print the kernel sequence of the last complete frame in an ncu launch list (gpu__time_duration CSV), per-kernel totals."""
import csv
import sys
from collections import OrderedDict

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 5]
hdr = rows[0]
ki, vi = hdr.index("Kernel Name"), hdr.index("Metric Value")
seq = [(r[ki].replace("rt::<unnamed>::", "").replace("void ", "").split("(")[0], float(r[vi].replace(",", "")) / 1e6) for r in rows[1:]]
start = [i for i, (k, _) in enumerate(seq) if "wave_gen" in k]
nb = int(sys.argv[2]) if len(sys.argv) > 2 else 2  # bands per frame
i0, i1 = start[-1 - nb], start[-1]
tot = OrderedDict()
for k, v in seq[i0:i1]:
    tot[k] = tot.get(k, 0.0) + v
    if "-v" in sys.argv:
        print("%-40s %8.3f" % (k, v))
for k, v in tot.items():
    print("%-40s %8.3f ms" % (k, v))
print("%-40s %8.3f ms  (%d launches)" % ("frame", sum(tot.values()), i1 - i0))
