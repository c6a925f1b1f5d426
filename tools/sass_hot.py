#!/usr/bin/env python
"""Condense `ncu --page source --csv --print-source sass` output: runs of SASS instructions with the same execution
count (~ basic blocks), their share of executed instructions and of stall samples, average active threads, op mix.
usage: ncu -i x.ncu-rep --page source --csv --print-source sass --kernel-name regex:K --launch-count 1 | tools/sass_hot.py [min_share]"""
import csv
import sys

thr = float(sys.argv[1]) if len(sys.argv) > 1 else 0.004
rows = list(csv.reader(sys.stdin))
h = next(r for r in rows if "Address" in r and "Source" in r)
ia, isrc, ie, it, iss = (h.index(k) for k in ("Address", "Source", "Instructions Executed", "Thread Instructions Executed", "# Samples"))
data = []
for r in rows:
    if len(r) > it and r[ie].isdigit():
        data.append((r[ia], r[isrc].strip(), int(r[ie]), int(r[it]), int(r[iss])))
    elif r and r[0] == "Kernel Name" or (len(r) > 1 and r[0] == "Function Name"):
        data.append(("--", "FUNC " + " ".join(r[:2])[:100], -1, 0, 0))
tot = sum(d[2] for d in data if d[2] > 0)
tott = sum(d[3] for d in data)
tots = sum(d[4] for d in data)
print("warp insts %d  thread insts %d (avg %.1f thr/inst)  samples %d  sass lines %d" % (tot, tott, tott / max(tot, 1), tots, len(data)))
start = 0
for i in range(1, len(data) + 1):
    if i == len(data) or data[i][2] != data[start][2]:
        s, e = start, i
        start = i
        n, ex = e - s, data[s][2]
        if ex < 0:
            print(data[s][1])
            continue
        if ex * n <= thr * tot:
            continue
        th = sum(d[3] for d in data[s:e])
        sm = sum(d[4] for d in data[s:e])
        ops = {}
        for d in data[s:e]:
            tok = d[1].split()
            op = (tok[1] if tok[0].startswith("@") and len(tok) > 1 else tok[0]).split(".")[0]
            ops[op] = ops.get(op, 0) + 1
        top = " ".join("%s:%d" % kv for kv in sorted(ops.items(), key=lambda kv: -kv[1])[:9])
        print("sass %4d-%4d n=%3d exec=%10d  %4.1f%% insts  %4.1f%% samples  thr %4.1f  %s" % (s, e, n, ex, 100.0 * ex * n / tot, 100.0 * sm / max(tots, 1), th / max(ex * n, 1), top))
