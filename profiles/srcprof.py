#!/usr/bin/env python3
"""Summarise an `ncu --set full --import-source on` capture per SASS instruction: samples, executed count, top stalls.
usage: srcprof.py file.ncu-rep [min_fraction] [kernel-name regex]"""
import csv, subprocess, sys
rep = sys.argv[1]; frac = float(sys.argv[2]) if len(sys.argv) > 2 else 0.004
sel = ["--kernel-name", "regex:" + sys.argv[3]] if len(sys.argv) > 3 else []
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"] + sel, capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
h = rows[hi]; data = [r for r in rows[hi + 1:] if len(r) == len(h) and r[0] != "Address"]
ia = h.index("Address"); isrc = h.index("Source"); isamp = h.index("# Samples"); iex = h.index("Instructions Executed"); ith = h.index("Avg. Threads Executed")
stall = [i for i, n in enumerate(h) if n.startswith("stall_") and "Not Issued" not in n]
tot = sum(int(r[isamp]) for r in data)
print("total samples", tot, "instructions executed", sum(int(r[iex]) for r in data))
base = int(data[0][ia], 16)
for r in data:
    s = int(r[isamp])
    if s > tot * frac:
        st = sorted(((int(r[i]), h[i][6:]) for i in stall), reverse=True)[:2]
        print(f"{int(r[ia],16)-base:#06x} {r[isrc].strip()[:58]:58s} {100*s/tot:5.1f}% ex={r[iex]:>9s} thr={r[ith]:>2s} {st}")
