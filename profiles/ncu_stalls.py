#!/usr/bin/env python
"""Per-instruction stall hot spots of one kernel from an ncu report (source page).
usage: ncu_stalls.py report.ncu-rep kernel-regex [instance] [threshold]"""
import csv
import subprocess
import sys

rep, kre = sys.argv[1], sys.argv[2]
inst = int(sys.argv[3]) if len(sys.argv) > 3 else 0
thr = float(sys.argv[4]) if len(sys.argv) > 4 else 0.015
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass",
                      "--kernel-name", "regex:" + kre], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr = [i for i, r in enumerate(rows) if r and r[0] == "Address"] + [len(rows)]
h = rows[hdr[inst]]
si, ie = h.index("# Samples"), h.index("Instructions Executed")
cols = {n: h.index(n) for n in h if n.startswith("stall_") and "Not Issued" not in n}


def I(x):
    try:
        return int(x)
    except ValueError:
        return 0


body = [r for r in rows[hdr[inst] + 1:hdr[inst + 1]] if len(r) > si]
tot = sum(I(r[si]) for r in body)
print("samples", tot, "instructions", len(body), "warp-instr executed", sum(I(r[ie]) for r in body))
agg = {n: sum(I(r[c]) for r in body) for n, c in cols.items()}
print("  ".join(f"{n[6:]}={v}" for n, v in sorted(agg.items(), key=lambda kv: -kv[1]) if v))
for k, r in enumerate(body):
    s = I(r[si])
    if s > thr * tot:
        top = max(cols, key=lambda n: I(r[cols[n]]))
        print(f"{k:5d} {r[1][:64]:64s} exec={I(r[ie])} samples={s} {top}")
