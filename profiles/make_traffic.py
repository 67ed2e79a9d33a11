#!/usr/bin/env python
"""DRAM traffic per launch (dram__bytes_read.sum + dram__bytes_write.sum) of every
kernel in one or more `ncu --set full` reports -> profiles/traffic.json, which
bench.py quotes as roofline.traffic.  Kernel names are reduced to the short names
bench.py uses (k_hessf_cols1 -> k_hessf_cols, k_hessf_rows_inv_tma -> k_hessf_rows_inv).
usage: make_traffic.py tag report.ncu-rep [report2.ncu-rep ...]"""
import csv
import json
import subprocess
import sys

UNIT = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}
tag, reps = sys.argv[1], sys.argv[2:]
SHORT = {"k_hessf_cols1": "k_hessf_cols", "k_hessf_rows_inv_tma": "k_hessf_rows_inv",
         "k_hessf_rows_inv2": "k_hessf_rows_inv"}
acc = {}
for rep in reps:
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    h, u = rows[0], rows[1]
    ki, ri, wi, ti = (h.index(n) for n in ("Kernel Name", "dram__bytes_read.sum", "dram__bytes_write.sum",
                                           "gpu__time_duration.sum"))
    for r in rows[2:]:
        name = r[ki].split("<")[0].split("(")[0].replace("void ", "").strip()
        name = SHORT.get(name, name)
        b = float(r[ri]) * UNIT[u[ri]] + float(r[wi]) * UNIT[u[wi]]
        t = float(r[ti]) * {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(u[ti], 1e-6)
        a = acc.setdefault(name, {"launches": 0, "dram_bytes": 0.0, "ms": 0.0})
        a["launches"] += 1
        a["dram_bytes"] += b
        a["ms"] += t
res = {"source": f"ncu --set full --clock-control none, {tag} (bench.py --steps 3 --warmup 3 --ncu-step: one PD step, "
                 "config 3 fp32, 1 x B200)", "kernels": {}}
for k, a in acc.items():
    res["kernels"][k] = {"launches_in_step": a["launches"],
                         "dram_bytes_per_launch": round(a["dram_bytes"] / a["launches"]),
                         "dram_bytes_per_step": round(a["dram_bytes"]),
                         "ncu_ms_per_step": round(a["ms"], 4)}
json.dump(res, open("profiles/traffic.json", "w"), indent=1)
print(json.dumps(res, indent=1))
