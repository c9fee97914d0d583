#!/usr/bin/env python
"""gpurun_out/ref_ncu_*.csv (scripts/ref_kernels_ncu.sh) -> profiles/ref_kernels_ncu.json"""
import csv
import glob
import json
import os
import re
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
out = {"when": time.strftime("%Y-%m-%d"), "how": "ncu --metrics gpu__time_duration.sum --clock-control none on the "
       "unmodified, shim-built baseline/_ref/ref_kernel_cu (one rank, one B200)"}
for p in sorted(glob.glob(os.path.join(ROOT, "gpurun_out", "ref_ncu_*x*.csv"))):
    size = re.search(r"ref_ncu_(\d+x\d+)\.csv", p).group(1)
    lines = [ln for ln in open(p) if ln.startswith('"')]
    per = {}
    for row in csv.DictReader(lines):
        k = row["Kernel Name"].split("(")[0]
        v = float(row["Metric Value"].replace(",", ""))
        unit = row["Metric Unit"]
        if row["Metric Name"] == "gpu__time_duration.sum":
            v = v / 1e6 if unit in ("ns", "nsecond") else v / 1e3 if unit in ("us", "usecond") else v
            per.setdefault(k, {})["ms"] = round(v, 5)
        else:
            scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(unit, 1)
            per.setdefault(k, {})[row["Metric Name"].split("__")[1].replace(".sum", "")] = int(v * scale)
    out[size] = {"kernels": per, "total_ms": round(sum(k["ms"] for k in per.values()), 5)}
json.dump(out, open(os.path.join(ROOT, "profiles", "ref_kernels_ncu.json"), "w"), indent=1)
print(json.dumps(out, indent=1))
