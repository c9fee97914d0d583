#!/usr/bin/env python
"""Top stall locations of an ncu report's source page: python scripts/ncu_hot.py rep.ncu-rep [N]"""
import csv, subprocess, sys
rep = sys.argv[1]; N = int(sys.argv[2]) if len(sys.argv) > 2 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hi]; ci = {h: i for i, h in enumerate(hdr)}
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
data = []
for k, r in enumerate(rows[hi + 1:]):
    if len(r) < len(hdr): continue
    data.append((float(r[ci["# Samples"]] or 0), k, r))
tot = sum(d[0] for d in data)
print("total samples", tot, "instructions", len(data))
for v, k, r in sorted(data, key=lambda x: -x[0])[:N]:
    top = sorted(((float(r[ci[s]] or 0), s) for s in stalls), reverse=True)[:2]
    print(f"{int(v):6d} {100*v/tot:5.1f}%  #{k:5d} exec={r[ci['Instructions Executed']]:>8s}  {r[ci['Source']].strip()[:70]:70s} {top[0][1]}={int(top[0][0])} {top[1][1]}={int(top[1][0])}")
