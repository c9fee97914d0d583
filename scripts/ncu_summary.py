#!/usr/bin/env python
"""Turns an .ncu-rep (brought back from the GPU box in gpurun_out/) into a small JSON summary for
profiles/ncu/: headline metrics of the first captured launch, warp-stall sample distribution and the
executed-instruction mix. Usage: ncu_summary.py report.ncu-rep "what was run" > profiles/ncu/x.json"""
import collections
import csv
import io
import json
import subprocess
import sys

rep, note = sys.argv[1], (sys.argv[2] if len(sys.argv) > 2 else "")
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, first = rows[0], rows[1], rows[2]
want = ["Kernel Name", "gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
        "launch__shared_mem_per_block_dynamic", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__cycles_elapsed.max",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "lts__t_sector_hit_rate.pct", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active"]
metrics = {k: (first[hdr.index(k)] + (" " + units[hdr.index(k)] if units[hdr.index(k)] else "")) for k in want if k in hdr}
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
hi = [i for i, r in enumerate(rows) if r and r[0] == "Address"]
stalls, ops, total = collections.Counter(), collections.Counter(), 0
if hi:
    h = rows[hi[0]]
    end = hi[1] - 1 if len(hi) > 1 else len(rows)
    ia, isrc = h.index("Instructions Executed"), h.index("Source")
    sc = [i for i, x in enumerate(h) if x.startswith("stall_") and "Not Issued" not in x]
    for r in rows[hi[0] + 1:end]:
        if len(r) <= ia or not r[ia].isdigit():
            continue
        c = int(r[ia])
        t = r[isrc].split()
        op = (t[1] if t[0].startswith("@") else t[0]).split(".")
        ops[".".join(op[:2]) if op[0] in ("IDP", "FFMA", "UTMALDG", "UTMASTG", "SYNCS", "LDS", "STS", "LDG", "STG") else op[0]] += c
        total += c
        for i in sc:
            if r[i].isdigit():
                stalls[h[i][6:]] += int(r[i])
ts = sum(stalls.values()) or 1
print(json.dumps({"report": rep.split("/")[-1], "what": note, "metrics": metrics,
                  "warp_stall_samples_pct": {k: round(100 * v / ts, 1) for k, v in stalls.most_common(10)},
                  "executed_warp_instructions": total,
                  "instruction_mix_pct": {k: round(100 * v / max(total, 1), 1) for k, v in ops.most_common(16)}}, indent=1))
