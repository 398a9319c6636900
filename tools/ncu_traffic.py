#!/usr/bin/env python3
"""Per-kernel DRAM bytes and durations from an `ncu --set full` report -> JSON (profiles/r1_traffic.json).
Usage: ncu_traffic.py report.ncu-rep "how the report was made" > out.json"""
import collections
import csv
import io
import json
import re
import subprocess
import sys

rep, how = sys.argv[1], sys.argv[2] if len(sys.argv) > 2 else ""
txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv", "--metrics",
                      "gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,smsp__inst_executed.sum,"
                      "smsp__issue_active.avg.pct_of_peak_sustained_active,sm__warps_active.avg.pct_of_peak_sustained_active"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
hdr, units = rows[0], rows[1]
col = {n: hdr.index(n) for n in hdr}
scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "us": 1, "ms": 1e3, "ns": 1e-3, "s": 1e6}
def val(r, name):
    u = units[col[name]]
    return float(r[col[name]].replace(",", "")) * scale.get(u, 1)
agg = collections.OrderedDict()
for r in rows[2:]:
    k = re.sub(r"[(<].*", "", r[col["Kernel Name"]]).replace("void ", "").strip()
    a = agg.setdefault(k, {"launches": 0, "dram_bytes": 0.0, "us": 0.0, "warp_instructions": 0.0, "issue_active_pct": [], "warps_active_pct": []})
    a["launches"] += 1
    a["dram_bytes"] += val(r, "dram__bytes_read.sum") + val(r, "dram__bytes_write.sum")
    a["us"] += val(r, "gpu__time_duration.sum")
    a["warp_instructions"] += val(r, "smsp__inst_executed.sum")
    a["issue_active_pct"].append(val(r, "smsp__issue_active.avg.pct_of_peak_sustained_active"))
    a["warps_active_pct"].append(val(r, "sm__warps_active.avg.pct_of_peak_sustained_active"))
for a in agg.values():
    a["us"] = round(a["us"], 3)
    a["issue_active_pct"] = round(sum(a["issue_active_pct"]) / len(a["issue_active_pct"]), 1)
    a["warps_active_pct"] = round(sum(a["warps_active_pct"]) / len(a["warps_active_pct"]), 1)
print(json.dumps({"source": how,
                  "note": "dram_bytes = dram__bytes_read.sum + dram__bytes_write.sum summed over the kernel's launches in the capture; "
                          "us = gpu__time_duration.sum under ncu (serialised, cold cache)",
                  "kernels": agg}, indent=1))
