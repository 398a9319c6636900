#!/usr/bin/env python3
"""Aggregate `ncu --page source --csv --print-source cuda,sass` by CUDA source line: share of
executed warp instructions and of stall samples, top stall reason.
Usage: ncu -i rep --page source --csv --print-source cuda,sass --kernel-name K | ncu_line_summary.py [min_pct]"""
import csv
import collections
import sys

minp = float(sys.argv[1]) if len(sys.argv) > 1 else 0.5
rows = list(csv.reader(sys.stdin))
hdr = next(r for r in rows if r and r[0] == "Line No" and len(r) > 10)
iL, iSrc, iAddr, iI, iS = 0, 1, 2, hdr.index("Instructions Executed"), hdr.index("# Samples")
stall = [(k, i) for i, k in enumerate(hdr) if k.startswith("stall_") and "Not Issued" not in k]
agg = collections.OrderedDict()
fname = ""
for r in rows:
    if r and r[0] == "File Path":
        fname = r[1].rsplit("/", 1)[-1]
    if len(r) < iS + 1 or not r[0].isdigit():        # per-line rows carry the line's totals
        continue
    # badly escaped source text may add fields: metrics are located relative to the "-","-" pair
    try:
        o = next(i for i in range(2, len(r) - 1) if r[i] == "-" and r[i + 1] == "-") - 2
        n, sm = int(r[iI + o]), int(r[iS + o])
    except (StopIteration, ValueError, IndexError):
        continue
    key = (fname, int(r[0]))
    a = agg.setdefault(key, {"src": ",".join(r[1:2 + o]).strip(), "n": 0, "s": 0, "st": collections.Counter()})
    a["n"] += n; a["s"] += sm
    for k, i in stall:
        if i + o < len(r) and r[i + o].isdigit():
            a["st"][k] += int(r[i + o])
tn = sum(a["n"] for a in agg.values()); ts = sum(a["s"] for a in agg.values())
print(f"warp instr {tn}  samples {ts}")
for (f, l), a in sorted(agg.items(), key=lambda kv: (kv[0][0], kv[0][1])):
    pn, ps = 100 * a["n"] / max(tn, 1), 100 * a["s"] / max(ts, 1)
    if pn >= minp or ps >= minp:
        top = ",".join(f"{k[6:]}:{v}" for k, v in a["st"].most_common(2))
        print(f"{f}:{l:5d} {pn:5.1f}%i {ps:5.1f}%s  {top:28s} {a['src'][:100]}")
