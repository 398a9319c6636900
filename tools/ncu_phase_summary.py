#!/usr/bin/env python3
"""Summarise `ncu --page source --csv` output of one kernel: executed warp instructions, stall
samples and average active threads between consecutive BAR.SYNC instructions (= kernel phases),
plus the opcode mix.  Usage: ncu -i rep --page source --csv --kernel-name K | ncu_phase_summary.py"""
import csv
import collections
import sys

rows = list(csv.reader(sys.stdin))
hdr = next(r for r in rows if r and r[0] == "Address")
data = [r for r in rows if r and r[0].startswith("0x")]
iS, iI, iT, iSamp = (hdr.index(k) for k in ("Source", "Instructions Executed", "Thread Instructions Executed", "# Samples"))
tot = sum(int(r[iI]) for r in data)
tots = sum(int(r[iSamp]) for r in data)
print(f"total warp instr {tot}  thread instr {sum(int(r[iT]) for r in data)}  SASS lines {len(data)}  samples {tots}")
seg = acc = accs = acct = start = 0
ops = collections.Counter()
for k, r in enumerate(data):
    n = int(r[iI])
    acc += n; accs += int(r[iSamp]); acct += int(r[iT])
    op = r[iS].split()[0] if not r[iS].strip().startswith("@") else r[iS].split()[1]
    ops[op.split(".")[0]] += n
    if "BAR.SYNC" in r[iS] or k == len(data) - 1:
        print(f"  phase {seg}: SASS {start}-{k}  instr {acc} ({100 * acc / max(tot, 1):.1f}%)  samples {accs} "
              f"({100 * accs / max(tots, 1):.1f}%)  avg threads {acct / max(acc, 1):.1f}")
        seg += 1; acc = accs = acct = 0; start = k + 1
print("  opcode mix:", ", ".join(f"{o} {100 * c / max(tot, 1):.1f}%" for o, c in ops.most_common(14)))
