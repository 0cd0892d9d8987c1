#!/usr/bin/env python3
"""Per-CUDA-source-line share of executed warp instructions and stall samples from an .ncu-rep captured with
--import-source on (kernels compiled with -lineinfo).  usage: tools/ncu_lines.py report.ncu-rep [kernel substring] [top N]"""
import csv
import subprocess
import sys

rep = sys.argv[1]
want = sys.argv[2] if len(sys.argv) > 2 else ""
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
launches, cur, hdr = [], None, None
for r in csv.reader(out.splitlines()):
    if r and r[0] == "Function Name":
        cur = {"name": r[1], "lines": []}
        launches.append(cur)
    elif r and r[0] == "Line No":
        hdr = r
    elif cur is not None and hdr is not None and r and r[0].isdigit() and len(r) == len(hdr):
        try:   # source lines holding quotes (inline asm) can break the CSV row: skip those
            int(r[hdr.index("# Samples")]), int(r[hdr.index("Instructions Executed")]), int(r[hdr.index("Thread Instructions Executed")])
        except ValueError:
            continue
        cur["lines"].append(r)
iS, iI, iT = hdr.index("# Samples"), hdr.index("Instructions Executed"), hdr.index("Thread Instructions Executed")
for L in launches:
    if want not in L["name"]:
        continue
    rows = L["lines"]
    tS = sum(int(r[iS]) for r in rows) or 1
    tI = sum(int(r[iI]) for r in rows) or 1
    tT = sum(int(r[iT]) for r in rows)
    print(f"==== {L['name'][:90]}: {tI} warp instr, {tT / tI:.1f} active lanes, {tS} samples")
    for r in sorted(rows, key=lambda r: -int(r[iI]))[:top]:
        print(f"{r[0]:>5} I {100 * int(r[iI]) / tI:5.1f}%  S {100 * int(r[iS]) / tS:5.1f}%  act {int(r[iT]) / max(int(r[iI]), 1):5.1f}  {r[1].strip()[:120]}")
    if want:
        break
