#!/usr/bin/env python3
"""Executed warp instructions per SASS instruction of one kernel launch from an .ncu-rep (--set full --import-source on):
prints every instruction with its share, so loop bodies can be read off.  usage: tools/ncu_sass_hot.py rep kernel_regex [launch index]"""
import csv
import subprocess
import sys

rep, kern = sys.argv[1], sys.argv[2]
which = int(sys.argv[3]) if len(sys.argv) > 3 else 0
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass", "--kernel-name", "regex:" + kern],
                     capture_output=True, text=True).stdout
launches, cur, hdr = [], None, None
for r in csv.reader(out.splitlines()):
    if r and r[0] == "Kernel Name":
        cur = []
        launches.append(cur)
    elif r and r[0] == "Address":
        hdr = r
    elif cur is not None and hdr and len(r) == len(hdr):
        cur.append(r)
rows = launches[which]
iI, iS, iT = hdr.index("Instructions Executed"), hdr.index("# Samples"), hdr.index("Thread Instructions Executed")
tI = sum(int(r[iI]) for r in rows)
tS = sum(int(r[iS]) for r in rows) or 1
print(f"{len(launches)} launches; launch {which}: {tI} warp instructions, {len(rows)} SASS instructions")
for k, r in enumerate(rows):
    n = int(r[iI])
    print(f"{k:4d} {100 * n / tI:5.2f}% S{100 * int(r[iS]) / tS:5.1f}% act{int(r[iT]) / max(n, 1):5.1f} {r[1].strip()[:90]}")
