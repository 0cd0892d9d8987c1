#!/usr/bin/env python3
"""Compact table of an `ncu --csv --metrics ...` log: one line per launch.  usage: python tools/ncu_table.py file.csv"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr = None; data = {}
for r in rows:
    if len(r) > 10 and r[0] == 'ID': hdr = r; continue
    if hdr and len(r) == len(hdr):
        d = dict(zip(hdr, r))
        data.setdefault((int(d['ID']), d['Kernel Name'].split('(')[0], d['Grid Size']), {})[d['Metric Name']] = d['Metric Value']
tot = {}
for (i, name, grid), v in sorted(data.items()):
    us = float(v.get('gpu__time_duration.sum', 0)) / 1e3
    ins = float(v.get('smsp__inst_executed.sum', 0)) / 1e6
    rd = float(v.get('dram__bytes_read.sum', 0)) / 1e6; wr = float(v.get('dram__bytes_write.sum', 0)) / 1e6
    print(f"{i:3d} {name:24s} {grid:18s} {us:8.1f} us {ins:7.1f} Minst  rd {rd:7.1f} MB wr {wr:7.1f} MB  warps {v.get('sm__warps_active.avg.pct_of_peak_sustained_active','')} issue {v.get('smsp__issue_active.avg.pct_of_peak_sustained_active','')}")
    t = tot.setdefault(name, [0, 0, 0, 0]); t[0] += us; t[1] += ins; t[2] += rd; t[3] += wr
for k, t in tot.items(): print(f"TOTAL {k:24s} {t[0]:8.1f} us {t[1]:7.1f} Minst rd {t[2]:7.1f} wr {t[3]:7.1f}")
