#!/usr/bin/env python3
"""Summarise an .ncu-rep: per-kernel headline metrics and (optionally) per-opcode stall samples.
usage: tools/ncu_summary.py report.ncu-rep [kernel_name_for_opcode_breakdown ...]"""
import collections, csv, subprocess, sys

rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr, units = rows[0], rows[1]
idx = {h: i for i, h in enumerate(hdr)}
KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "launch__grid_size", "launch__registers_per_thread",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum", "smsp__thread_inst_executed_per_inst_executed.ratio",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
        "l1tex__throughput.avg.pct_of_peak_sustained_active", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum"]
seen = collections.OrderedDict()
for r in rows[2:]:
    name = r[idx["Kernel Name"]].split("(")[0]
    seen.setdefault(name, []).append(r)
for name, rs in seen.items():
    r = rs[len(rs) // 2]
    print(f"==== {name}  ({len(rs)} captured)")
    for k in KEYS:
        if k in idx:
            print(f"   {k:78s} {r[idx[k]]:>16s} {units[idx[k]]}")
    st = [(float(r[i].replace(",", "")), h) for h, i in idx.items() if "issue_stalled" in h and "per_issue_active" in h and r[i]]
    print("   stalls/issue: " + ", ".join(f"{h.split('issue_stalled_')[1].split('_per_')[0]}={v:.2f}" for v, h in sorted(st, reverse=True)[:6]))
for kn in sys.argv[2:]:
    src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", kn], capture_output=True, text=True).stdout
    rows = list(csv.reader(src.splitlines()))
    h = None
    for i, r in enumerate(rows):
        if "Source" in r and "# Samples" in r:
            h, start = r, i + 1
            break
    if h is None:
        continue
    si, ns, ie = h.index("Source"), h.index("# Samples"), h.index("Instructions Executed")
    inst, cur, prev = [], [], -1
    for r in rows[start:]:
        if len(r) < len(h):
            continue
        try:
            addr = int(r[0], 16)
        except ValueError:
            continue
        if addr < prev and cur:
            inst.append(cur); cur = []
        prev = addr; cur.append(r)
    if cur:
        inst.append(cur)
    k = inst[0]
    tot = sum(int(r[ns]) for r in k) or 1
    toti = sum(int(r[ie]) for r in k) or 1
    by = collections.defaultdict(lambda: [0, 0])
    for r in k:
        t = r[si].split()
        op = (t[1] if t[0].startswith("@") else t[0]).split(".")[0]
        by[op][0] += int(r[ns]); by[op][1] += int(r[ie])
    print(f"---- {kn}: opcode share of stall samples / of executed instructions ({toti} warp instr)")
    for op, (s, i) in sorted(by.items(), key=lambda kv: -kv[1][0])[:14]:
        print(f"   {op:10s} samples {100*s/tot:5.1f}%   instr {100*i/toti:5.1f}%")
    print("   hottest lines:")
    for r in sorted(k, key=lambda r: -int(r[ns]))[:10]:
        print(f"   {100*int(r[ns])/tot:5.1f}%  x{r[ie]:>9s}  {r[si][:80]}")
