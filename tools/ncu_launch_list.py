#!/usr/bin/env python3
"""Turn the CSV log of
  ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,dram__bytes_read.sum,dram__bytes_write.sum
      --clock-control none --csv --log-file raw.csv python bench.py --steps 1 --warmup 1 --skip-cpu
into the per-launch list and the per-kernel shares committed under profiles/ (one extraction step = the launches from one
k_pyr_level0 to the next).  usage: tools/ncu_launch_list.py raw.csv out_prefix "comment" """
import collections
import csv
import sys

raw, prefix, comment = sys.argv[1], sys.argv[2], (sys.argv[3] if len(sys.argv) > 3 else "")
rows = []
with open(raw, newline="") as f:
    lines = [ln for ln in f if not ln.startswith("==")]
rd = csv.DictReader(lines)
launch = collections.OrderedDict()
for r in rd:
    i = int(r["ID"])
    d = launch.setdefault(i, {"kernel": r["Kernel Name"].split("(")[0].split("::")[-1]})
    v = float(r["Metric Value"].replace(",", ""))
    u = r["Metric Unit"]
    m = r["Metric Name"]
    if m == "gpu__time_duration.sum":
        v *= {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}.get(u, 1.0)
    if m.startswith("dram__bytes"):
        v *= {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(u, 1.0)
    d[m] = v
L = [d for d in launch.values() if d["kernel"].startswith("k_")]   # the library's kernels (torch fills / copies of bench.py dropped)
starts = [i for i, d in enumerate(L) if d["kernel"] == "k_pyr_level0"]
# the LAST complete step (warm caches, steady state)
s0 = starts[-2] if len(starts) >= 2 and starts[-1] + 26 > len(L) else starts[-1]
s1 = next((s for s in starts if s > s0), len(L))
step = L[s0:s1]
with open(prefix + "_launches.csv", "w") as f:
    f.write(f"# {comment}\n")
    f.write("kernel,time_us,warp_instructions,issue_active_pct,dram_read_bytes,dram_write_bytes\n")
    for d in step:
        f.write(f"{d['kernel']},{d['gpu__time_duration.sum']:.1f},{int(d['smsp__inst_executed.sum'])},"
                f"{d['smsp__issue_active.avg.pct_of_peak_sustained_active']:.1f},{int(d['dram__bytes_read.sum'])},{int(d['dram__bytes_write.sum'])}\n")
agg = collections.OrderedDict()
for d in step:
    a = agg.setdefault(d["kernel"], [0, 0.0, 0, 0.0, 0.0])
    a[0] += 1; a[1] += d["gpu__time_duration.sum"]; a[2] += int(d["smsp__inst_executed.sum"])
    a[3] += d["dram__bytes_read.sum"] / 1e6; a[4] += d["dram__bytes_write.sum"] / 1e6
tot = sum(a[1] for a in agg.values())
with open(prefix + "_launch_shares.csv", "w") as f:
    f.write(f"# per-kernel totals of {prefix.split('/')[-1]}_launches.csv ({comment})\n")
    f.write("kernel,launches,total_us,share,warp_instructions,dram_read_MB,dram_write_MB\n")
    for k, a in agg.items():
        f.write(f"{k},{a[0]},{a[1]:.1f},{a[1] / tot:.4f},{a[2]},{a[3]:.1f},{a[4]:.1f}\n")
    f.write(f"TOTAL,{len(step)},{tot:.1f},1.0,,,\n")
print(open(prefix + "_launch_shares.csv").read())
