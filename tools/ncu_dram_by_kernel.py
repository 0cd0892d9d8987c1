import csv,sys,collections
lines=[l for l in open(sys.argv[1]) if not l.startswith("==")]
L=collections.OrderedDict()
for r in csv.DictReader(lines):
    d=L.setdefault(int(r["ID"]),{"k":r["Kernel Name"].split("(")[0].split("::")[-1]})
    v=float(r["Metric Value"].replace(",","")); u=r["Metric Unit"]; m=r["Metric Name"]
    if m.startswith("dram"): v*={"byte":1,"Kbyte":1e3,"Mbyte":1e6,"Gbyte":1e9}.get(u,1)
    if m.startswith("gpu__time"): v*={"ns":1e-3,"us":1,"ms":1e3}.get(u,1)
    d[m]=v
L=[d for d in L.values() if d["k"].startswith("k_")]
st=[i for i,d in enumerate(L) if d["k"].startswith("k_stereo_filter")]
step=L[st[-2]+1:st[-1]+1]
agg=collections.OrderedDict()
for d in step:
    a=agg.setdefault(d["k"],[0,0,0,0]); a[0]+=1; a[1]+=d["gpu__time_duration.sum"]; a[2]+=d["dram__bytes_read.sum"]/1e6; a[3]+=d["dram__bytes_write.sum"]/1e6
print("kernel,launches,total_us,dram_read_MB,dram_write_MB")
for k,a in agg.items(): print(f"{k},{a[0]},{a[1]:.1f},{a[2]:.1f},{a[3]:.1f}")
print("TOTAL,%d,%.1f,%.1f,%.1f"%(len(step),sum(a[1] for a in agg.values()),sum(a[2] for a in agg.values()),sum(a[3] for a in agg.values())))
