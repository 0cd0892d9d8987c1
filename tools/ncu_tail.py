import csv,sys,collections
lines=[l for l in open(sys.argv[1]) if not l.startswith('==')]
rd=csv.DictReader(lines)
rows=[]
for r in rd:
    if r['Metric Name']=='gpu__time_duration.sum':
        v=float(r['Metric Value'].replace(',','')); u=r['Metric Unit']; v*={'ns':1e-3,'us':1,'ms':1e3}.get(u,1)
        rows.append((r['Kernel Name'].split('(')[0].split('::')[-1],v, r.get('Grid Size',''), r.get('Block Size','')))
# print last N launches
N=int(sys.argv[2]) if len(sys.argv)>2 else 40
for k,v,g,b in rows[-N:]: print(f"{k:28s} {v:9.1f} us  grid {g} block {b}")
