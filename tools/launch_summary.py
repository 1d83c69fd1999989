"""Per-kernel totals of an ncu launch list (`ncu --metrics gpu__time_duration.sum --csv --log-file x.csv ...`):
average duration x launches = total, share of the listed time.  usage: python tools/launch_summary.py x.csv"""
import csv,sys,collections
rows=[r for r in csv.reader(open(sys.argv[1])) if len(r)>10]
hdr=rows[0]; ki=hdr.index('Kernel Name'); vi=hdr.index('Metric Value'); ui=hdr.index('Metric Unit')
t=collections.OrderedDict(); n=collections.Counter()
for r in rows[1:]:
    v=float(r[vi].replace(',','')); u=r[ui]
    v*= {'ns':1e-3,'us':1,'ms':1e3,'usecond':1,'nsecond':1e-3,'msecond':1e3}.get(u,1)
    k=r[ki].split('(')[0][:60]; t[k]=t.get(k,0)+v; n[k]+=1
tot=sum(t.values())
for k,v in sorted(t.items(), key=lambda x:-x[1]): print(f"{v/n[k]:9.1f} us x {n[k]:4d} = {v:10.1f} us {100*v/tot:5.1f}%  {k}")
