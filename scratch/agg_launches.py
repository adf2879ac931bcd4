import csv, collections, sys
f = sys.argv[1]; nlast = int(sys.argv[2]) if len(sys.argv) > 2 else 330
lines=[l for l in open(f) if not l.startswith('==')]
rows=[]
for x in csv.DictReader(lines):
    try: rows.append((int(x['ID']), x['Kernel Name'], float(x['Metric Value'].replace(',','')), x['Grid Size']))
    except Exception as e: pass
last=rows[-nlast:]
agg=collections.defaultdict(lambda:[0,0.0])
for i,n,v,g in last:
    k=n.split('(')[0]; agg[k][0]+=1; agg[k][1]+=v
tot=sum(v for _,v in agg.values())
for k,(c,v) in sorted(agg.items(), key=lambda kv:-kv[1][1]): print(f"{k:50s} {c:4d} {v/1e3:9.1f} us  {v/tot*100:5.1f}%  avg {v/c/1e3:7.2f} us")
print("total us", tot/1e3)
for i,n,v,g in last[20:32]: print(i,n[:40],v/1e3,g)
