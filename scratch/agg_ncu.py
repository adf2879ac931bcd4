import csv, sys, collections, re
rows = list(csv.reader(open(sys.argv[1], errors='replace')))
hdr = None; agg = collections.defaultdict(lambda: [0, 0.0]); total = 0.0
for r in rows:
    if hdr is None:
        if 'Kernel Name' in r: hdr = r
        continue
    d = dict(zip(hdr, r))
    if d.get('Metric Name') != 'gpu__time_duration.sum': continue
    v = float(d['Metric Value'].replace(',', '')); u = d['Metric Unit']
    v = v / 1000.0 if u in ('ns', 'nsecond') else (v if u in ('us', 'usecond') else v * 1000.0)
    name = re.sub(r'\(.*', '', d['Kernel Name'])
    grid = d.get('Grid Size', '')
    agg[(name, grid)][0] += 1; agg[(name, grid)][1] += v; total += v
print(f"total {total:.1f} us over {sum(a[0] for a in agg.values())} launches")
for (name, grid), (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:25]:
    print(f"{t:10.1f} us {100*t/total:5.1f}%  n={n:5d}  avg {t/n:8.2f} us  grid {grid:>12}  {name[:90]}")
