"""Top source lines by warp-stall samples of an ncu capture (--import-source on). usage: ncu_top_lines.py rep [n]"""
import collections, csv, io, subprocess, sys
rep = sys.argv[1]; n = int(sys.argv[2]) if len(sys.argv) > 2 else 40
page = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(page)))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Line No")
hdr = rows[hi]
isamp, iexec = hdr.index("# Samples"), hdr.index("Instructions Executed")
stall_cols = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
srcpage = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda"], capture_output=True, text=True).stdout
src = {}
for r in csv.reader(io.StringIO(srcpage)):
    if r and r[0].isdigit(): src[int(r[0])] = ",".join(r[1:])[:110]
cur = -1; per = collections.Counter(); ex = collections.Counter(); st = collections.defaultdict(collections.Counter); tot = 0; tote = 0
sass = collections.Counter(); sass_txt = {}
for r in rows[hi + 1:]:
    if r and r[0].isdigit(): cur = int(r[0]); continue
    if len(r) < len(hdr) or r[0] != "" or not r[2].startswith("0x"): continue
    s = int(r[isamp] or 0); e = int(r[iexec] or 0)
    tot += s; tote += e; per[cur] += s; ex[cur] += e
    sass[(cur, r[2])] += s; sass_txt[(cur, r[2])] = r[3][:60]
    for i in stall_cols:
        v = int(r[i] or 0)
        if v: st[cur][hdr[i][6:]] += v
print("total samples", tot, "instructions", tote)
for l, s in per.most_common(n):
    top = ", ".join(f"{k}:{v}" for k, v in st[l].most_common(3))
    print(f"{s:7d} {100*s/tot:5.1f}%  exec {ex[l]:9d}  L{l}: {src.get(l,'').strip()}   [{top}]")
print("--- top SASS")
for (l, a), s in sass.most_common(25): print(f"{s:7d} L{l} {a} {sass_txt[(l,a)]}")
