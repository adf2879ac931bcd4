"""Aggregate the warp-stall samples of a decode_mega_kernel ncu capture by source line and by code region.

Uses ncu's own source correlation (--print-source cuda,sass) and the source text embedded in the report
(--import-source on), so the report can be read after the working tree has moved on.
usage: python scratch/ncu_lines.py gpurun_out/r2_mega_v9.ncu-rep profiles/r2_mega_v9_stalls_by_region.txt
"""
import collections, csv, io, os, subprocess, sys

rep, out = sys.argv[1], sys.argv[2]
page = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(page)))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Line No")
hdr = rows[hi]
isamp, iexec = hdr.index("# Samples"), hdr.index("Instructions Executed")
stall_cols = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
srcpage = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda"], capture_output=True, text=True).stdout
src = {}
for r in csv.reader(io.StringIO(srcpage)):
    if r and r[0].isdigit():
        src[int(r[0])] = ",".join(r[1:])
nlines = max(src)


def find(s):
    for i in range(1, nlines + 1):
        if s in src.get(i, ""):
            return i
    return 10 ** 9


marks = sorted([(1, "prims (mbarrier / LDS / LL words)"), (find("struct MgLane {"), "dots"), (find("float mg_row_sum"), "reduce"),
                (find("void mg_attention_ll("), "attention"), (find("void mg_prefetch_kv("), "kv-prefetch"),
                (find("void mg_quant_block_ll("), "dist-quant"), (find("void mg_copy_xq("), "xq copy"), (find("void mg_prologue("), "prologue fn"),
                (find("decode_mega_kernel(const MegaArgs a)"), "kernel init"), (find("for (int p = 0; p < a.n_phases; p++) {"), "phase head"),
                (find("// ---- this lane's slice of the quantised row"), "slice"), (find("// ---- this warp's tiles of the phase"), "tile loop"),
                (find("// ---- epilogue: combine the K-split partials"), "epilogue")])


def region(l):
    r = marks[0][1]
    for a, n in marks:
        if l >= a:
            r = n
    return r


cur = -1
per = collections.Counter(); st = collections.defaultdict(collections.Counter)
exr = collections.Counter(); smr = collections.Counter(); strg = collections.defaultdict(collections.Counter); tot = tote = 0
for r in rows[hi + 1:]:
    if r and r[0].isdigit():
        cur = int(r[0])
        continue
    if len(r) < len(hdr) or r[0] != "" or not r[2].startswith("0x"):
        continue
    s = int(r[isamp] or 0); e = int(r[iexec] or 0)
    tot += s; tote += e
    rg = region(cur); exr[rg] += e; smr[rg] += s; per[cur] += s
    for i in stall_cols:
        v = int(r[i] or 0)
        if v:
            st[cur][hdr[i][6:]] += v; strg[rg][hdr[i][6:]] += v
o = [f"# {os.path.basename(rep)}: ncu --set full source-level warp-stall sampling of decode_mega_kernel (llama3-8b q4_k_m, b=1)",
     f"## by code region: executed warp-instructions ({tote} total), stall samples ({tot} total), top stall reasons"]
for k, v in smr.most_common():
    top = ", ".join(f"{a}={b}" for a, b in strg[k].most_common(4))
    o.append(f"{k:36s} exec {100 * exr[k] / max(1, tote):5.1f}%  samples {100 * v / max(1, tot):5.1f}%  [{top}]")
o += ["", "## hottest source lines"]
for l, s in per.most_common(28):
    top = ", ".join(f"{a}={b}" for a, b in st[l].most_common(3))
    o.append(f"{100 * s / max(1, tot):5.1f}%  L{l:4d} [{top}] {src.get(l, '').strip()[:100]}")
open(out, "w").write("\n".join(o) + "\n")
print("\n".join(o))
