"""Aggregate an ncu source page (warp-stall sampling) of decode_mega_kernel by code region and source line."""
import re, csv, collections, subprocess, sys, io, os
rep, out = sys.argv[1], sys.argv[2]
src_path = '/root/repo/localai_b200/csrc/kernels_mega.cu'
subprocess.run("cd /root/repo/localai_b200/csrc && nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -c kernels_mega.cu -o /tmp/mega.o && cd /tmp && rm -f *.cubin && cuobjdump -xelf all mega.o >/dev/null && nvdisasm --print-line-info kernels_mega.sm_100a.cubin > /tmp/mega_lines.txt 2>/dev/null", shell=True, check=True)
lines = open('/tmp/mega_lines.txt').read().split('\n')
start = [i for i, l in enumerate(lines) if l.startswith('.text._ZN2b218decode_mega_kernelILb0E')][0]
amap = {}; ln = None
for l in lines[start + 1:]:
    if l.startswith('.text.'): break
    m = re.search(r'//## File ".*", line (\d+)', l)
    if m: ln = int(m.group(1)); continue
    m = re.match(r'\s+/\*([0-9a-f]+)\*/\s+(.*?);', l)
    if m: amap[int(m.group(1), 16)] = ln
page = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(page)))
hdr = rows[1]; ia = hdr.index('Address'); isamp = hdr.index('# Samples'); iexec = hdr.index('Instructions Executed')
stall_cols = [i for i, h in enumerate(hdr) if h.startswith('stall_') and 'Not Issued' not in h]
src = open(src_path).read().split('\n')
def find(s): return [i + 1 for i, l in enumerate(src) if s in l][0]
marks = sorted([(1, "prims"), (find("struct MgLane {"), "dots"), (find("float mg_row_sum"), "reduce"),
    (find("void mg_attention_ll("), "attention"),
    (find("void mg_prefetch_kv("), "kv-prefetch"), (find("void mg_quant_block_ll("), "dist-quant"), (find("void mg_copy_xq("), "xq copy"), (find("void mg_prologue("), "prologue fn"),
    (find("decode_mega_kernel(const MegaArgs a)"), "kernel init + boundary copy"), (find("for (int p = 0; p < a.n_phases; p++) {"), "phase head"),
    (find("// ---- this lane's slice of the quantised row"), "slice"), (find("// ---- this warp's tiles of the phase"), "tile loop"),
    (find("// ---- epilogue: combine the K-split partials"), "epilogue"), (find("// ---- phase boundary"), "boundary")])
def region(l):
    r = "prims"
    for a, n in marks:
        if l >= a: r = n
    return r
base = None; per = collections.Counter(); st = collections.defaultdict(collections.Counter)
exr = collections.Counter(); smr = collections.Counter(); strg = collections.defaultdict(collections.Counter); tot = tote = 0
for r in rows[2:]:
    if len(r) < len(hdr): continue
    a = int(r[ia], 16) if r[ia].startswith('0x') else int(r[ia])
    if base is None: base = a
    l = amap.get(a - base, -1)
    s = int(r[isamp] or 0); e = int(r[iexec] or 0); tot += s; tote += e
    rg = region(l); exr[rg] += e; smr[rg] += s; per[l] += s
    for i in stall_cols:
        v = int(r[i] or 0)
        if v: st[l][hdr[i][6:]] += v; strg[rg][hdr[i][6:]] += v
o = [f"# {os.path.basename(rep)}: ncu --set full source-level warp-stall sampling of decode_mega_kernel (llama3-8b q4_k_m, b=1)",
     f"## by code region: executed warp-instructions ({tote} total), stall samples ({tot} total), top stall reasons"]
for k, v in smr.most_common():
    top = ", ".join(f"{a}={b}" for a, b in strg[k].most_common(4))
    o.append(f"{k:28s} exec {100 * exr[k] / tote:5.1f}%  samples {100 * v / tot:5.1f}%  [{top}]")
o += ["", "## hottest source lines"]
for l, s in per.most_common(24):
    top = ", ".join(f"{a}={b}" for a, b in st[l].most_common(3))
    o.append(f"{100 * s / tot:5.1f}%  L{l:4d} [{top}] {src[l - 1].strip()[:95] if l > 0 else ''}")
open(out, 'w').write("\n".join(o) + "\n"); print("\n".join(o))
