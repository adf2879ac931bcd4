"""Summarise an ncu report of decode_mega_kernel into profiles/: key raw metrics + stall samples by source line/region."""
import csv, collections, io, re, subprocess, sys
rep, out = sys.argv[1], sys.argv[2]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
h, units, v = rows[0], rows[1], rows[2]
keys = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'dram__throughput.avg.pct_of_peak_sustained_elapsed',
        'lts__t_sector_hit_rate.pct', 'sm__inst_executed.sum', 'sm__inst_executed.avg.per_cycle_elapsed', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'launch__registers_per_thread', 'launch__grid_size', 'launch__block_size', 'launch__shared_mem_per_block_dynamic', 'sm__throughput.avg.pct_of_peak_sustained_elapsed',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'smsp__cycles_active.avg', 'sm__cycles_elapsed.max']
lines = [f"# {rep}: decode_mega_kernel, one launch (llama3-8b q4_k_m, ctx ~132), ncu --set full --clock-control none", "", "## raw metrics"]
for k in keys:
    if k in h:
        i = h.index(k); lines.append(f"{k:62s} {v[i]:>16s} {units[i]}")
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
srows = list(csv.reader(io.StringIO(src)))
hdr = srows[1]; ia = hdr.index('Address'); isamp = hdr.index('# Samples'); iexec = hdr.index('Instructions Executed')
stall_cols = [i for i, x in enumerate(hdr) if x.startswith('stall_') and 'Not Issued' not in x]
tot = collections.Counter(); n = 0; ex = 0
for r in srows[2:]:
    if len(r) < len(hdr): continue
    n += int(r[isamp] or 0); ex += int(r[iexec] or 0)
    for i in stall_cols:
        x = int(r[i] or 0)
        if x: tot[hdr[i][6:]] += x
lines += ["", f"## warp stall sampling, whole kernel ({n} samples, {ex} warp-instructions executed)"]
s = sum(tot.values())
for k, x in tot.most_common(12): lines.append(f"{k:20s} {x:8d} {100 * x / s:5.1f}%")
open(out, "w").write("\n".join(lines) + "\n")
print("\n".join(lines))
