#!/usr/bin/env python3
"""Extract the DRAM traffic of one kernel launch from an ncu --set full report into a small JSON that bench.py reads.

usage: python tools/ncu_traffic.py gpurun_out/r2_mega_v9.ncu-rep profiles/r2_mega_traffic.json [summary file it came with]
"""
import csv
import io
import json
import subprocess
import sys

rep, out = sys.argv[1], sys.argv[2]
src = sys.argv[3] if len(sys.argv) > 3 else rep
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
h, units, v = rows[0], rows[1], rows[2]
scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}


def metric(name):
    i = h.index(name)
    return float(v[i].replace(",", "")) * scale[units[i]]


d = {"kernel": v[h.index("Kernel Name")], "dram_bytes_read": metric("dram__bytes_read.sum"), "dram_bytes_write": metric("dram__bytes_write.sum"),
     "gpu_time_ms": float(v[h.index("gpu__time_duration.sum")].replace(",", "")) * {"ms": 1.0, "us": 1e-3, "ns": 1e-6, "s": 1e3}[units[h.index("gpu__time_duration.sum")]],
     "source": src, "how": "ncu --set full --clock-control none, one launch; dram__bytes_read.sum + dram__bytes_write.sum"}
json.dump(d, open(out, "w"), indent=1)
print(json.dumps(d, indent=1))
