import sys, os, json
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/tools')
import bench
from localai_b200 import capi
arch = sys.argv[1] if len(sys.argv) > 1 else "llama3-8b"
preset = sys.argv[2] if len(sys.argv) > 2 else "q4_k_m"
steps = int(sys.argv[3]) if len(sys.argv) > 3 else 4
path, info = bench.ensure_model(arch, preset, 0, 0, lambda: None)
e = capi.Engine(path, n_ctx=1024, n_batch=512, n_parallel=1)
ms, l = e.bench_decode(1, 128, steps, 3)
print("ms/step", ms, "launches", l)
