import sys, os
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/tools')
import bench
from localai_b200 import capi
path, info = bench.ensure_model("llama3-8b", "q4_k_m", 0, 0, lambda: None)
nb = int(os.environ.get("NB", "64")); steps = int(os.environ.get("STEPS", "8"))
e = capi.Engine(path, n_ctx=64 * 512, n_batch=2048, n_parallel=64)
ms, l = e.bench_decode(nb, 128, steps, 3)
print("b", nb, "ms/step", ms, "launches", l, "tok/s", nb * 1e3 / ms)
