import sys, os
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/tools')
import bench
from localai_b200 import capi
path, info = bench.ensure_model("llama3-8b", "q4_k_m", 0, 0, lambda: None)
e = capi.Engine(path, n_ctx=1024, n_batch=512, n_parallel=1)
ms, l = e.bench_decode(1, 128, 6, 3)
print("ms/step", ms)
