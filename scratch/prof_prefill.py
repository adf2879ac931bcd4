import sys, os
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/tools')
import bench
from localai_b200 import capi
path, info = bench.ensure_model("llama3-8b", "q4_k_m", 0, 0, lambda: None)
e = capi.Engine(path, n_ctx=2304, n_batch=2048, n_parallel=1)
it = int(sys.argv[1]) if len(sys.argv) > 1 else 2
ms, l = e.bench_prefill(2048, it, 1)
print("prefill ms", ms, "launches", l, "tok/s", 2048e3/ms)
