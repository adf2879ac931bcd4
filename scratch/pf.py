import sys; sys.path.insert(0,'/root/repo'); sys.path.insert(0,'/root/repo/tools')
import bench
from localai_b200 import capi
path, info = bench.ensure_model("llama3-8b", "q4_k_m", 0, 0, lambda: None)
e = capi.Engine(path, n_ctx=4096, n_batch=2048, n_parallel=1)
for i in range(2): print("prefill 2048 ms", e.bench_prefill(2048, 3, 1))
