import sys, os
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/tools')
import bench
from localai_b200 import capi
os.environ["B200LLM_NO_GRAPH"] = "1"
path, info = bench.ensure_model("llama3-8b", "q4_k_m", 0, 0, lambda: None)
e = capi.Engine(path, n_ctx=64 * 512, n_batch=512, n_parallel=64)
ms, l = e.bench_decode(64, 128, 2, 1)
print("ms/step", ms, "launches", l)
