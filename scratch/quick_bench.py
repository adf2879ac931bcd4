import sys, os
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/tools')
import bench
from localai_b200 import capi
path, info = bench.ensure_model("llama3-8b", "q4_k_m", 0, 0, lambda: None)
e = capi.Engine(path, n_ctx=1024, n_batch=512, n_parallel=1)
ms, l = e.bench_decode(1, 128, 64, 4)
print(os.environ.get("B200LLM_NO_MEGA"), "ms/step", ms, "launches", l, "tok/s", 1e3/ms, "frac of 8TB/s", 4.634e9/(ms*1e-3)/8e12)
for which,name in [(4,"gate Q4_K 14336x4096"),(3,"wo Q4_K 4096x4096"),(7,"lm_head Q6_K")]:
    k_ms, k_bytes, k_type = e.bench_gemv(which, 1, 8)
    print(name, "old gemv kernel: us", k_ms*1e3, "GB/s", k_bytes/(k_ms*1e-3)/1e9)
