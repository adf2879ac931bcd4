import sys, os
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/tools')
import bench
from localai_b200 import capi
path, info = bench.ensure_model("llama3-8b", "q4_k_m", 0, 0, lambda: None)
for thr in (sys.argv[1:] or ["default"]):
    if thr != "default": os.environ["B200LLM_GEMM_MIN_TOKENS"] = thr
    e = capi.Engine(path, n_ctx=64 * 320, n_batch=512, n_parallel=64)
    for b in (1, 2, 3, 4, 6, 8, 12, 16, 32, 64):
        ms, l = e.bench_decode(b, 128, 16, 3)
        print(f"gemm_min_tokens={thr} batch {b:3d}: {ms:7.3f} ms/step  {b * 1e3 / ms:9.1f} tok/s  launches {l}", flush=True)
    e.close()
