import sys, os
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/tools')
import bench
from localai_b200 import capi
path, info = bench.ensure_model("llama3-8b", "q4_k_m", 0, 0, lambda: None)
e = capi.Engine(path, n_ctx=64 * 512, n_batch=2048, n_parallel=64)
nt = int(os.environ.get("NB", "64"))
tot = 0
for which, name in [(8, "wq|wk|wv"), (3, "wo"), (9, "gate|up"), (6, "down"), (7, "lm_head")]:
    ms, b, t = e.bench_gemv(which, nt, 4)
    print(f"{name:10s} n_tok {nt}: {ms*1e3:8.2f} us  {b/1e6:7.1f} MB  {b/(ms*1e-3)/1e12:5.2f} TB/s")
    tot += ms * (1 if which == 7 else 32)
print("sum over a step (32 layers + lm_head): %.3f ms" % tot)
