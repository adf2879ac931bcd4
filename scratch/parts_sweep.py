import sys, os, itertools, subprocess
code = '''
import sys, os; sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tools")
import bench
from localai_b200 import capi
path, info = bench.ensure_model("llama3-8b", "q4_k_m", 0, 0, lambda: None)
e = capi.Engine(path, n_ctx=4096, n_batch=128, n_parallel=1)
for ctx in (128, 256, 512, 1024, 2048, 3800):
    ms, l = e.bench_decode(1, ctx, 32, 3)
    print("RESULT parts", os.environ["B200LLM_ATT_PARTS"], "ctx", ctx, "ms", round(ms, 4))
'''
for parts in ["1", "2", "4"]:
    env = dict(os.environ, B200LLM_ATT_PARTS=parts)
    r = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True, timeout=600)
    print("\n".join(l for l in r.stdout.splitlines() if l.startswith("RESULT")) or r.stderr[-300:])
