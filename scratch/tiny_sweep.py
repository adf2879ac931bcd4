import sys, os, itertools, subprocess
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/tools')
arch, preset = sys.argv[1], sys.argv[2]
code = f'''
import sys; sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tools")
import bench
from localai_b200 import capi
path, info = bench.ensure_model("{arch}", "{preset}", 0, 0, lambda: None)
e = capi.Engine(path, n_ctx=1024, n_batch=128, n_parallel=1)
ms, l = e.bench_decode(1, 128, 64, 5)
print("RESULT", ms)
'''
for parts, ndq in itertools.product(["1", "2", "4"], ["0", "1"]):
    env = dict(os.environ, B200LLM_ATT_PARTS=parts, B200LLM_NO_DIST_QUANT=ndq)
    r = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True, timeout=300)
    ms = [l for l in r.stdout.splitlines() if l.startswith("RESULT")]
    print(arch, "att_parts", parts, "no_dist_quant", ndq, ms[0] if ms else r.stderr[-300:])
