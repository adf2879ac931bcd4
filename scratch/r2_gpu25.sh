cd /root/repo
timeout 300 python scratch/qgemm_dbg.py 2>&1 | tail -16
timeout 600 python -m pytest tests/test_parity_gpu.py -x -q -k "qgemm or gemm_tcgen05" 2>&1 | tail -3
timeout 300 python scratch/b64_bench.py 2>&1 | tail -1
timeout 300 python - <<'PY'
import sys; sys.path.insert(0,'/root/repo'); sys.path.insert(0,'/root/repo/tools')
import bench
from localai_b200 import capi
path, info = bench.ensure_model("llama3-8b", "q4_k_m", 0, 0, lambda: None)
e = capi.Engine(path, n_ctx=4096, n_batch=2048, n_parallel=1)
import numpy as np
for i in range(3):
    ms = e.bench_prefill(2048, 3, 1)
    print("prefill 2048 ms", ms)
PY
