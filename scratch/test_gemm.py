import sys, os, time
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/tools')
import numpy as np
import make_gguf as mg
from localai_b200 import capi
from oracle import oracle as O
rng = np.random.default_rng(0)
for wt, rows, cols, n_tok in [(12, 256, 256, 128), (12, 512, 1024, 70), (14, 300, 512, 130), (8, 256, 2048, 5), (13, 1024, 4096, 512), (1, 256, 512, 33)]:
    if wt == 1:
        raw = (rng.standard_normal((rows, cols)) / np.sqrt(cols)).astype(np.float16).view(np.uint8).reshape(-1)
    else:
        raw = mg.random_blocks(rng, wt, rows * cols // mg.BLOCK[wt][0], 1 / np.sqrt(cols)).reshape(-1)
    w = O.dequantize(wt, raw, rows * cols).reshape(rows, cols).astype(np.float64)
    x = rng.standard_normal((n_tok, cols)).astype(np.float32)
    ref = x.astype(np.float64) @ w.T
    t = time.time(); got = capi.mul_mat(wt, raw, rows, cols, x, path=2); dt = time.time() - t
    err = np.abs(got - ref).max() / np.abs(ref).max()
    print(f"type {wt} [{rows}x{cols}] x {n_tok}: rel err {err:.3e} finite {np.isfinite(got).all()} ({dt*1e3:.1f} ms)", flush=True)
