import sys, os
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/tools')
import numpy as np
import make_gguf as mg
from localai_b200 import capi
from oracle import oracle as O
rng = np.random.default_rng(0)
for wt in (12, 8):
    for n_tok in (4, 8, 9, 16, 20):
        rows, cols = 64, 512
        raw = mg.random_blocks(rng, wt, rows*cols//mg.BLOCK[wt][0], 0.05)
        x = rng.standard_normal((n_tok, cols)).astype(np.float32)
        got = capi.mul_mat(wt, raw, rows, cols, x); ref = O.mul_mat(wt, raw, rows, cols, x)
        err = np.abs(got-ref).max(axis=1)/np.abs(ref).max()
        print("mul_mat type", wt, "n_tok", n_tok, "per-token err", ["%.0e"%e for e in err])
path = "/tmp/dbg_mini_q4_k.gguf"
mg.write_gguf(path, mg.ARCHS["mini"], "q4_k", seed=7, quiet=True)
for n in (6, 8, 9, 12, 20):
    toks = (np.arange(n)*37+100).astype(np.int32)
    ref = O.OracleModel(path, n_ctx=64); dbg = ref.enable_debug()
    wl = ref.eval(toks, logits_all=True)
    with capi.Engine(path, n_ctx=64) as e:
        e.arm_debug_hidden()
        gl = e.eval(toks, logits_all=True)
        got = e.debug_hidden()
    errs = [float(np.abs(got[l]-dbg[l]).max()/np.abs(dbg[l]).max()) for l in range(dbg.shape[0])]
    print("n", n, "layer errs", ["%.1e"%x for x in errs], "per-token logit err", ["%.0e"%v for v in np.abs(gl-wl).max(axis=1)/np.abs(wl).max()])
