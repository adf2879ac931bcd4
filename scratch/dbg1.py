import sys, os
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/tools')
import numpy as np
import make_gguf as mg
from localai_b200 import capi
from oracle import oracle as O
for arch, preset in [("micro","q8_0"),("mini","q4_k"),("mini","q6_k"),("mini","f16")]:
    path = f"/tmp/dbg_{arch}_{preset}.gguf"
    mg.write_gguf(path, mg.ARCHS[arch], preset, seed=7, quiet=True)
    for toks in ([100], [100, 200, 300, 400, 500]):
        ref = O.OracleModel(path, n_ctx=64); dbg = ref.enable_debug()
        wl = ref.eval(np.array(toks, dtype=np.int32))
        with capi.Engine(path, n_ctx=64) as e:
            e.arm_debug_hidden()
            gl = e.eval(toks)
            got = e.debug_hidden()
        errs = [float(np.abs(got[l]-dbg[l]).max()/np.abs(dbg[l]).max()) for l in range(dbg.shape[0])]
        print(arch, preset, len(toks), "layer errs", ["%.1e"%x for x in errs], "logit err %.1e" % (np.abs(gl-wl).max()/np.abs(wl).max()))
