import sys, os, time
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/tools')
import numpy as np
import make_gguf as mg
from localai_b200 import capi
from oracle import oracle as O
cases = [("micro","q8_0"),("mini","q4_k_m"),("mini","q5_k"),("mini","q6_k")]
if len(sys.argv) > 1: cases = [tuple(sys.argv[1].split(":"))]
for arch, preset in cases:
    path = f"/tmp/mega_{arch}_{preset}.gguf"
    mg.write_gguf(path, mg.ARCHS[arch], preset, seed=7, quiet=True)
    ref = O.OracleModel(path, n_ctx=160)
    prompt = np.random.default_rng(9).integers(0, ref.n_vocab, 40).astype(np.int32)
    want, _, marg = ref.generate_greedy(prompt, 48)
    with capi.Engine(path, n_ctx=160, n_batch=32) as e:
        t=time.time(); got, tp, td = e.decode_greedy(prompt, 48); dt=time.time()-t
        # logits of a single decode step through the mega path vs oracle
        e.eval(prompt)
    print(arch, preset, "match" if got.tolist()==want.tolist() else "MISMATCH", got[:8].tolist(), want[:8].tolist(), "decode ms/token %.3f"%(td/48), flush=True)
