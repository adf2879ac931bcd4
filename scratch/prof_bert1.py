"""One-text-per-call BERT-base latency: wall per call vs the sum of kernel times (run under ncu for the latter)."""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "tools"))
sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
import make_bert_gguf as mb
from localai_b200 import capi
path = "/tmp/bert_base_f16.gguf"
if not os.path.exists(path):
    mb.write_bert_gguf(path, mb.BERT_ARCHS["bert-base"], seed=0)
rng = np.random.default_rng(0)
with capi.Bert(path) as b:
    for n in (512, 128, 32):
        ids = [2] + rng.integers(5, 300, n - 2).tolist() + [3]
        b.embed_tokens(ids)
        t = time.time()
        for _ in range(50 if "NCU" not in os.environ else 1):
            b.embed_tokens(ids)
        dt = (time.time() - t) / (50 if "NCU" not in os.environ else 1)
        print(f"n={n}: {dt * 1e3:.3f} ms per call")
