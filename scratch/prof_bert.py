"""Two N x 512 BERT-base forwards (N = argv[1], default 16), for an ncu launch list."""
import os, sys
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
    seqs = [[2] + rng.integers(5, 300, 510).tolist() + [3] for _ in range(int(sys.argv[1]) if len(sys.argv) > 1 else 16)]
    b.embed_batch(seqs)
    b.embed_batch(seqs)
