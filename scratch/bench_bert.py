"""BERT-base F16 embeddings throughput (BASELINE.json config 4: 256 sequences x 512 tokens), device-timed per call."""
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
    seqs = [[2] + rng.integers(5, 300, 510).tolist() + [3] for _ in range(256)]
    import ctypes as C
    L = capi.lib()
    for per_call in (1, 16, 32, 64):
        calls = []
        for i in range(0, 256, per_call):  # host buffers ready, as a gRPC handler would hold them
            ss = seqs[i:i + per_call]
            ids = np.ascontiguousarray(np.concatenate(ss), dtype=np.int32)
            offs = np.concatenate([[0], np.cumsum([len(x) for x in ss])]).astype(np.int32)
            calls.append((ids, offs, np.empty((len(ss), b.n_embd), dtype=np.float32)))
        b.embed_batch(seqs[:per_call])
        t = time.time()
        for ids, offs, out in calls:
            L.b2_bert_embed_batch(b._h, ids.ctypes.data, offs.ctypes.data, len(offs) - 1, out.ctypes.data)
        dt = time.time() - t
        print(f"bert-base f16, 256 x 512 tokens, {per_call} sequences per call: {256 / dt:.1f} seq/s ({dt * 1e3 / 256:.3f} ms/seq, wall incl. copies)")
