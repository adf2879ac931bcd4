"""Tensor-parallel decode timing on the 8B file (cheap to generate): one GPU vs tp=2 in both exchange modes, and the
phase timeline of shard 0 under the fused mode (B200LLM_MEGA_DBG=1)."""
import sys, os, ctypes as C
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/tools')
import numpy as np, bench
from localai_b200 import capi
arch = sys.argv[1] if len(sys.argv) > 1 else "llama3-8b"
tp = int(sys.argv[2]) if len(sys.argv) > 2 else 2
path, info = bench.ensure_model(arch, "q4_k_m", 0, 0, lambda: None)
if os.environ.get("B200LLM_MEGA_DBG") is None:
    with capi.Engine(path, n_ctx=512, n_batch=32) as e:
        ms, l = e.bench_decode(1, 128, 32, 4); print(f"{arch} one GPU: {ms:.3f} ms/step")
with capi.Engine(path, n_ctx=512, n_batch=32, tp_size=tp) as e:
    for mode, name in ((1, "fused"), (0, "nccl")):
        e.tp_set_mode(mode)
        ms, l = e.bench_decode(1, 128, 32, 4)
        print(f"{arch} tp={tp} {name}: {ms:.3f} ms/step", e.tp_info())
    if os.environ.get("B200LLM_MEGA_DBG"):
        e.tp_set_mode(1)
        ms, l = e.bench_decode(1, 128, 8, 3)
        L = capi.lib(); L.b2_mega_debug.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
        n = L.b2_mega_debug(e._h, None, 0)
        buf = np.zeros(n, dtype=np.int64); L.b2_mega_debug(e._h, buf.ctypes.data_as(C.c_void_p), n)
        nph = n // 16 // 148; d = buf.reshape(148, nph, 16).astype(np.float64)
        kinds = ["qkv", "att", "wo", "gateup", "down"]
        print("fused, profiling build:", ms, "ms/step; per phase kind (ns): start spread | input ready after start mean/max | tiles+sync mean/max | epilogue mean/max | first start -> last end | period")
        for k in range(5):
            idx = np.arange(k + 10, nph - 6, 5)
            s0, s1, s2, s3 = d[:, idx, 0], d[:, idx, 1], d[:, idx, 2], d[:, idx, 3]
            print(f"{kinds[k]:7s} | {(s0.max(0) - s0.min(0)).mean():7.0f} | {(s1 - s0).mean():7.0f} / {(s1 - s0).max(0).mean():7.0f} | {(s2 - s1).mean():7.0f} / {(s2 - s1).max(0).mean():7.0f} | "
                  f"{(s3 - s2).mean():6.0f} / {(s3 - s2).max(0).mean():6.0f} | {(s3.max(0) - s0.min(0)).mean():7.0f} | {(d[:, idx + 1, 0].min(0) - s0.min(0)).mean():7.0f}")
