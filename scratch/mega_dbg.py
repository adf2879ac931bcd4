"""Per-phase timeline of the decode kernel (profiling build, B200LLM_MEGA_DBG=1): global-timer stamps of every CTA at
phase start / input ready / past the mid-phase sync / phase end, for one decode step of llama3-8b q4_k_m."""
import sys, os, ctypes as C
os.environ["B200LLM_MEGA_DBG"] = "1"
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/tools')
import numpy as np, bench
from localai_b200 import capi
arch, preset = (sys.argv[1], sys.argv[2]) if len(sys.argv) > 2 else ("llama3-8b", "q4_k_m")
path, info = bench.ensure_model(arch, preset, 0, 0, lambda: None)
e = capi.Engine(path, n_ctx=1024, n_batch=512, n_parallel=1)
ms, l = e.bench_decode(1, 128, 8, 3)
print("ms/step (profiling build)", ms, "launches", l)
L = capi.lib(); L.b2_mega_debug.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
n = L.b2_mega_debug(e._h, None, 0)
buf = np.zeros(n, dtype=np.int64); L.b2_mega_debug(e._h, buf.ctypes.data_as(C.c_void_p), n)
nph = n // 16 // 148; d = buf.reshape(148, nph, 16).astype(np.float64)
os.makedirs('gpurun_out', exist_ok=True); np.save('gpurun_out/r2_mega_dbg_raw.npy', buf.reshape(148, nph, 16))
t00 = d[:, 0, 0].min()
kinds = ["qkv", "att", "wo", "gateup", "down"]
print("all times in ns (global timer); per phase kind, mean over layers 2..30")
print("kind    | start: spread over CTAs | input ready after start (mean / max-CTA) | tiles+sync (mean/max) | epilogue (mean/max) | phase: first start -> last end | next phase first start - this first start")
for k in range(5):
    idx = np.arange(k + 10, nph - 6, 5)
    s0, s1, s2, s3 = d[:, idx, 0], d[:, idx, 1], d[:, idx, 2], d[:, idx, 3]
    spread = (s0.max(0) - s0.min(0)).mean()
    rin = (s1 - s0); til = (s2 - s1); epi = (s3 - s2)
    whole = (s3.max(0) - s0.min(0)).mean()
    nxt = (d[:, idx + 1, 0].min(0) - s0.min(0)).mean()
    print(f"{kinds[k]:7s} | {spread:7.0f} | {rin.mean():7.0f} / {rin.max(0).mean():7.0f} | {til.mean():7.0f} / {til.max(0).mean():7.0f} | {epi.mean():6.0f} / {epi.max(0).mean():6.0f} | {whole:7.0f} | {nxt:7.0f}"
          f"   warp0: tiles {d[:, idx, 6].mean():.1f} wait {d[:, idx, 4].mean():.0f} cyc compute {d[:, idx, 5].mean():.0f} cyc; rows/CTA {d[:, idx, 12].mean():.1f}")
lay = (d[:, 15:nph - 6:5, 0].min(0)); print("layer period (ns):", np.diff(lay).mean(), " => 32 layers:", np.diff(lay).mean() * 32 / 1e6, "ms")
print("lm head: first start -> last end (ns):", d[:, -1, 3].max() - d[:, -1, 0].min())
print("whole kernel (ns):", d[:, -1, 3].max() - t00)
# who is last? histogram of the CTA that ends each phase last
last = d[:, 10:nph - 6, 3].argmax(0); cnt = np.bincount(last, minlength=148)
print("CTAs that most often finish a phase last:", np.argsort(-cnt)[:10].tolist(), cnt[np.argsort(-cnt)[:10]].tolist())
first_in = d[:, 10:nph - 6, 1].argmax(0); cnt = np.bincount(first_in, minlength=148)
print("CTAs whose input is ready last most often:", np.argsort(-cnt)[:10].tolist(), cnt[np.argsort(-cnt)[:10]].tolist())

# is the spread systematic?  per-CTA mean of (phase end - earliest phase end) for the two big phases, over layers
for k, name in ((3, "gateup"), (4, "down"), (0, "qkv"), (2, "wo")):
    idx = np.arange(k + 10, nph - 6, 5)
    late = d[:, idx, 3] - d[:, idx, 3].min(0)            # how long after the first CTA this CTA ends the phase
    til = d[:, idx, 2] - d[:, idx, 1]
    print(f"{name}: lateness at phase end (ns): mean over CTAs {late.mean():.0f}; per-CTA means: std {late.mean(1).std():.0f}, min {late.mean(1).min():.0f}, max {late.mean(1).max():.0f}; "
          f"within-CTA std over layers {late.std(1).mean():.0f}; tile time per-CTA means: std {til.mean(1).std():.0f} of {til.mean():.0f}")
    worst = np.argsort(-late.mean(1))[:12]; print("   latest CTAs:", worst.tolist(), late.mean(1)[worst].round(0).tolist())
