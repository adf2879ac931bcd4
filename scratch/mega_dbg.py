import sys, os, ctypes as C
os.environ["B200LLM_MEGA_DBG"] = "1"
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/tools')
import numpy as np, bench
from localai_b200 import capi
arch, preset = (sys.argv[1], sys.argv[2]) if len(sys.argv) > 2 else ("llama3-8b", "q4_k_m")
path, info = bench.ensure_model(arch, preset, 0, 0, lambda: None)
e = capi.Engine(path, n_ctx=1024, n_batch=512, n_parallel=1)
ms, l = e.bench_decode(1, 128, 8, 3)
print("ms/step", ms, "launches", l)
L = capi.lib(); L.b2_mega_debug.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
n = L.b2_mega_debug(e._h, None, 0)
buf = np.zeros(n, dtype=np.int64); L.b2_mega_debug(e._h, buf.ctypes.data_as(C.c_void_p), n)
nph = n // 16 // 148; d = buf.reshape(148, nph, 16)
kinds = ["qkv","att","wo","gateup","down"]
pro = (d[:,:,1]-d[:,:,0]); stg = (d[:,:,2]-d[:,:,1]); bar = (d[:,:,3]-d[:,:,2])
print("cycles per phase kind (mean over CTAs and layers; max over CTAs mean over layers) -- clock ~1.9 GHz")
for k in range(5):
    idx = np.arange(k, nph-1, 5)
    print(f"        warp0: tiles/phase {d[:,idx,7].mean():.1f} wait/phase {d[:,idx,4].mean():.0f} compute/phase {d[:,idx,5].mean():.0f} (per tile {d[:,idx,5].sum()/max(1,d[:,idx,7].sum()):.0f}) epilogue {d[:,idx,6].mean():.0f}")
    print(f"        prologue parts: setup+xload {d[:,idx,8].mean():.0f} prologue-fn {d[:,idx,10].mean():.0f} [tile loop: adv-cons {d[:,idx,9].mean():.0f} fence+issue {d[:,idx,11].mean():.0f} adv-prod {d[:,idx,15].mean():.0f}] slice {d[:,idx,12].mean():.0f} | tile loop w0 {d[:,idx,14].mean():.0f} wait-for-other-warps {d[:,idx,13].mean():.0f}")
    print(f"{kinds[k]:7s} prologue {pro[:,idx].mean():8.0f}  stages mean {stg[:,idx].mean():8.0f} max-cta {stg[:,idx].max(axis=0).mean():8.0f}  barrier mean {bar[:,idx].mean():8.0f} min-cta {bar[:,idx].min(axis=0).mean():8.0f}")
print("lm_head  prologue", pro[:,-1].mean(), "stages mean", stg[:,-1].mean(), "max", stg[:,-1].max())
tot = (d[:,-1,3]-d[:,0,0]); print("total cycles per CTA mean", tot.mean(), "=> ms at 1.9GHz", tot.mean()/1.9e6)

idx = np.arange(1, nph-1, 5)
a = d[:128][:, idx, :]
print("split attention (CTAs 0-127) cumulative cycles: loads-issued %.0f  q-sync %.0f  scores-written %.0f  mini-barrier %.0f  softmax %.0f  partial-out %.0f  end %.0f" % tuple(a[:,:,k].mean() for k in (8,9,10,11,12,13,14)))
print("split attention: entry->before loads %.0f, after K loads %.0f, after V loads %.0f" % tuple(a[:,:,k].mean() for k in (5,6,8)))
print("raw CTA0 layer3:", d[0, 1+5*3, :].tolist())
print("raw CTA77 layer9:", d[77, 1+5*9, :].tolist())
