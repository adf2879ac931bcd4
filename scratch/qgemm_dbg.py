"""Role timing of qgemm_kernel (profiling build -DQG_DBG): where the MMA thread and dequant warp 0 spend their cycles."""
import os, sys, ctypes as C
os.environ["B200LLM_LIB"] = "/root/repo/localai_b200/lib/libb200llm_qgdbg.so"
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/tools')
import numpy as np
from localai_b200 import capi
import make_gguf as mg
L = capi.lib(); L.b2_qgemm_debug.argtypes = [C.c_void_p, C.c_int]
for (wtype, rows, cols, n_tok) in [(capi.B2_T_Q4_K, 28672, 4096, 64), (capi.B2_T_Q4_K, 4096, 4096, 64), (capi.B2_T_Q6_K, 4096, 14336, 64), (capi.B2_T_Q4_K, 28672, 4096, 16)]:
    rng = np.random.default_rng(1)
    raw = mg.random_blocks(rng, wtype, rows * cols // 256, 0.02).reshape(-1)
    x = rng.standard_normal((n_tok, cols)).astype(np.float32)
    capi.mul_mat(wtype, raw, rows, cols, x, path=6)
    buf = np.zeros(148 * 16, dtype=np.int64); L.b2_qgemm_debug(buf.ctypes.data_as(C.c_void_p), buf.size)
    d = buf.reshape(148, 16).astype(np.float64)
    print(f"type {wtype} {rows}x{cols} n={n_tok}: chunks/CTA {d[:,9].mean():.1f}; dequant warp total cycles {d[:,8].mean():.0f} (max {d[:,8].max():.0f})")
    print("  MMA thread cycles: wait d_empty %.0f | wait b_full %.0f | wait a_full %.0f | issue+commit %.0f" % tuple(d[:, :4].mean(0)))
    print("  dequant warp 0   : wait raw_full %.0f | compute %.0f | wait a_empty %.0f | tmem st+arrive %.0f" % tuple(d[:, 4:8].mean(0)))
    print("  producer         : wait b_empty %.0f | wait raw_empty %.0f | issue %.0f" % tuple(d[:, 10:13].mean(0)))
    print("  CTA              : prologue %.0f | main loops done at %.0f (max %.0f) | with fix-up %.0f (max %.0f)" % (d[:,13].mean(), d[:,14].mean(), d[:,14].max(), d[:,15].mean(), d[:,15].max()))
    st = d[:, 9].mean() * 2
    print("  per stage        : raw wait %.0f compute %.0f a_empty %.0f st %.0f ; mma a_full wait %.0f issue %.0f" % (d[:,4].mean()/st*2, d[:,5].mean()/st, d[:,6].mean()/st, d[:,7].mean()/st, d[:,2].mean()/st, d[:,3].mean()/st))
