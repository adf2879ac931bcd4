#!/usr/bin/env python3
"""Mint tests/golden/dequant_*.npz and quant_q8_0.npz with gguf-py (llama.cpp's own Python package, 0.19.0).

Runs in the build container (needs `gguf`).  Each dequant file holds raw GGUF blocks -- random ones plus
edge cases (zero / maximal sub-scales, negative and subnormal super-scales, all-0 and all-1 quants) -- and the
float32 values gguf.quants.dequantize() produces for them.  The oracle (oracle/ggml_ref.c) and the CUDA
kernels must both reproduce these bit for bit.  quant_q8_0.npz pins quantize_row_q8_0 the same way (gguf-py
documents its Q8_0 quantiser as bit-exact with ggml's reference implementation).
"""
import os
import sys

import gguf
import numpy as np

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import make_gguf as mg  # noqa: E402

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden")
QT = gguf.GGMLQuantizationType
TYPES = {"q8_0": (mg.Q8_0, QT.Q8_0), "q4_k": (mg.Q4_K, QT.Q4_K), "q5_k": (mg.Q5_K, QT.Q5_K), "q6_k": (mg.Q6_K, QT.Q6_K)}


def edge_blocks(gtype, rng):
    bs, bb = mg.BLOCK[gtype]
    out = []
    base = mg.random_blocks(rng, gtype, 8, 0.05)
    f16 = lambda v: np.frombuffer(np.float16(v).tobytes(), dtype=np.uint8)  # noqa: E731
    for i, d in enumerate([0.0, -0.0, 6.1e-5, -3.0e-3, 5.96e-8, 1.0, -2.5, 65504.0 / 4096]):
        b = base[i].copy()
        if gtype == mg.Q8_0:
            b[0:2] = f16(d)
            if i == 0:
                b[2:] = 0x80  # -128
            if i == 1:
                b[2:] = 0x7F
        elif gtype in (mg.Q4_K, mg.Q5_K):
            b[0:2] = f16(d)
            b[2:4] = f16(-d if i % 2 else d * 0.5)
            if i == 2:
                b[4:16] = 0xFF  # all scales/mins 63
            if i == 3:
                b[4:16] = 0x00
            if i == 4:
                b[16:] = 0xFF
            if i == 5:
                b[16:] = 0x00
        else:  # Q6_K: ql[128] qh[64] scales[16] d
            b[208:210] = f16(d)
            if i == 2:
                b[192:208] = 0x80  # scale -128
            if i == 3:
                b[192:208] = 0x7F
            if i == 4:
                b[0:192] = 0xFF
            if i == 5:
                b[0:192] = 0x00
        out.append(b)
    return np.stack(out)


def main():
    os.makedirs(GOLD, exist_ok=True)
    rng = np.random.default_rng(1234)
    for name, (gt, qt) in TYPES.items():
        blocks = np.concatenate([mg.random_blocks(rng, gt, 56, 0.02), edge_blocks(gt, rng)], axis=0)
        per256 = 256 // mg.BLOCK[gt][0]
        blocks = blocks[: blocks.shape[0] // per256 * per256]
        vals = gguf.quants.dequantize(blocks.reshape(-1), qt).astype(np.float32).reshape(-1)
        np.savez_compressed(os.path.join(GOLD, f"dequant_{name}.npz"), blocks=blocks, values=vals, ggml_type=gt)
        print(name, blocks.shape, vals.shape, "finite:", bool(np.isfinite(vals).all()))
    h = rng.standard_normal(2048).astype(np.float16)
    np.savez_compressed(os.path.join(GOLD, "dequant_f16.npz"), blocks=h.view(np.uint8), values=h.astype(np.float32),
                        ggml_type=mg.F16)
    # Q8_0 quantisation known answers
    x = (rng.standard_normal(4096) * np.repeat(rng.uniform(1e-3, 30, 128), 32)).astype(np.float32)
    x[0:32] = 0.0
    x[32:64] = np.linspace(-127, 127, 32, dtype=np.float32) * 0.5  # exact .5 ties after scaling
    x[64] = 1e-30
    q = gguf.quants.quantize(x, QT.Q8_0).reshape(-1)
    np.savez_compressed(os.path.join(GOLD, "quant_q8_0.npz"), x=x, blocks=q)
    print("quant_q8_0", x.shape, q.shape)
    mint_legacy()
    mint_small_k()


LEGACY = {"q4_0": (mg.Q4_0, QT.Q4_0), "q4_1": (mg.Q4_1, QT.Q4_1), "q5_0": (mg.Q5_0, QT.Q5_0), "q5_1": (mg.Q5_1, QT.Q5_1)}


def mint_legacy():
    """dequant_{q4_0,q4_1,q5_0,q5_1}.npz: quantised Gaussian data plus raw random bytes (every bit pattern of the
    quants and high bits) under edge-case scales.  Own random stream, so the files above never change."""
    rng = np.random.default_rng(4321)
    f16 = lambda v: np.frombuffer(np.float16(v).tobytes(), dtype=np.uint8)  # noqa: E731
    for name, (gt, qt) in LEGACY.items():
        bb = mg.BLOCK[gt][1]
        blocks = [mg.random_blocks(rng, gt, 48, 0.02)]
        raw = rng.integers(0, 256, size=(16, bb), dtype=np.uint8)
        for i, d in enumerate([0.0, -0.0, 6.1e-5, -3.0e-3, 5.96e-8, 1.0, -2.5, 15.99, 0.25, -0.125, 3e-4, 7.0, -1e-2, 0.5, 2.0, -4.0]):
            raw[i, 0:2] = f16(d)
            if gt in (mg.Q4_1, mg.Q5_1):
                raw[i, 2:4] = f16(-d * 3 if i % 2 else d * 0.5)
            if i == 2:
                raw[i, -16:] = 0xFF
            if i == 3:
                raw[i, -16:] = 0x00
        blocks.append(raw)
        blocks = np.concatenate(blocks, axis=0)
        vals = gguf.quants.dequantize(blocks.reshape(-1), qt).astype(np.float32).reshape(-1)
        np.savez_compressed(os.path.join(GOLD, f"dequant_{name}.npz"), blocks=blocks, values=vals, ggml_type=gt)
        print(name, blocks.shape, vals.shape, "finite:", bool(np.isfinite(vals).all()))


def mint_small_k():
    """dequant_{q2_k,q3_k}.npz: random blocks plus raw random bytes under edge-case super-block scales."""
    rng = np.random.default_rng(987)
    f16 = lambda v: np.frombuffer(np.float16(v).tobytes(), dtype=np.uint8)  # noqa: E731
    for name, gt, qt in (("q2_k", mg.Q2_K, QT.Q2_K), ("q3_k", mg.Q3_K, QT.Q3_K)):
        bb = mg.BLOCK[gt][1]
        raw = rng.integers(0, 256, size=(8, bb), dtype=np.uint8)
        for i, d in enumerate([0.0, -0.0, 6.1e-5, -3.0e-3, 5.96e-8, 1.0, -2.5, 0.25]):
            if gt == mg.Q2_K:
                raw[i, 80:82] = f16(d)
                raw[i, 82:84] = f16(-d * 3 if i % 2 else d * 0.5)
                if i == 2:
                    raw[i, 0:80] = 0xFF
                if i == 3:
                    raw[i, 0:80] = 0x00
            else:
                raw[i, 108:110] = f16(d)
                if i == 2:
                    raw[i, 0:108] = 0xFF
                if i == 3:
                    raw[i, 0:108] = 0x00
        blocks = np.concatenate([mg.random_blocks(rng, gt, 24, 0.02), raw], axis=0)
        vals = gguf.quants.dequantize(blocks.reshape(-1), qt).astype(np.float32).reshape(-1)
        np.savez_compressed(os.path.join(GOLD, f"dequant_{name}.npz"), blocks=blocks, values=vals, ggml_type=gt)
        print(name, blocks.shape, vals.shape, "finite:", bool(np.isfinite(vals).all()))


def mint_iq4():
    """dequant_{iq4_nl,iq4_xs}.npz: random blocks plus raw random bytes under edge-case scales (gguf-py dequantiser)."""
    rng = np.random.default_rng(4321)
    f16 = lambda v: np.frombuffer(np.float16(v).tobytes(), dtype=np.uint8)  # noqa: E731
    for name, gt, qt in (("iq4_nl", mg.IQ4_NL, QT.IQ4_NL), ("iq4_xs", mg.IQ4_XS, QT.IQ4_XS)):
        bb = mg.BLOCK[gt][1]
        raw = rng.integers(0, 256, size=(8, bb), dtype=np.uint8)
        for i, d in enumerate([0.0, -0.0, 6.1e-5, -3.0e-3, 5.96e-8, 1.0, -2.5, 0.25]):
            raw[i, 0:2] = f16(d)
            if i == 2:
                raw[i, 2:] = 0xFF
            if i == 3:
                raw[i, 2:] = 0x00
        blocks = np.concatenate([mg.random_blocks(rng, gt, 24, 0.02), raw], axis=0)
        vals = gguf.quants.dequantize(blocks.reshape(-1), qt).astype(np.float32).reshape(-1)
        np.savez_compressed(os.path.join(GOLD, f"dequant_{name}.npz"), blocks=blocks, values=vals, ggml_type=gt)
        print(name, blocks.shape, vals.shape, "finite:", bool(np.isfinite(vals).all()))


if __name__ == "__main__":
    if "--iq4-only" in sys.argv:
        sys.exit(mint_iq4())
    if "--small-k-only" in sys.argv:
        sys.exit(mint_small_k())
    if "--legacy-only" in sys.argv:
        sys.exit(mint_legacy())
    sys.exit(main())
