#!/usr/bin/env python3
"""Synthetic random-init GGUF writer (SURVEY.md §8d "Synthetic inputs").

Writes llama-architecture GGUF v3 files whose quantised tensors are generated
*directly as blocks* (gguf-py cannot quantise K-quants): random nibbles/bits,
6-bit sub-block scales, small finite fp16 super-block scales chosen so that the
dequantised weights have std ~= gain / sqrt(fan_in).  The type mix of the
``q4_k_m`` preset follows llama.cpp's published Q4_K_M rule (Q4_K default,
Q6_K for output.weight and for attn_v / ffn_down on "use_more_bits" layers,
Q5_K for the remaining attn_v on 70B-class models).

The file is streamed tensor by tensor, so an 8B (4.9 GB) or 70B (42.5 GB) file
never has to fit in RAM.  Everything is deterministic in ``--seed``.

Stand-alone on purpose: numpy only (runs on the GPU box, where /root/reference
does not exist).  Tests cross-check the output with gguf-py's GGUFReader.
"""
from __future__ import annotations

import argparse
import json
import os
import struct
import sys
from dataclasses import dataclass, field

import numpy as np

GGUF_MAGIC = 0x46554747
GGUF_VERSION = 3
ALIGN = 32

# ggml type ids
F32, F16, Q8_0, Q4_K, Q5_K, Q6_K = 0, 1, 8, 12, 13, 14
Q4_0, Q4_1, Q5_0, Q5_1 = 2, 3, 6, 7  # legacy 32-element blocks (tensor-core path only on the device)
LEGACY = (Q4_0, Q4_1, Q5_0, Q5_1)
Q2_K, Q3_K = 10, 11  # K-quants without an integer decode kernel on the device (tensor-core path only)
IQ4_NL, IQ4_XS = 20, 23  # 4-bit indices into a fixed non-linear table (tensor-core path only)
TYPE_NAMES = {F32: "F32", F16: "F16", Q8_0: "Q8_0", Q4_K: "Q4_K", Q5_K: "Q5_K", Q6_K: "Q6_K", Q4_0: "Q4_0", Q4_1: "Q4_1",
              Q5_0: "Q5_0", Q5_1: "Q5_1", Q2_K: "Q2_K", Q3_K: "Q3_K", IQ4_NL: "IQ4_NL", IQ4_XS: "IQ4_XS"}
BLOCK = {F32: (1, 4), F16: (1, 2), Q8_0: (32, 34), Q4_K: (256, 144), Q5_K: (256, 176), Q6_K: (256, 210), Q4_0: (32, 18),
         Q4_1: (32, 20), Q5_0: (32, 22), Q5_1: (32, 24), Q2_K: (256, 84), Q3_K: (256, 110),
         IQ4_NL: (32, 18), IQ4_XS: (256, 136)}

# gguf metadata value types
T_U8, T_I8, T_U16, T_I16, T_U32, T_I32, T_F32, T_BOOL, T_STR, T_ARR, T_U64, T_I64, T_F64 = range(13)


# --------------------------------------------------------------------------- model presets
@dataclass
class Arch:
    name: str
    n_layer: int
    n_embd: int
    n_head: int
    n_head_kv: int
    n_ff: int
    n_vocab: int
    n_ctx_train: int = 2048
    rope_theta: float = 10000.0
    rms_eps: float = 1e-5
    tokenizer: str = "spm"  # "spm" | "bpe"
    is_70b: bool = False
    rope_scaling: str = "none"   # "none" | "linear" | "yarn"  (llama.rope.scaling.type)
    rope_factor: float = 1.0     # llama.rope.scaling.factor
    rope_orig_ctx: int = 0       # llama.rope.scaling.original_context_length
    rope_freqs: bool = False     # write rope_freqs.weight (per-dimension frequency divisors, the Llama-3.1 scheme)
    gguf_arch: str = "llama"     # general.architecture and the prefix of the hyper-parameter keys ("llama" | "qwen2")
    qkv_bias: bool = False       # attn_q/k/v.bias tensors (qwen2)
    fused_qkv: bool = False      # phi3: attn_qkv.weight = [q; k; v] rows, ffn_up.weight = [gate; up] rows, no ffn_gate
    head_dim: int = field(init=False)

    def __post_init__(self):
        self.head_dim = self.n_embd // self.n_head


ARCHS = {
    # unit-test sized models: every quant type has K % 256 == 0
    "micro": Arch("micro", 2, 256, 4, 2, 512, 4352, 512, 10000.0, 1e-5, "spm"),
    "mini": Arch("mini", 3, 512, 8, 2, 1024, 4608, 1024, 500000.0, 1e-5, "bpe"),
    "small": Arch("small", 8, 1024, 16, 4, 2816, 8192, 1024, 500000.0, 1e-5, "bpe"),          # deeper: error growth over 8 layers
    # RoPE scaling variants of "micro" (the tokens and tensors are the same; only the rotary tables differ)
    "micro-linear": Arch("micro-linear", 2, 256, 4, 2, 512, 4352, 512, 10000.0, 1e-5, "spm", False, "linear", 4.0, 128),
    "micro-yarn": Arch("micro-yarn", 2, 256, 4, 2, 512, 4352, 512, 10000.0, 1e-5, "spm", False, "yarn", 4.0, 128),
    "micro-l31": Arch("micro-l31", 2, 256, 4, 2, 512, 4352, 512, 10000.0, 1e-5, "spm", False, "none", 1.0, 0, True),
    # qwen2: QKV biases, NEOX rotary pairs (i, i + head_dim / 2), its own BPE pre-tokeniser (digits one at a time)
    "mini-qwen2": Arch("mini-qwen2", 3, 512, 8, 2, 1024, 4608, 1024, 1000000.0, 1e-6, "bpe", False, "none", 1.0, 0, False, "qwen2", True),
    "mini-qwen2-g7": Arch("mini-qwen2-g7", 2, 1792, 14, 2, 1024, 4608, 1024, 1000000.0, 1e-6, "bpe", False, "none", 1.0, 0, False, "qwen2", True),  # 7 query heads per KV head (Qwen2.5-7B's grouping), head_dim 128
    # phi3: fused attn_qkv and gate|up tensors, NEOX rotary pairs, SentencePiece vocabulary, no biases
    "mini-phi3": Arch("mini-phi3", 3, 512, 8, 2, 1024, 4352, 1024, 10000.0, 1e-5, "spm", False, "none", 1.0, 0, False, "phi3", False, True),
    "micro-g8": Arch("micro-g8", 2, 512, 8, 1, 1024, 4352, 512, 10000.0, 1e-5, "spm"),  # 8 query heads per KV head (the 70B grouping), head_dim 64
    "mini-70b": Arch("mini-70b", 4, 2048, 16, 8, 4096, 8192, 1024, 500000.0, 1e-5, "bpe", True),  # 70B type mix (Q5_K attn_v), 8 KV heads; shards 2 / 4 / 8 ways
    # BASELINE.json configs
    "tinyllama-1.1b": Arch("tinyllama-1.1b", 22, 2048, 32, 4, 5632, 32000, 2048, 10000.0, 1e-5, "spm"),
    "llama3-8b": Arch("llama3-8b", 32, 4096, 32, 8, 14336, 128256, 8192, 500000.0, 1e-5, "bpe"),
    "llama3-70b": Arch("llama3-70b", 80, 8192, 64, 8, 28672, 128256, 8192, 500000.0, 1e-5, "bpe", True),
}


def use_more_bits(i: int, n: int) -> bool:
    return i < n // 8 or i >= 7 * n // 8 or (i - n // 8) % 3 == 2


def tensor_type(preset: str, arch: Arch, name: str, layer: int) -> int:
    if name.endswith("_norm.weight"):
        return F32
    if preset == "f32":
        return F32
    if preset == "f16":
        return F16
    if preset == "q8_0":
        return Q8_0
    if preset in ("q4_0", "q4_1", "q5_0", "q5_1"):  # as llama-quantize: the output matrix gets Q6_K
        return Q6_K if name == "output.weight" else {"q4_0": Q4_0, "q4_1": Q4_1, "q5_0": Q5_0, "q5_1": Q5_1}[preset]
    if preset in ("q2_k", "q3_k"):  # token embeddings and the output matrix keep more bits, as llama-quantize does
        if name == "output.weight":
            return Q6_K
        if name == "token_embd.weight":
            return Q4_K
        return Q2_K if preset == "q2_k" else Q3_K
    if preset in ("iq4_nl", "iq4_xs"):  # output Q6_K as llama-quantize does; token embeddings Q4_K (needs a quantiser here)
        if name == "output.weight":
            return Q6_K
        if name == "token_embd.weight":
            return Q4_K
        return IQ4_NL if preset == "iq4_nl" else IQ4_XS
    if preset in ("q4_k", "q5_k", "q6_k"):
        return {"q4_k": Q4_K, "q5_k": Q5_K, "q6_k": Q6_K}[preset]
    if preset == "q4_k_m":
        if name == "output.weight":
            return Q6_K
        if name.endswith("attn_v.weight"):
            if use_more_bits(layer, arch.n_layer):
                return Q6_K
            return Q5_K if arch.is_70b else Q4_K
        if name.endswith("ffn_down.weight"):
            return Q6_K if use_more_bits(layer, arch.n_layer) else Q4_K
        return Q4_K
    if preset == "mixed":  # exercise every kernel inside one small model
        order = [Q4_K, Q6_K, Q5_K, Q8_0, F16]
        if name == "output.weight":
            return Q6_K
        if name == "token_embd.weight":
            return Q4_K
        return order[(hash_name(name) + layer) % len(order)]
    raise ValueError(f"unknown preset {preset}")


def hash_name(s: str) -> int:
    h = 0
    for c in s.encode():
        h = (h * 131 + c) & 0xFFFFFFFF
    return h


# --------------------------------------------------------------------------- random blocks
def _f16_bytes(x: np.ndarray) -> np.ndarray:
    return x.astype(np.float16).view(np.uint8).reshape(x.shape[0], 2)


def _pack_k_scales(sc: np.ndarray, mn: np.ndarray) -> np.ndarray:
    """Pack 8 six-bit scales + 8 six-bit mins into the 12-byte K-quant layout."""
    sc = sc.astype(np.uint8)
    mn = mn.astype(np.uint8)
    out = np.empty((sc.shape[0], 12), dtype=np.uint8)
    out[:, 0:4] = (sc[:, 0:4] & 63) | ((sc[:, 4:8] >> 4) << 6)
    out[:, 4:8] = (mn[:, 0:4] & 63) | ((mn[:, 4:8] >> 4) << 6)
    out[:, 8:12] = (sc[:, 4:8] & 0x0F) | ((mn[:, 4:8] & 0x0F) << 4)
    return out


def random_blocks(rng: np.random.Generator, gtype: int, n_blocks: int, sigma: float) -> np.ndarray:
    """Return uint8 [n_blocks, block_bytes] of random blocks with dequantised std ~ sigma."""
    jitter = rng.uniform(0.8, 1.25, size=n_blocks).astype(np.float32)
    if gtype in LEGACY:  # quantise Gaussian data with llama.cpp's own Python quantiser (gguf-py)
        x = (rng.standard_normal((n_blocks, 32)).astype(np.float32) * (sigma * jitter)[:, None])
        return quantize_rows(x, gtype)
    if gtype == Q2_K:  # scales[16] (scale | min << 4), qs[64], d, dmin: value = d*sc*q - dmin*m, q in 0..3
        scales = rng.integers(0, 256, size=(n_blocks, 16), dtype=np.uint8)
        qs = rng.integers(0, 256, size=(n_blocks, 64), dtype=np.uint8)
        d = (sigma / 11.5) * jitter
        return np.concatenate([scales, qs, _f16_bytes(d), _f16_bytes(d * 1.5)], axis=1)
    if gtype == Q3_K:  # hmask[32], qs[64], scales[12] (6-bit, -32), d: value = d*(sc-32)*q, q in -4..3
        body = rng.integers(0, 256, size=(n_blocks, 108), dtype=np.uint8)
        d = (sigma / 43.0) * jitter
        return np.concatenate([body, _f16_bytes(d)], axis=1)
    if gtype == IQ4_NL:  # d, qs[16]: value = d * level[nibble]; the 16 levels have rms ~ 68
        qs = rng.integers(0, 256, size=(n_blocks, 16), dtype=np.uint8)
        sign = rng.integers(0, 2, size=n_blocks).astype(np.float32) * 2 - 1  # the levels are not centred: mix both signs of d
        return np.concatenate([_f16_bytes((sigma / 68.0) * jitter * sign), qs], axis=1)
    if gtype == IQ4_XS:  # d, scales_h, scales_l[4], qs[128]: value = d * (ls - 32) * level[nibble], ls 6-bit
        body = rng.integers(0, 256, size=(n_blocks, 134), dtype=np.uint8)
        return np.concatenate([_f16_bytes((sigma / (68.0 * 18.5)) * jitter), body], axis=1)
    if gtype == Q8_0:
        qs = rng.integers(-127, 128, size=(n_blocks, 32), dtype=np.int8)
        d = (sigma / 73.3) * jitter
        return np.concatenate([_f16_bytes(d), qs.view(np.uint8)], axis=1)
    if gtype in (Q4_K, Q5_K):
        qmean = 7.5 if gtype == Q4_K else 15.5
        qstd = 4.61 if gtype == Q4_K else 9.23
        sc = rng.integers(24, 64, size=(n_blocks, 8))
        mn = np.clip(sc + rng.integers(-3, 4, size=(n_blocks, 8)), 0, 63)
        d = (sigma / (44.0 * qstd)) * jitter
        dmin = d * qmean
        scales = _pack_k_scales(sc, mn)
        qs = rng.integers(0, 256, size=(n_blocks, 128), dtype=np.uint8)
        if gtype == Q4_K:
            return np.concatenate([_f16_bytes(d), _f16_bytes(dmin), scales, qs], axis=1)
        qh = rng.integers(0, 256, size=(n_blocks, 32), dtype=np.uint8)
        return np.concatenate([_f16_bytes(d), _f16_bytes(dmin), scales, qh, qs], axis=1)
    if gtype == Q6_K:
        ql = rng.integers(0, 256, size=(n_blocks, 128), dtype=np.uint8)
        qh = rng.integers(0, 256, size=(n_blocks, 64), dtype=np.uint8)
        sc = rng.integers(32, 128, size=(n_blocks, 16)).astype(np.int8)
        flip = rng.random(size=(n_blocks, 16)) < 0.1
        sc = np.where(flip, -sc, sc).astype(np.int8)
        d = (sigma / (80.0 * 18.47)) * jitter
        return np.concatenate([ql, qh, sc.view(np.uint8), _f16_bytes(d)], axis=1)
    raise ValueError(gtype)


def random_tensor_bytes(rng, gtype: int, rows: int, cols: int, sigma: float, chunk_rows: int = 4096):
    """Yield the tensor's bytes in row chunks (rows x cols logical, cols innermost)."""
    if gtype == F32:
        for r0 in range(0, rows, chunk_rows):
            r = min(chunk_rows, rows - r0)
            yield (rng.standard_normal((r, cols), dtype=np.float32) * sigma).tobytes()
        return
    if gtype == F16:
        for r0 in range(0, rows, chunk_rows):
            r = min(chunk_rows, rows - r0)
            yield (rng.standard_normal((r, cols), dtype=np.float32) * sigma).astype(np.float16).tobytes()
        return
    bs, _ = BLOCK[gtype]
    assert cols % bs == 0, f"cols {cols} not a multiple of block {bs}"
    bpr = cols // bs
    for r0 in range(0, rows, chunk_rows):
        r = min(chunk_rows, rows - r0)
        yield random_blocks(rng, gtype, r * bpr, sigma).tobytes()


# --------------------------------------------------------------------------- simple float -> block quantisers
# Used only for the token embedding and the (structured) output matrix.  They are plain round-to-nearest
# quantisers, NOT llama.cpp's search-based ones: any valid block is a valid test input, and the dequantised
# values (which is what every implementation consumes) are what matters.
def quantize_rows(x: np.ndarray, gtype: int) -> np.ndarray:
    """x: float32 [rows, cols] -> uint8 [rows, row_bytes]."""
    rows, cols = x.shape
    if gtype == F32:
        return x.astype(np.float32).view(np.uint8).reshape(rows, -1)
    if gtype == F16:
        return x.astype(np.float16).view(np.uint8).reshape(rows, -1)
    if gtype in LEGACY:
        import gguf
        return np.ascontiguousarray(gguf.quants.quantize(np.ascontiguousarray(x, dtype=np.float32), gguf.GGMLQuantizationType(gtype)))
    if gtype == Q8_0:
        b = x.reshape(-1, 32)
        d = np.abs(b).max(axis=1, keepdims=True) / 127.0
        with np.errstate(divide="ignore", invalid="ignore"):
            inv = np.where(d == 0, 0, 1.0 / d)
        q = np.clip(np.rint(b * inv), -127, 127).astype(np.int8)
        return np.concatenate([d.astype(np.float16).view(np.uint8).reshape(-1, 2), q.view(np.uint8)], axis=1).reshape(rows, -1)
    if gtype in (Q4_K, Q5_K):
        qmax = 15 if gtype == Q4_K else 31
        b = x.reshape(-1, 8, 32)
        mn = np.minimum(b.min(axis=2), 0.0)
        mx = b.max(axis=2)
        scale = np.maximum((mx - mn) / qmax, 1e-12)
        d = np.maximum(scale.max(axis=1, keepdims=True) / 63.0, 1e-12).astype(np.float16).astype(np.float32)
        dmin = np.maximum((-mn).max(axis=1, keepdims=True) / 63.0, 1e-12).astype(np.float16).astype(np.float32)
        sc = np.clip(np.rint(scale / d), 1, 63)
        m = np.clip(np.rint(-mn / dmin), 0, 63)
        q = np.clip(np.rint((b + (dmin * m)[:, :, None]) / (d * sc)[:, :, None]), 0, qmax).astype(np.uint8)
        scales = _pack_k_scales(sc, m)
        ql = q & 0x0F
        qs = (ql[:, 0::2, :] | (ql[:, 1::2, :] << 4)).reshape(-1, 128)  # sub-block 2j -> low nibbles, 2j+1 -> high
        parts = [_f16_bytes(d[:, 0]), _f16_bytes(dmin[:, 0]), scales]
        if gtype == Q5_K:
            hb = (q >> 4) & 1  # [n, 8, 32]: bit j of qh[l] belongs to sub-block j
            qh = np.zeros((q.shape[0], 32), dtype=np.uint8)
            for j in range(8):
                qh |= (hb[:, j, :] << j).astype(np.uint8)
            parts.append(qh)
        parts.append(qs)
        return np.concatenate(parts, axis=1).reshape(rows, -1)
    if gtype == Q6_K:
        b = x.reshape(-1, 16, 16)
        amax = np.abs(b).max(axis=2)
        scale = np.maximum(amax / 31.0, 1e-12)
        d = np.maximum(scale.max(axis=1, keepdims=True) / 127.0, 1e-12).astype(np.float16).astype(np.float32)
        sc = np.clip(np.rint(scale / d), 1, 127)
        q = (np.clip(np.rint(b / (d * sc)[:, :, None]), -32, 31) + 32).astype(np.uint8).reshape(-1, 2, 4, 32)  # [n, half, quarter, l]
        lo = q & 0x0F
        ql = np.concatenate([lo[:, :, 0, :] | (lo[:, :, 2, :] << 4), lo[:, :, 1, :] | (lo[:, :, 3, :] << 4)], axis=2)  # [n, 2, 64]
        hi = q >> 4
        qh = hi[:, :, 0, :] | (hi[:, :, 1, :] << 2) | (hi[:, :, 2, :] << 4) | (hi[:, :, 3, :] << 6)  # [n, 2, 32]
        return np.concatenate([ql.reshape(-1, 128), qh.reshape(-1, 64), sc.astype(np.int8).view(np.uint8), _f16_bytes(d[:, 0])],
                              axis=1).reshape(rows, -1)
    raise ValueError(gtype)


_EMB_BLOCK_ROWS = 256


def embedding_rows(seed: int, r0: int, r1: int, cols: int) -> np.ndarray:
    """Deterministic N(0,1) embedding floats for rows [r0, r1): generated in fixed 256-row blocks so that any
    row range can be re-created independently (the structured output matrix re-reads shifted rows)."""
    out = np.empty((r1 - r0, cols), dtype=np.float32)
    b0, b1 = r0 // _EMB_BLOCK_ROWS, (r1 - 1) // _EMB_BLOCK_ROWS
    for b in range(b0, b1 + 1):
        blk = np.random.default_rng([seed, 7_000_003, b]).standard_normal((_EMB_BLOCK_ROWS, cols), dtype=np.float32)
        lo, hi = max(r0, b * _EMB_BLOCK_ROWS), min(r1, (b + 1) * _EMB_BLOCK_ROWS)
        out[lo - r0:hi - r0] = blk[lo - b * _EMB_BLOCK_ROWS:hi - b * _EMB_BLOCK_ROWS]
    return out


NEXT_TOKEN_SHIFT = 1009  # structured models greedily continue t -> (t + 1009) mod n_vocab


def structured_tensor_bytes(kind: str, seed: int, gtype: int, rows: int, cols: int, scale: float, chunk_rows: int = 2048):
    """token_embd: quantise(e[r]);  output: quantise(scale * e[(v - SHIFT) mod V]).

    With output rows aligned to (shifted) embedding rows the logit of token t+SHIFT stands ~5 sigma above the rest,
    so the greedy continuation has a healthy top-1 margin instead of the near-ties a purely random output matrix
    gives (SURVEY.md 7 "hard parts": argmax flips on 1e-3 noise).  All transformer layers stay random."""
    for r0 in range(0, rows, chunk_rows):
        r1 = min(rows, r0 + chunk_rows)
        if kind == "embd":
            x = embedding_rows(seed, r0, r1, cols)
        else:
            s0 = (r0 - NEXT_TOKEN_SHIFT) % rows
            n = r1 - r0
            if s0 + n <= rows:
                x = embedding_rows(seed, s0, s0 + n, cols)
            else:
                x = np.concatenate([embedding_rows(seed, s0, rows, cols), embedding_rows(seed, 0, s0 + n - rows, cols)])
        yield quantize_rows(x * np.float32(scale), gtype).tobytes()


# --------------------------------------------------------------------------- vocab
def load_vocab(kind: str, n_vocab: int, vocab_dir: str):
    """Return dict(tokens, scores, types, merges, bos, eos, model, pre) padded to n_vocab.

    The trained base vocabularies live in tests/golden/vocab_{spm,bpe}.json (made by
    tools/train_vocab.py); if absent a purely synthetic byte-level vocabulary is used.
    """
    path = os.path.join(vocab_dir, f"vocab_{kind}.json")
    if os.path.exists(path):
        with open(path, encoding="utf-8") as f:
            v = json.load(f)
    else:
        v = _fallback_vocab(kind)
    tokens, scores, types = list(v["tokens"]), list(v["scores"]), list(v["types"])
    if len(tokens) > n_vocab:
        raise ValueError(f"base vocab {len(tokens)} > n_vocab {n_vocab}")
    i = len(tokens)
    while len(tokens) < n_vocab:
        tokens.append(f"<|pad_{i}|>")
        scores.append(-1e9 if kind == "spm" else 0.0)
        types.append(5)  # UNUSED
        i += 1
    v = dict(v)
    v.update(tokens=tokens, scores=scores, types=types)
    return v


def _bytes_to_unicode():
    bs = list(range(ord("!"), ord("~") + 1)) + list(range(0xA1, 0xAC + 1)) + list(range(0xAE, 0xFF + 1))
    cs = bs[:]
    n = 0
    for b in range(256):
        if b not in bs:
            bs.append(b)
            cs.append(256 + n)
            n += 1
    return {b: chr(c) for b, c in zip(bs, cs)}


def _fallback_vocab(kind: str):
    if kind == "spm":
        tokens = ["<unk>", "<s>", "</s>"] + [f"<0x{b:02X}>" for b in range(256)]
        types = [2, 3, 3] + [6] * 256
        scores = [0.0] * len(tokens)
        for i, p in enumerate(["▁", "e", "t", "a", "o", "▁t", "he", "▁the", "in", "▁a"]):
            tokens.append(p)
            types.append(1)
            scores.append(-float(i + 1))
        return dict(model="llama", pre="default", tokens=tokens, scores=scores, types=types, merges=[],
                    bos=1, eos=2, unk=0, add_bos=True)
    b2u = _bytes_to_unicode()
    tokens = [b2u[b] for b in range(256)]
    merges = ["Ġ t", "h e", "Ġt he", "i n", "Ġ a"]
    for m in merges:
        tokens.append(m.replace(" ", ""))
    types = [1] * len(tokens)
    tokens += ["<|begin_of_text|>", "<|end_of_text|>", "<|eot_id|>"]
    types += [3, 3, 3]
    return dict(model="gpt2", pre="llama-bpe", tokens=tokens, scores=[0.0] * len(tokens), types=types,
                merges=merges, bos=len(tokens) - 3, eos=len(tokens) - 2, unk=-1, add_bos=True)


# --------------------------------------------------------------------------- gguf writer
def _s(b: bytes) -> bytes:
    return struct.pack("<Q", len(b)) + b


def kv_bytes(key: str, vtype: int, val) -> bytes:
    out = _s(key.encode()) + struct.pack("<I", vtype)
    return out + _val_bytes(vtype, val)


def _val_bytes(vtype: int, val) -> bytes:
    if vtype == T_U32:
        return struct.pack("<I", val)
    if vtype == T_I32:
        return struct.pack("<i", val)
    if vtype == T_F32:
        return struct.pack("<f", val)
    if vtype == T_BOOL:
        return struct.pack("<B", 1 if val else 0)
    if vtype == T_STR:
        return _s(val.encode("utf-8"))
    if vtype == T_U64:
        return struct.pack("<Q", val)
    if vtype == T_ARR:
        etype, items = val
        out = struct.pack("<IQ", etype, len(items))
        if etype == T_STR:
            return out + b"".join(_s(x.encode("utf-8")) for x in items)
        if etype == T_F32:
            return out + np.asarray(items, dtype="<f4").tobytes()
        if etype == T_I32:
            return out + np.asarray(items, dtype="<i4").tobytes()
        raise ValueError(etype)
    raise ValueError(vtype)


def tensor_plan(arch: Arch, preset: str, gain_out: float):
    """[(name, rows, cols, ggml_type, sigma, is_norm)] in file order."""
    d, ff, hd = arch.n_embd, arch.n_ff, arch.head_dim
    kvd = arch.n_head_kv * hd
    plan = [("token_embd.weight", arch.n_vocab, d, tensor_type(preset, arch, "token_embd.weight", 0), 1.0, False)]
    for i in range(arch.n_layer):
        p = f"blk.{i}."
        names = [
            ("attn_norm.weight", 1, d), ("attn_q.weight", d, d), ("attn_k.weight", kvd, d),
            ("attn_v.weight", kvd, d), ("attn_output.weight", d, d), ("ffn_norm.weight", 1, d),
            ("ffn_gate.weight", ff, d), ("ffn_up.weight", ff, d), ("ffn_down.weight", d, ff),
        ]
        if arch.fused_qkv:
            names = [
                ("attn_norm.weight", 1, d), ("attn_qkv.weight", d + 2 * kvd, d), ("attn_output.weight", d, d),
                ("ffn_norm.weight", 1, d), ("ffn_up.weight", 2 * ff, d), ("ffn_down.weight", d, ff),
            ]
        for nm, rows, cols in names:
            t = tensor_type(preset, arch, p + nm, i)
            norm = nm.endswith("_norm.weight")
            plan.append((p + nm, rows, cols, t, 0.0 if norm else 1.0 / np.sqrt(cols), norm))
            if arch.qkv_bias and nm in ("attn_q.weight", "attn_k.weight", "attn_v.weight"):
                plan.append((p + nm.replace(".weight", ".bias"), 1, rows, F32, 0.0, True))
    if arch.rope_freqs:
        plan.append(("rope_freqs.weight", 1, hd // 2, F32, 0.0, True))
    plan.append(("output_norm.weight", 1, d, F32, 0.0, True))
    plan.append(("output.weight", arch.n_vocab, d, tensor_type(preset, arch, "output.weight", 0),
                 gain_out / np.sqrt(d), False))
    return plan


def llama31_rope_factors(head_dim: int, theta: float, orig_ctx: int, factor: float = 8.0, low: float = 1.0, high: float = 4.0):
    """Per-dimension frequency divisors of the Llama-3.1 long-context scheme (what convert_hf_to_gguf stores as
    rope_freqs.weight): wavelengths shorter than orig_ctx / high keep their frequency, longer than orig_ctx / low are
    divided by `factor`, in between a smooth blend."""
    out = np.empty(head_dim // 2, dtype=np.float64)
    for i in range(head_dim // 2):
        freq = theta ** (-2.0 * i / head_dim)
        wavelen = 2 * np.pi / freq
        if wavelen < orig_ctx / high:
            out[i] = 1.0
        elif wavelen > orig_ctx / low:
            out[i] = factor
        else:
            smooth = (orig_ctx / wavelen - low) / (high - low)
            out[i] = 1.0 / ((1 - smooth) / factor + smooth)
    return out


def tensor_nbytes(rows, cols, t):
    bs, bb = BLOCK[t]
    return rows * (cols // bs) * bb


def _write_tensor(job):
    """Generate one tensor (its own RNG stream) and write it at its offset in the file."""
    path, off, idx, name, rows, cols, t, sigma, is_norm, seed, structured = job[:11]
    rng = np.random.default_rng([seed, idx])
    n = 0
    with open(path, "r+b") as f:
        f.seek(off)
        if name.endswith(".bias"):
            w = (0.3 * rng.standard_normal(cols)).astype(np.float32)
            f.write(w.tobytes())
            n = w.nbytes
        elif name == "rope_freqs.weight":
            w = llama31_rope_factors(cols * 2, *job[11]).astype(np.float32)
            f.write(w.tobytes())
            n = w.nbytes
        elif is_norm:
            w = (1.0 + 0.1 * rng.standard_normal(cols)).astype(np.float32)
            f.write(w.tobytes())
            n = w.nbytes
        elif structured and name in ("token_embd.weight", "output.weight"):
            kind = "embd" if name == "token_embd.weight" else "out"
            for chunk in structured_tensor_bytes(kind, seed, t, rows, cols, sigma):
                f.write(chunk)
                n += len(chunk)
        else:
            for chunk in random_tensor_bytes(rng, t, rows, cols, sigma):
                f.write(chunk)
                n += len(chunk)
    assert n == tensor_nbytes(rows, cols, t), (name, n)
    return n


def write_gguf(path: str, arch: Arch, preset: str, seed: int = 0, gain_out: float = 4.0,
               vocab_dir: str | None = None, n_ctx_train: int | None = None, quiet: bool = False,
               structured: bool = True, workers: int = 0) -> dict:
    vocab_dir = vocab_dir or os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden")
    vocab = load_vocab(arch.tokenizer, arch.n_vocab, vocab_dir)
    plan = tensor_plan(arch, preset, gain_out)

    ftype = {"f32": 0, "f16": 1, "q8_0": 7, "q4_k_m": 15, "q4_k": 14, "q5_k": 16, "q6_k": 18, "mixed": 15,
             "q4_0": 2, "q4_1": 3, "q5_0": 8, "q5_1": 9, "q2_k": 10, "q3_k": 12, "iq4_nl": 25, "iq4_xs": 30}[preset]  # llama_ftype
    kvs = [
        kv_bytes("general.architecture", T_STR, arch.gguf_arch),
        kv_bytes("general.name", T_STR, f"{arch.name}-{preset}-random-seed{seed}"),
        kv_bytes("general.file_type", T_U32, ftype),
        kv_bytes("general.quantization_version", T_U32, 2),
        kv_bytes("general.alignment", T_U32, ALIGN),
        kv_bytes(arch.gguf_arch + ".context_length", T_U32, n_ctx_train or arch.n_ctx_train),
        kv_bytes(arch.gguf_arch + ".embedding_length", T_U32, arch.n_embd),
        kv_bytes(arch.gguf_arch + ".block_count", T_U32, arch.n_layer),
        kv_bytes(arch.gguf_arch + ".feed_forward_length", T_U32, arch.n_ff),
        kv_bytes(arch.gguf_arch + ".attention.head_count", T_U32, arch.n_head),
        kv_bytes(arch.gguf_arch + ".attention.head_count_kv", T_U32, arch.n_head_kv),
        kv_bytes(arch.gguf_arch + ".attention.layer_norm_rms_epsilon", T_F32, arch.rms_eps),
        kv_bytes(arch.gguf_arch + ".rope.freq_base", T_F32, arch.rope_theta),
        kv_bytes(arch.gguf_arch + ".rope.dimension_count", T_U32, arch.head_dim),
        kv_bytes(arch.gguf_arch + ".vocab_size", T_U32, arch.n_vocab),
        kv_bytes("tokenizer.ggml.model", T_STR, vocab["model"]),
        kv_bytes("tokenizer.ggml.pre", T_STR, "qwen2" if arch.gguf_arch == "qwen2" else vocab["pre"]),
        kv_bytes("tokenizer.ggml.tokens", T_ARR, (T_STR, vocab["tokens"])),
        kv_bytes("tokenizer.ggml.scores", T_ARR, (T_F32, vocab["scores"])),
        kv_bytes("tokenizer.ggml.token_type", T_ARR, (T_I32, vocab["types"])),
        kv_bytes("tokenizer.ggml.bos_token_id", T_U32, vocab["bos"]),
        kv_bytes("tokenizer.ggml.eos_token_id", T_U32, vocab["eos"]),
        kv_bytes("tokenizer.ggml.add_bos_token", T_BOOL, bool(vocab.get("add_bos", True)) and arch.gguf_arch != "qwen2"),
        kv_bytes("tokenizer.ggml.add_eos_token", T_BOOL, False),
    ]
    if arch.rope_scaling != "none":
        kvs.append(kv_bytes(arch.gguf_arch + ".rope.scaling.type", T_STR, arch.rope_scaling))
        kvs.append(kv_bytes(arch.gguf_arch + ".rope.scaling.factor", T_F32, arch.rope_factor))
        if arch.rope_orig_ctx:
            kvs.append(kv_bytes(arch.gguf_arch + ".rope.scaling.original_context_length", T_U32, arch.rope_orig_ctx))
    if vocab.get("unk", -1) >= 0:
        kvs.append(kv_bytes("tokenizer.ggml.unknown_token_id", T_U32, vocab["unk"]))
    if vocab["merges"]:
        kvs.append(kv_bytes("tokenizer.ggml.merges", T_ARR, (T_STR, vocab["merges"])))
    if vocab.get("eot", -1) >= 0:
        kvs.append(kv_bytes("tokenizer.ggml.eot_token_id", T_U32, vocab["eot"]))

    infos = []
    off = 0
    for name, rows, cols, t, _, _ in plan:
        dims = [cols] if rows == 1 and (name.endswith("_norm.weight") or name == "rope_freqs.weight" or name.endswith(".bias")) else [cols, rows]
        infos.append(_s(name.encode()) + struct.pack("<I", len(dims)) + b"".join(struct.pack("<Q", x) for x in dims)
                     + struct.pack("<IQ", t, off))
        nb = tensor_nbytes(rows, cols, t)
        off += (nb + ALIGN - 1) // ALIGN * ALIGN

    header = struct.pack("<IIQQ", GGUF_MAGIC, GGUF_VERSION, len(plan), len(kvs))
    meta = header + b"".join(kvs) + b"".join(infos)
    pad = (-len(meta)) % ALIGN
    # Every tensor has its own generator (seed, index) and a known offset, so tensors are written independently: big files
    # (the 8B / 70B configurations) are filled by a pool of worker processes, byte-identical to the serial order.
    data_off = len(meta) + pad
    jobs, off = [], data_off
    type_bytes: dict[str, int] = {}
    for idx, (name, rows, cols, t, sigma, is_norm) in enumerate(plan):
        nb = tensor_nbytes(rows, cols, t)
        jobs.append((path, off, idx, name, rows, cols, t, sigma, is_norm, seed, structured, (arch.rope_theta, arch.rope_orig_ctx or 64)))
        off += (nb + ALIGN - 1) // ALIGN * ALIGN
        type_bytes[TYPE_NAMES[t]] = type_bytes.get(TYPE_NAMES[t], 0) + nb
    total = off
    with open(path, "wb") as f:
        f.write(meta + b"\0" * pad)
        f.truncate(total)  # the alignment gaps between tensors read as zeros
    n_workers = workers if workers > 0 else (min(os.cpu_count() or 1, 32) if total > (1 << 30) else 1)
    if n_workers > 1:
        import multiprocessing as mp

        jobs.sort(key=lambda j: -tensor_nbytes(j[4], j[5], j[6]))  # largest first: better balance
        with mp.get_context("fork").Pool(n_workers) as pool:
            for _ in pool.imap_unordered(_write_tensor, jobs, chunksize=1):
                pass
    else:
        for j in jobs:
            _write_tensor(j)
    # bytes streamed per decoded token = every weight except the embedding table
    # (one row gathered) -- SURVEY.md §8(d)
    stream = sum(tensor_nbytes(r, c, t) for (nm, r, c, t, _, _) in plan if nm != "token_embd.weight")
    info = dict(path=path, arch=arch.name, preset=preset, seed=seed, file_bytes=total, n_tensors=len(plan),
                structured=structured, next_token_shift=NEXT_TOKEN_SHIFT if structured else None,
                type_bytes=type_bytes, decode_stream_bytes=stream)
    if not quiet:
        print(json.dumps(info))
    return info


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("--arch", required=True, choices=sorted(ARCHS))
    ap.add_argument("--preset", required=True,
                    choices=["f32", "f16", "q8_0", "q4_k_m", "q4_k", "q5_k", "q6_k", "mixed", "q4_0", "q4_1", "q5_0", "q5_1", "q2_k", "q3_k", "iq4_nl", "iq4_xs"])
    ap.add_argument("--out", required=True)
    ap.add_argument("--seed", type=int, default=0)
    ap.add_argument("--gain-out", type=float, default=4.0, help="output.weight std multiplier (argmax margin)")
    ap.add_argument("--n-layer", type=int, default=None, help="override layer count (debug)")
    ap.add_argument("--unstructured", action="store_true", help="purely random output matrix (near-flat logits)")
    args = ap.parse_args(argv)
    arch = ARCHS[args.arch]
    if args.n_layer:
        arch = Arch(**{**{k: getattr(arch, k) for k in
                          ["name", "n_layer", "n_embd", "n_head", "n_head_kv", "n_ff", "n_vocab", "n_ctx_train",
                           "rope_theta", "rms_eps", "tokenizer", "is_70b"]}, "n_layer": args.n_layer})
    write_gguf(args.out, arch, args.preset, args.seed, args.gain_out, structured=not args.unstructured)


if __name__ == "__main__":
    sys.exit(main())
