#!/usr/bin/env python3
"""Random-init BERT encoder as a GGUF file (general.architecture = "bert", llama.cpp's tensor and key names) for the
embeddings path: BASELINE.json config 4 (BERT-base F16, 256 x 512 tokens) and small test models.

No checkpoint can be fetched here, so weights are seeded Gaussians; the WordPiece vocabulary is synthetic (special
tokens, every printable ASCII character as word-initial and as continuation piece, a few hundred words and suffixes),
in llama.cpp's convention: word-initial pieces carry a leading "▁", continuation pieces are bare.
"""
from __future__ import annotations

import argparse
import os
import struct
import sys
from dataclasses import dataclass

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import make_gguf as mg  # noqa: E402


@dataclass
class BertArch:
    name: str
    n_layer: int
    n_embd: int
    n_head: int
    n_ff: int
    n_vocab: int
    n_ctx: int = 512
    eps: float = 1e-12


BERT_ARCHS = {
    "bert-mini": BertArch("bert-mini", 3, 256, 4, 1024, 2048),       # head_dim 64
    "minilm-l3": BertArch("minilm-l3", 3, 384, 12, 1536, 2048),      # head_dim 32 (all-MiniLM shape, 384-d output)
    "bert-base": BertArch("bert-base", 12, 768, 12, 3072, 30522),
}

WORDS = ("the of and to in a is that for it as was with be by on not he this are or his from at which but have an had "
         "they you were their one all we can her has there been if more when will would who so no local ai model "
         "embedding vector search query document quick brown fox jumps over lazy dog hello world gpu memory token").split()
SUFFIXES = "s ed ing ly er est ion tion ment ness able al ic ize".split()


def make_vocab(n_vocab: int):
    toks = ["[PAD]", "[UNK]", "[CLS]", "[SEP]", "[MASK]"]
    types = [3] * 5
    chars = [chr(c) for c in range(33, 127)] + ["é", "ü", "ß", "日", "本", "語"]
    for c in chars:
        c = c.lower()
        for t in ("▁" + c, c):
            if t not in toks:
                toks.append(t)
                types.append(1)
    for w in WORDS:
        if "▁" + w not in toks:
            toks.append("▁" + w)
            types.append(1)
    for s in SUFFIXES:
        if s not in toks:
            toks.append(s)
            types.append(1)
    assert len(toks) <= n_vocab, len(toks)
    i = 0
    while len(toks) < n_vocab:
        toks.append(f"[unused{i}]")
        types.append(5)
        i += 1
    return toks, types


def plan(a: BertArch, wtype: int):
    """(name, rows, cols, ggml type, sigma or None for norm weights / 0 for biases)"""
    d, ff = a.n_embd, a.n_ff
    s = 1.0 / np.sqrt(d)
    out = [("token_embd.weight", a.n_vocab, d, wtype, 1.0), ("token_types.weight", 2, d, mg.F32, 0.3),
           ("position_embd.weight", a.n_ctx, d, mg.F32, 0.3), ("token_embd_norm.weight", 1, d, mg.F32, None),
           ("token_embd_norm.bias", 1, d, mg.F32, 0.0)]
    for i in range(a.n_layer):
        p = f"blk.{i}."
        for nm in ("attn_q", "attn_k", "attn_v", "attn_output"):
            out += [(p + nm + ".weight", d, d, wtype, s * (2.0 if nm in ("attn_q", "attn_k") else 1.0)), (p + nm + ".bias", 1, d, mg.F32, 0.0)]
        out += [(p + "attn_output_norm.weight", 1, d, mg.F32, None), (p + "attn_output_norm.bias", 1, d, mg.F32, 0.0),
                (p + "ffn_up.weight", ff, d, wtype, s), (p + "ffn_up.bias", 1, ff, mg.F32, 0.0),
                (p + "ffn_down.weight", d, ff, wtype, 1.0 / np.sqrt(ff)), (p + "ffn_down.bias", 1, d, mg.F32, 0.0),
                (p + "layer_output_norm.weight", 1, d, mg.F32, None), (p + "layer_output_norm.bias", 1, d, mg.F32, 0.0)]
    return out


def write_bert_gguf(path: str, a: BertArch, wtype: int = mg.F16, seed: int = 0, quiet: bool = True):
    toks, types = make_vocab(a.n_vocab)
    K, T = mg.kv_bytes, mg
    kvs = [
        K("general.architecture", T.T_STR, "bert"), K("general.name", T.T_STR, f"{a.name}-random-seed{seed}"),
        K("general.file_type", T.T_U32, 1 if wtype == mg.F16 else 0), K("general.alignment", T.T_U32, mg.ALIGN),
        K("bert.context_length", T.T_U32, a.n_ctx), K("bert.embedding_length", T.T_U32, a.n_embd),
        K("bert.block_count", T.T_U32, a.n_layer), K("bert.feed_forward_length", T.T_U32, a.n_ff),
        K("bert.attention.head_count", T.T_U32, a.n_head), K("bert.attention.layer_norm_epsilon", T.T_F32, a.eps),
        K("bert.attention.causal", T.T_BOOL, False), K("bert.pooling_type", T.T_U32, 1),
        K("tokenizer.ggml.model", T.T_STR, "bert"), K("tokenizer.ggml.tokens", T.T_ARR, (T.T_STR, toks)),
        K("tokenizer.ggml.token_type", T.T_ARR, (T.T_I32, types)), K("tokenizer.ggml.token_type_count", T.T_U32, 2),
        K("tokenizer.ggml.bos_token_id", T.T_U32, 2), K("tokenizer.ggml.eos_token_id", T.T_U32, 3),
        K("tokenizer.ggml.unknown_token_id", T.T_U32, 1), K("tokenizer.ggml.seperator_token_id", T.T_U32, 3),
        K("tokenizer.ggml.padding_token_id", T.T_U32, 0), K("tokenizer.ggml.cls_token_id", T.T_U32, 2),
        K("tokenizer.ggml.mask_token_id", T.T_U32, 4),
    ]
    pl = plan(a, wtype)
    infos, off = [], 0
    for name, rows, cols, t, _ in pl:
        dims = [cols] if rows == 1 else [cols, rows]
        infos.append(mg._s(name.encode()) + struct.pack("<I", len(dims)) + b"".join(struct.pack("<Q", x) for x in dims)
                     + struct.pack("<IQ", t, off))
        off += (mg.tensor_nbytes(rows, cols, t) + mg.ALIGN - 1) // mg.ALIGN * mg.ALIGN
    meta = struct.pack("<IIQQ", mg.GGUF_MAGIC, mg.GGUF_VERSION, len(pl), len(kvs)) + b"".join(kvs) + b"".join(infos)
    with open(path, "wb") as f:
        f.write(meta + b"\0" * ((-len(meta)) % mg.ALIGN))
        for idx, (name, rows, cols, t, sigma) in enumerate(pl):
            rng = np.random.default_rng([seed, 77, idx])
            if sigma is None:
                data = (1.0 + 0.1 * rng.standard_normal(cols)).astype(np.float32).tobytes()
            elif sigma == 0.0:
                data = (0.05 * rng.standard_normal(cols)).astype(np.float32).tobytes()
            else:
                data = b"".join(mg.random_tensor_bytes(rng, t, rows, cols, sigma))
            assert len(data) == mg.tensor_nbytes(rows, cols, t), name
            f.write(data + b"\0" * ((-len(data)) % mg.ALIGN))
    if not quiet:
        print(f"{path}: {a.name}, {len(pl)} tensors, {os.path.getsize(path) / 1e6:.1f} MB")
    return path


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--arch", default="bert-base", choices=sorted(BERT_ARCHS))
    ap.add_argument("--type", default="f16", choices=["f16", "f32", "q8_0"])
    ap.add_argument("--seed", type=int, default=0)
    ap.add_argument("out")
    a = ap.parse_args()
    write_bert_gguf(a.out, BERT_ARCHS[a.arch], {"f16": mg.F16, "f32": mg.F32, "q8_0": mg.Q8_0}[a.type], a.seed, quiet=False)
