#!/usr/bin/env python3
"""Mint tests/golden/tok_bpe_qwen2.json: the BPE vocabulary of tests/golden/vocab_bpe.json under the qwen2 pre-tokeniser.

The qwen2 pre-tokeniser is the llama-3 expression with one digit per word ("\\p{N}" for "\\p{N}{1,3}"), words that are whole
tokens go through the merge loop like any other (no ignore_merges), and no BOS is added.  The ids come from the HF
`tokenizers` library, built from the committed vocabulary and merges; nothing of /root/reference is read.
"""
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = os.path.join(HERE, "..", "tests", "golden")
QWEN2_RE = ("(?i:'s|'t|'re|'ve|'m|'ll|'d)|[^\\r\\n\\p{L}\\p{N}]?\\p{L}+|\\p{N}| ?[^\\s\\p{L}\\p{N}]+[\\r\\n]*|"
            "\\s*[\\r\\n]+|\\s+(?!\\S)|\\s+")


def main():
    from tokenizers import Regex, Tokenizer, decoders, models, pre_tokenizers

    vocab = json.load(open(os.path.join(GOLD, "vocab_bpe.json"), encoding="utf-8"))
    specials = {t for t, ty in zip(vocab["tokens"], vocab["types"]) if ty == 3}
    table = {t: i for i, t in enumerate(vocab["tokens"]) if t not in specials}
    merges = [tuple(m.split(" ")) for m in vocab["merges"]]
    tok = Tokenizer(models.BPE(vocab=table, merges=merges, ignore_merges=False))
    tok.pre_tokenizer = pre_tokenizers.Sequence([
        pre_tokenizers.Split(Regex(QWEN2_RE), behavior="isolated", invert=False),
        pre_tokenizers.ByteLevel(add_prefix_space=False, use_regex=False)])
    tok.decoder = decoders.ByteLevel()
    texts = [g["text"] for g in json.load(open(os.path.join(GOLD, "tok_bpe.json"), encoding="utf-8"))]
    texts = [t for t in texts if not any(s in t for s in specials)]
    texts += ["2024-11-07 12:34:56", "x1y22z333 4444", "٣٤٥ and ४५६ digits", "v1.2.3-rc10"]
    gold = [dict(text=t, ids=tok.encode(t, add_special_tokens=False).ids) for t in texts]
    with open(os.path.join(GOLD, "tok_bpe_qwen2.json"), "w", encoding="utf-8") as f:
        json.dump(gold, f, ensure_ascii=False)
    print("samples", len(gold))


if __name__ == "__main__":
    sys.exit(main())
