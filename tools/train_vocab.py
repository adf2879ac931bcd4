#!/usr/bin/env python3
"""Train the two small base vocabularies used by tools/make_gguf.py and mint the tokeniser golden vectors.

Runs ONLY in the build container (reads /root/reference/docs as training text; needs `sentencepiece` and
`tokenizers`).  Outputs, all committed:
  tests/golden/vocab_spm.json   SentencePiece BPE model (byte fallback, llama-style) -> tokenizer.ggml.model="llama"
  tests/golden/vocab_bpe.json   byte-level BPE with the llama-3 pre-tokeniser regex    -> "gpt2" / pre "llama-bpe"
  tests/golden/tok_spm.json     {text, ids} pairs produced by the sentencepiece library itself
  tests/golden/tok_bpe.json     {text, ids} pairs produced by the HF tokenizers library itself
make_gguf.py pads these vocabularies with unused tokens up to the model's n_vocab (32 000 / 128 256).
There are no tokeniser known-answer tests in the reference; these goldens come from the two
upstream libraries llama.cpp's tokenisers are modelled on.
"""
import glob
import io
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = os.path.join(HERE, "..", "tests", "golden")
LLAMA3_RE = ("(?i:'s|'t|'re|'ve|'m|'ll|'d)|[^\\r\\n\\p{L}\\p{N}]?\\p{L}+|\\p{N}{1,3}| ?[^\\s\\p{L}\\p{N}]+[\\r\\n]*|"
             "\\s*[\\r\\n]+|\\s+(?!\\S)|\\s+")

SAMPLES = [
    "Hello world", " Hello world", "Hello  world  ", "", " ", "  ", "\n", "\n\n", " \n \n", "\t\ttabs\tand spaces  \n",
    "The quick brown fox jumps over the lazy dog.", "I'm sure they've said we'll go; he'd know, it's DON'T and I'M.",
    "numbers 1 12 123 1234 12345 3.14159 1,000,000", "snake_case camelCase kebab-case PATH=/usr/local/bin:$PATH",
    "naïve café résumé coöperate", "日本語のテキストと中文文本", "Привет, мир! Ελληνικά עברית العربية",
    "emoji 🙂🚀👨‍👩‍👧‍👦 done", "x = {'a': [1, 2, 3], \"b\": None}  # comment", "```yaml\nname: gpt-4\nbackend: llama-cpp\n```",
    "trailing space ", "multiple\r\nline\r\nendings\r\n", "a" * 50, "ﬁ ligature and   nbsp   em-space",
    "<s>literal</s> <|begin_of_text|>special<|eot_id|> tokens <unk>",
]


def corpus():
    files = sorted(glob.glob("/root/reference/docs/content/**/*.md", recursive=True)) + ["/root/reference/README.md"]
    text = []
    for f in files:
        with open(f, encoding="utf-8", errors="replace") as fh:
            text.append(fh.read())
    return "\n".join(text)


def train_spm(text, vocab_size=4096):
    import sentencepiece as spm

    model = io.BytesIO()
    spm.SentencePieceTrainer.train(
        sentence_iterator=iter(text.split("\n")), model_writer=model, model_type="bpe", vocab_size=vocab_size,
        byte_fallback=True, normalization_rule_name="identity", remove_extra_whitespaces=False, add_dummy_prefix=True,
        split_digits=True, allow_whitespace_only_pieces=True, character_coverage=0.9995, unk_id=0, bos_id=1, eos_id=2,
        pad_id=-1, num_threads=1, train_extremely_large_corpus=False, max_sentence_length=16384)
    sp = spm.SentencePieceProcessor(model_proto=model.getvalue())
    n = sp.get_piece_size()
    tokens, scores, types = [], [], []
    for i in range(n):
        tokens.append(sp.id_to_piece(i))
        scores.append(float(sp.get_score(i)))
        if sp.is_unknown(i):
            types.append(2)
        elif sp.is_control(i):
            types.append(3)
        elif sp.is_byte(i):
            types.append(6)
        elif sp.is_unused(i):
            types.append(5)
        else:
            types.append(1)
    vocab = dict(model="llama", pre="default", tokens=tokens, scores=scores, types=types, merges=[], bos=1, eos=2, unk=0,
                 add_bos=True)
    return vocab, sp


def train_bpe(text, vocab_size=4096):
    from tokenizers import Regex, Tokenizer, decoders, models, pre_tokenizers, trainers

    tok = Tokenizer(models.BPE(ignore_merges=True))
    tok.pre_tokenizer = pre_tokenizers.Sequence([
        pre_tokenizers.Split(Regex(LLAMA3_RE), behavior="isolated", invert=False),
        pre_tokenizers.ByteLevel(add_prefix_space=False, use_regex=False)])
    tok.decoder = decoders.ByteLevel()
    specials = ["<|begin_of_text|>", "<|end_of_text|>", "<|eot_id|>"]
    trainer = trainers.BpeTrainer(vocab_size=vocab_size - len(specials), initial_alphabet=pre_tokenizers.ByteLevel.alphabet(),
                                  special_tokens=[], show_progress=False)
    tok.train_from_iterator(text.split("\n"), trainer)
    tok.add_special_tokens(specials)
    js = json.loads(tok.to_str())
    vocab_map = js["model"]["vocab"]
    merges = [m if isinstance(m, str) else " ".join(m) for m in js["model"]["merges"]]
    n = tok.get_vocab_size(with_added_tokens=True)
    tokens = [None] * n
    for t, i in vocab_map.items():
        tokens[i] = t
    for at in js["added_tokens"]:
        tokens[at["id"]] = at["content"]
    assert all(t is not None for t in tokens)
    types = [3 if t in specials else 1 for t in tokens]
    vocab = dict(model="gpt2", pre="llama-bpe", tokens=tokens, scores=[0.0] * n, types=types, merges=merges,
                 bos=tokens.index(specials[0]), eos=tokens.index(specials[1]), eot=tokens.index(specials[2]), unk=-1,
                 add_bos=True)
    return vocab, tok


def main():
    text = corpus()
    os.makedirs(GOLD, exist_ok=True)
    paragraphs = [p for p in text.split("\n\n") if 20 < len(p) < 600]
    step = max(1, len(paragraphs) // 60)
    samples = SAMPLES + paragraphs[::step][:60]

    spm_vocab, sp = train_spm(text)
    with open(os.path.join(GOLD, "vocab_spm.json"), "w", encoding="utf-8") as f:
        json.dump(spm_vocab, f, ensure_ascii=False)
    # sentencepiece treats "<s>" etc. in raw text as plain text; the engine is called with parse_special=False here
    gold = [dict(text=s, ids=[1] + sp.encode(s)) for s in samples]
    with open(os.path.join(GOLD, "tok_spm.json"), "w", encoding="utf-8") as f:
        json.dump(gold, f, ensure_ascii=False)

    bpe_vocab, tok = train_bpe(text)
    with open(os.path.join(GOLD, "vocab_bpe.json"), "w", encoding="utf-8") as f:
        json.dump(bpe_vocab, f, ensure_ascii=False)
    gold = [dict(text=s, ids=[bpe_vocab["bos"]] + tok.encode(s, add_special_tokens=False).ids) for s in samples]
    with open(os.path.join(GOLD, "tok_bpe.json"), "w", encoding="utf-8") as f:
        json.dump(gold, f, ensure_ascii=False)
    print("spm vocab", len(spm_vocab["tokens"]), "bpe vocab", len(bpe_vocab["tokens"]), "merges", len(bpe_vocab["merges"]),
          "samples", len(samples))


if __name__ == "__main__":
    sys.exit(main())
