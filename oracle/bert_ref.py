"""TEST INFRASTRUCTURE ONLY -- numpy float32 restatement of the BERT embeddings path:
`bert.New` / `Embeddings` / `TokenEmbeddings` (backend/go/llm/bert/bert.go:18,30,33) -> go-bert.cpp @ 710044b ->
bert.cpp `bert_eval` (embeddings sum -> LayerNorm -> N x {QKV + bias, softmax(QK^T / sqrt(hd)) V, output dense + bias,
residual + LayerNorm, dense + bias, GELU, dense + bias, residual + LayerNorm} -> mean over tokens -> L2 normalise), and
llama.cpp's build_bert / WordPiece tokeniser for the same model stored as GGUF (general.architecture = "bert").

Parity: UNPINNED -- neither go-bert.cpp nor llama.cpp is in the reference tree and no model file can be fetched; the
reference's own test at this boundary is behavioural (core/http/app_test.go:856-881: vector length, same input ->
same vector).  The product's implementation is localai_b200/csrc/bert.cu; nothing here is imported by it.
"""
from __future__ import annotations

import numpy as np


class BertRef:
    def __init__(self, path: str):
        import gguf

        r = gguf.GGUFReader(path)
        self.t = {}
        for t in r.tensors:
            w = gguf.quants.dequantize(t.data, t.tensor_type).astype(np.float32)
            self.t[t.name] = w.reshape([int(x) for x in reversed(t.shape)])

        def kv(key):
            f = r.fields[key]
            return f.parts[f.data[0]]

        self.n_layer = int(kv("bert.block_count")[0])
        self.n_embd = int(kv("bert.embedding_length")[0])
        self.n_head = int(kv("bert.attention.head_count")[0])
        self.eps = float(kv("bert.attention.layer_norm_epsilon")[0])
        self.n_ctx = int(kv("bert.context_length")[0])
        f = r.fields["tokenizer.ggml.tokens"]
        self.tokens = [bytes(f.parts[i]).decode("utf-8") for i in f.data]
        self.tok_id = {}
        for i, s in enumerate(self.tokens):
            self.tok_id.setdefault(s, i)
        self.cls, self.sep, self.unk = int(kv("tokenizer.ggml.bos_token_id")[0]), int(kv("tokenizer.ggml.eos_token_id")[0]), int(
            kv("tokenizer.ggml.unknown_token_id")[0])

    # ---- WordPiece (llama.cpp llm_tokenizer_wpm: lower-case, split on white space, punctuation and CJK characters
    # stand alone, greedy longest match with a leading "▁" on the first piece of a word; a word with an unmatched rest is
    # [UNK]); [CLS] ... [SEP] around it
    @staticmethod
    def _is_punct(ch: str) -> bool:
        c = ord(ch)
        return (33 <= c <= 47) or (58 <= c <= 64) or (91 <= c <= 96) or (123 <= c <= 126)

    @staticmethod
    def _is_cjk(ch: str) -> bool:
        c = ord(ch)
        return (0x4E00 <= c <= 0x9FFF) or (0x3400 <= c <= 0x4DBF) or (0x3040 <= c <= 0x30FF) or (0xF900 <= c <= 0xFAFF)

    def words(self, text: str):
        out, cur = [], ""
        for ch in text:
            if ch.isascii() and ch.isupper():
                ch = ch.lower()
            if ch in " \t\n\r":
                if cur:
                    out.append(cur)
                cur = ""
            elif self._is_punct(ch) or self._is_cjk(ch):
                if cur:
                    out.append(cur)
                out.append(ch)
                cur = ""
            else:
                cur += ch
        if cur:
            out.append(cur)
        return out

    def tokenize(self, text: str, max_tokens: int | None = None):
        ids = [self.cls]
        for w in self.words(text):
            piece_ids, i, ok = [], 0, True
            while i < len(w):
                j = len(w)
                while j > i:
                    cand = ("▁" if i == 0 else "") + w[i:j]
                    if cand in self.tok_id:
                        break
                    j -= 1
                if j == i:
                    ok = False
                    break
                piece_ids.append(self.tok_id[cand])
                i = j
            ids += piece_ids if ok else [self.unk]
        ids.append(self.sep)
        n = max_tokens or self.n_ctx
        if len(ids) > n:  # bert.cpp truncates to the model's maximum; the closing [SEP] is kept
            ids = ids[: n - 1] + [self.sep]
        return ids

    # ---- forward
    def _ln(self, x, w, b):
        mu = x.mean(axis=-1, keepdims=True, dtype=np.float32)
        xc = x - mu
        var = (xc * xc).mean(axis=-1, keepdims=True, dtype=np.float32)
        return (xc / np.sqrt(var + np.float32(self.eps))) * w + b

    @staticmethod
    def _gelu(x):
        return (0.5 * x * (1.0 + np.tanh(np.float32(0.7978845608028654) * x * (1.0 + np.float32(0.044715) * x * x)))).astype(np.float32)

    def embed_tokens(self, ids, normalize: bool = True, return_hidden: bool = False):
        t = self.t
        ids = np.asarray(ids, dtype=np.int64)
        n, d, h = ids.size, self.n_embd, self.n_head
        hd = d // h
        x = t["token_embd.weight"][ids] + t["token_types.weight"][0] + t["position_embd.weight"][:n]
        x = self._ln(x.astype(np.float32), t["token_embd_norm.weight"], t["token_embd_norm.bias"])
        for i in range(self.n_layer):
            p = f"blk.{i}."
            q = x @ t[p + "attn_q.weight"].T + t[p + "attn_q.bias"]
            k = x @ t[p + "attn_k.weight"].T + t[p + "attn_k.bias"]
            v = x @ t[p + "attn_v.weight"].T + t[p + "attn_v.bias"]
            q, k, v = (a.reshape(n, h, hd).transpose(1, 0, 2) for a in (q, k, v))
            s = (q @ k.transpose(0, 2, 1)) / np.float32(np.sqrt(hd))
            s = s - s.max(axis=-1, keepdims=True)
            pr = np.exp(s)
            pr = pr / pr.sum(axis=-1, keepdims=True)
            att = (pr @ v).transpose(1, 0, 2).reshape(n, d).astype(np.float32)
            o = att @ t[p + "attn_output.weight"].T + t[p + "attn_output.bias"]
            x = self._ln(x + o, t[p + "attn_output_norm.weight"], t[p + "attn_output_norm.bias"]).astype(np.float32)
            u = self._gelu((x @ t[p + "ffn_up.weight"].T + t[p + "ffn_up.bias"]).astype(np.float32))
            dn = u @ t[p + "ffn_down.weight"].T + t[p + "ffn_down.bias"]
            x = self._ln(x + dn, t[p + "layer_output_norm.weight"], t[p + "layer_output_norm.bias"]).astype(np.float32)
        if return_hidden:
            return x
        e = x.mean(axis=0, dtype=np.float32)
        if normalize:
            e = e / np.sqrt((e * e).sum(dtype=np.float32))
        return e.astype(np.float32)

    def embed(self, text: str):
        return self.embed_tokens(self.tokenize(text))
