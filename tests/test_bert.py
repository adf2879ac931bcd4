"""BERT embeddings backend (csrc/bert.cu; reference: backend/go/llm/bert/bert.go -> go-bert.cpp -> bert.cpp bert_eval).

CPU part: the WordPiece tokeniser through the C ABI vs the restatement in oracle/bert_ref.py, and error behaviour.
GPU part: hidden states and embeddings vs the float32 numpy restatement, plus the behaviour the reference's own test
pins (core/http/app_test.go:856-881: vector length, same input -> same vector).

Tolerance: the mat-muls run on fp16 operand tiles with fp32 accumulation (the file's F16 weights exactly; activations
rounded to fp16), everything else is fp32: after every layer the rows are re-normalised, so the operand rounding shows as an
absolute error on unit-variance values: max <= 8 % of the standard deviation, median <= 1.2 %, cosine(embedding) >= 0.9995.
The default attention kernel also rounds Q, K, P and V to bf16 for its MMAs (the usual flash-attention numerics); the
scores are exponentiated, so its worst element may be off by up to 15 % of the standard deviation -- the median bound
and the cosine bound are the same for both kernels."""
import os
import sys

import numpy as np
import pytest

from localai_b200 import capi
from oracle.bert_ref import BertRef

sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "tools"))

TEXTS = ["The quick brown fox, jumping over 2 lazy dogs! 日本語 xyzzy café", "", "   ", "Hello,World!!", "ÀB ünïcode ß", "a" * 700,
         "tokenization embeddings unusedword \x7f ", "LocalAI is a drop-in replacement REST API", "[CLS] is just text here"]


@pytest.fixture(scope="module")
def bert_files(tmp_path_factory):
    import make_bert_gguf as mb

    d = tmp_path_factory.mktemp("bert")
    return {n: mb.write_bert_gguf(str(d / f"{n}.gguf"), mb.BERT_ARCHS[n], seed=3) for n in ("bert-mini", "minilm-l3")}


def test_wordpiece_matches_restatement(bert_files):
    ref = BertRef(bert_files["bert-mini"])
    with capi.Bert(bert_files["bert-mini"], vocab_only=True) as b:
        assert b.info["n_embd"] == 256 and b.info["n_ctx"] == 512
        for t in TEXTS:
            got = b.tokenize(t)
            assert got == ref.tokenize(t), t
            assert got[0] == ref.cls and got[-1] == ref.sep and len(got) <= 512
        assert b.tokenize("the fox") == [ref.cls, ref.tok_id["▁the"], ref.tok_id["▁fox"], ref.sep]
        assert b.tokenize("foxes") == [ref.cls, ref.tok_id["▁fox"], ref.tok_id["e"], ref.tok_id["s"], ref.sep]
        assert b.tokenize("\x7f") == [ref.cls, ref.unk, ref.sep]  # no piece: the whole word is [UNK]
        with pytest.raises(RuntimeError, match="vocab_only"):
            b.embed("x")


def test_wordpiece_matches_hf_tokenizers(bert_files):
    """The WordPiece tokeniser (and its restatement) against the HF `tokenizers` library's: lower-casing normaliser, split on
    white space and punctuation, greedy longest match with [UNK] for a word with an unmatched rest.  The GGUF vocabulary
    marks word starts ("▁the") where BERT's marks continuations ("##e"); converted, the ids must be the same -- on ASCII
    text, where bert.cpp's ASCII-only case folding and punctuation table say what the full Unicode ones say."""
    tk = pytest.importorskip("tokenizers")
    ref = BertRef(bert_files["bert-mini"])
    vocab = {}
    for i, s in enumerate(ref.tokens):
        hf = s if (s.startswith("[") and s.endswith("]")) else (s[1:] if s.startswith("▁") else "##" + s)
        vocab.setdefault(hf, i)
    tok = tk.Tokenizer(tk.models.WordPiece(vocab=vocab, unk_token="[UNK]", max_input_chars_per_word=10000))
    tok.normalizer = tk.normalizers.BertNormalizer(clean_text=False, handle_chinese_chars=True, strip_accents=False, lowercase=True)
    tok.pre_tokenizer = tk.pre_tokenizers.BertPreTokenizer()
    texts = ["The quick brown fox, jumping over 2 lazy dogs! xyzzy", "Hello,World!!", "tokenization embeddings unusedword", "",
             "LocalAI is a drop-in replacement REST API", "[CLS] is just text here", "a" * 90, "  spaces\tand\nlines  ",
             "model's vector-search (query) & documents: 100% local_ai", "UPPER lower MiXeD 0123456789 ~`^|<>"]
    with capi.Bert(bert_files["bert-mini"], vocab_only=True) as b:
        for t in texts:
            want = [ref.cls] + tok.encode(t, add_special_tokens=False).ids + [ref.sep]
            assert ref.tokenize(t) == want, t
            assert b.tokenize(t) == want, t


def test_load_errors(model_files):
    with pytest.raises(RuntimeError, match="bert"):
        capi.Bert(model_files["micro_q8_0"], vocab_only=True)  # a llama GGUF is not a BERT encoder
    with pytest.raises(RuntimeError):
        capi.Bert("/nonexistent.gguf", vocab_only=True)


@pytest.mark.gpu
@pytest.mark.parametrize("att", ["mma", "fp32"])
@pytest.mark.parametrize("name", ["bert-mini", "minilm-l3"])
def test_hidden_states_and_embedding_match_restatement(bert_files, name, att, monkeypatch):
    """att: the tensor-core attention kernel (default) and the fp32 CUDA-core one kept beside it for A/B checks"""
    monkeypatch.setenv("B200LLM_BERT_ATT_SIMPLE", "1" if att == "fp32" else "0")
    monkeypatch.setenv("B200LLM_BERT_UNFUSED_LN", "1" if att == "fp32" else "0")  # and the LayerNorm kernel for wide rows
    ref = BertRef(bert_files[name])
    with capi.Bert(bert_files[name]) as b:
        for text in (TEXTS[0], TEXTS[7], "a" * 700, "x"):
            ids = ref.tokenize(text)
            want = ref.embed_tokens(ids, return_hidden=True)
            got = b.hidden(ids)
            err = np.abs(got - want)
            print(f"{name} {att} n={len(ids)}: max err {err.max():.4f}, median {np.median(err):.5f}, std {want.std():.3f}")
            max_tol = 0.15 if att == "mma" else 0.08
            assert err.max() <= max_tol * want.std() and np.median(err) <= 0.012 * want.std(), (name, len(ids), err.max(), want.std())
            e, we = b.embed_tokens(ids), ref.embed_tokens(ids)
            assert e.shape == (ref.n_embd,) and abs(np.linalg.norm(e) - 1.0) < 1e-5
            assert float(e @ we) >= 0.9995, float(e @ we)
            assert np.array_equal(e, b.embed(text))  # Embeddings(text) = TokenEmbeddings(tokenize(text))


@pytest.mark.gpu
def test_same_input_same_vector_and_batch_equals_single(bert_files):
    ref = BertRef(bert_files["minilm-l3"])
    with capi.Bert(bert_files["minilm-l3"]) as b:
        a1, a2 = b.embed("doc one"), b.embed("doc one")
        assert a1.shape == (384,) and np.array_equal(a1, a2)  # app_test.go:856-881
        assert not np.allclose(a1, b.embed("another document"), atol=1e-3)
        seqs = [ref.tokenize(t) for t in TEXTS]
        batch = b.embed_batch(seqs)  # ragged lengths 2..512 packed into one forward
        for s, row in zip(seqs, batch):
            single = b.embed_tokens(s)
            # same kernels, but the GEMM sees another row count, so its split-K slices and with them the fp32 summation
            # order differ; a last-bit difference can flip the fp16 rounding of an activation before the next GEMM
            # (2^-11 relative on that element), which is what remains: 1e-3 on unit vectors, cosine 1 - 1e-5
            assert np.abs(row - single).max() <= 1e-3 and float(row @ single) >= 1.0 - 1e-5
        with pytest.raises(RuntimeError, match="tokens"):
            b.embed_tokens(list(range(5)) * 200)  # 1000 tokens > context
        with pytest.raises(RuntimeError, match="out of range"):
            b.embed_tokens([1, 999999])


def test_engine_face_on_a_bert_gguf_vocab_only(bert_files):
    """b2_create on a BERT GGUF (what the llama-cpp backend is given with `embeddings: true`): metadata + WordPiece"""
    ref = BertRef(bert_files["bert-mini"])
    with capi.Engine(bert_files["bert-mini"], vocab_only=True) as e:
        assert e.info.arch == b"bert" and e.info.n_embd == 256 and e.info.tokenizer_model == b"bert"
        assert e.tokenize("The quick brown fox!") == ref.tokenize("The quick brown fox!")
        assert e.token_to_piece(ref.tok_id["▁fox"]) == "▁fox".encode()


@pytest.mark.gpu
def test_engine_embedding_request_on_a_bert_gguf(bert_files):
    """Embedding through b2_submit / b2_poll / b2_embed_result (grpc-server.cpp:2410-2432) = the bert backend's vector;
    anything but an Embedding request is refused"""
    text = "LocalAI serves embeddings"
    with capi.Bert(bert_files["minilm-l3"]) as b:
        want = b.embed(text)
    with capi.Engine(bert_files["minilm-l3"], embeddings=True) as e:
        rid = e.submit(prompt=text, embedding_mode=True)
        evs = list(e.events(rid))
        assert len(evs) == 1 and evs[0]["flags"] & capi.EV_FINAL and evs[0]["flags"] & capi.EV_EMBEDDING and evs[0]["n_embd"] == 384
        vec = e.embed_result(rid)
        e.release(rid)
        assert np.array_equal(vec, want)
        rid = e.submit(prompt_ids=e.tokenize(text), embedding_mode=True)
        list(e.events(rid))
        assert np.array_equal(e.embed_result(rid), want)
        with pytest.raises(RuntimeError, match="Embedding requests only"):
            e.submit(prompt=text, n_predict=4)
        with pytest.raises(RuntimeError, match="decoder"):
            e.eval([1, 2, 3])
    with capi.Engine(bert_files["minilm-l3"], embeddings=False) as e:
        rid = e.submit(prompt=text, embedding_mode=True)
        list(e.events(rid))
        assert not e.embed_result(rid).any()  # embeddings disabled -> zeros (grpc-server.cpp:1301-1310)


@pytest.mark.gpu
def test_fp16_attention_path_on_ragged_batches(bert_files, monkeypatch):
    """heads of 64 and more than one token tile: q | k | v leave the projection GEMM as fp16 rows with their biases and
    the attention kernel writes attn_output's operand tiles.  Against the restatement, and against the fp32-row path."""
    ref = BertRef(bert_files["bert-mini"])
    rng = np.random.default_rng(11)
    seqs = [[2] + rng.integers(5, 2000, n - 2).tolist() + [3] for n in (512, 3, 129, 64, 200, 511, 65, 128, 17, 300)]
    with capi.Bert(bert_files["bert-mini"]) as b:
        batch = b.embed_batch(seqs)
        for s, row in zip(seqs, batch):
            want = ref.embed_tokens(s)
            assert abs(np.linalg.norm(row) - 1.0) < 1e-5 and float(row @ want) >= 0.9995, (len(s), float(row @ want))
        again = b.embed_batch(seqs)
        assert np.array_equal(batch, again)
        h = b.hidden(seqs[4])  # one text of 200 tokens: two token tiles, the same path
        want = ref.embed_tokens(seqs[4], return_hidden=True)
        err = np.abs(h - want)
        assert err.max() <= 0.15 * want.std() and np.median(err) <= 0.012 * want.std()
        monkeypatch.setenv("B200LLM_BERT_ATT_F32", "1")
        other = b.embed_batch(seqs)
        assert not np.array_equal(batch, other)  # another kernel, another rounding ...
        assert np.abs(batch - other).max() <= 3e-3  # ... of the same numbers (unit vectors)


@pytest.mark.gpu
def test_single_text_graph_replay_equals_plain_launches(bert_files, monkeypatch):
    """one text per call: first call plain, second captures a CUDA graph for the 32-token length bucket, later ones
    replay it; padding rows must not leak into the vector"""
    rng = np.random.default_rng(4)
    seqs = [[2] + rng.integers(5, 200, n - 2).tolist() + [3] for n in (3, 31, 32, 33, 100, 512)]
    with capi.Bert(bert_files["bert-mini"]) as b:
        first = [b.embed_tokens(s) for s in seqs]      # plain launches (lengths 31 and 32 share a bucket: 32 captures)
        second = [b.embed_tokens(s) for s in seqs]     # capture + launch
        third = [b.embed_tokens(s) for s in reversed(seqs)][::-1]  # replay, other order
        for a, c, d in zip(first, second, third):
            assert np.array_equal(a, c) and np.array_equal(a, d)
        monkeypatch.setenv("B200LLM_BERT_NO_GRAPH", "1")  # no padding, no graph
        for s, a in zip(seqs, first):
            assert np.abs(b.embed_tokens(s) - a).max() <= 1e-6
