"""CPU-side tests of the product's host logic: the C-ABI library loads and exports every symbol declared in
include/b200llm.h, GGUF parsing + tokenisers work without a GPU (vocab_only), and compute entry points fail
loudly (no CPU fallback) when there is no sm_100 device."""
import ctypes
import json
import os
import re

import numpy as np
import pytest

from conftest import have_gpu
from localai_b200 import capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = os.path.join(ROOT, "tests", "golden")


def header_functions():
    src = open(os.path.join(ROOT, "include", "b200llm.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(b2_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    L = ctypes.CDLL(capi.LIB_PATH)
    names = header_functions()
    assert len(names) >= 20
    for n in names:
        assert hasattr(L, n), f"libb200llm.so does not export {n}"
    assert sorted(capi.EXPORTS) == names, "capi.EXPORTS out of sync with include/b200llm.h"
    assert capi.lib().b2_abi_version() == 1


def test_struct_sizes_match_header():
    # a few layout invariants the cgo / ctypes bindings rely on
    assert ctypes.sizeof(capi.LoadOpts) == 80
    assert ctypes.sizeof(capi.Event) == 72
    assert ctypes.sizeof(capi.Metrics) == 48


def test_vocab_only_reads_metadata(model_files):
    with capi.Engine(model_files["mini_q4_k_m"], vocab_only=True) as e:
        i = e.info
        assert (i.n_layer, i.n_embd, i.n_head, i.n_head_kv, i.n_ff, i.n_vocab) == (3, 512, 8, 2, 1024, 4608)
        assert i.arch == b"llama" and i.tokenizer_model == b"gpt2"
        with pytest.raises(capi.B2Error):
            e.eval([1, 2, 3])
        with pytest.raises(capi.B2Error):
            e.submit("hi")


def test_other_architectures_are_recognised(model_files):
    """qwen2 and phi3 GGUFs load (metadata keys carry the architecture's prefix; phi3 has no ffn_gate / attn_q tensors)"""
    for name, arch, tok in (("mini_qwen2", b"qwen2", b"gpt2"), ("mini_phi3", b"phi3", b"llama")):
        with capi.Engine(model_files[name], vocab_only=True) as e:
            i = e.info
            assert i.arch == arch and i.tokenizer_model == tok
            assert (i.n_layer, i.n_embd, i.n_head, i.n_head_kv, i.n_ff) == (3, 512, 8, 2, 1024)


def test_bad_files_are_rejected(tmp_path):
    p = tmp_path / "junk.gguf"
    p.write_bytes(b"not a gguf file at all, definitely" * 4)
    with pytest.raises(capi.B2Error, match="magic"):
        capi.Engine(str(p), vocab_only=True)
    with pytest.raises(capi.B2Error, match="cannot open"):
        capi.Engine(str(tmp_path / "missing.gguf"), vocab_only=True)


@pytest.mark.parametrize("kind,model,parse_special", [("spm", "micro_mixed", False), ("bpe", "mini_q4_k_m", True)])
def test_tokenizer_matches_upstream_library_goldens(model_files, kind, model, parse_special):
    gold = json.load(open(os.path.join(GOLD, f"tok_{kind}.json"), encoding="utf-8"))
    with capi.Engine(model_files[model], vocab_only=True) as e:
        for g in gold:
            ids = e.tokenize(g["text"], add_bos=True, parse_special=parse_special)
            assert ids == g["ids"], g["text"]


def test_qwen2_pretokenizer_matches_upstream_library_goldens(model_files):
    """One digit per word, no ignore_merges, no BOS (tools/make_qwen2_tok_goldens.py; ids from the HF tokenizers library)."""
    gold = json.load(open(os.path.join(GOLD, "tok_bpe_qwen2.json"), encoding="utf-8"))
    with capi.Engine(model_files["mini_qwen2"], vocab_only=True) as e:
        assert e.info.arch == b"qwen2"
        for g in gold:
            assert e.tokenize(g["text"], add_bos=True, parse_special=False) == g["ids"], g["text"]


@pytest.mark.parametrize("model", ["micro_mixed", "mini_q4_k_m"])
def test_detokenize_round_trip(model_files, model):
    texts = ["Hello world", "naïve café 日本語 🙂", "tabs\tand\nnewlines  ", "x = {'a': 1}"]
    with capi.Engine(model_files[model], vocab_only=True) as e:
        for t in texts:
            ids = e.tokenize(t, add_bos=False, parse_special=False)
            dec = b"".join(e.token_to_piece(i) for i in ids)
            if e.info.tokenizer_model == b"llama":
                assert dec == b" " + t.encode()  # SPM adds the dummy space prefix
            else:
                assert dec == t.encode()


def test_special_tokens_parsed_only_on_request(model_files):
    with capi.Engine(model_files["mini_q4_k_m"], vocab_only=True) as e:
        eot = e.info.eot_id
        assert eot > 0
        assert eot in e.tokenize("a<|eot_id|>b", add_bos=False, parse_special=True)
        assert eot not in e.tokenize("a<|eot_id|>b", add_bos=False, parse_special=False)
        assert e.token_to_piece(eot) == b"<|eot_id|>"
        assert e.tokenize("", add_bos=True) == [e.info.bos_id]
        assert e.tokenize("", add_bos=False) == []


@pytest.mark.skipif(have_gpu(), reason="checks the no-GPU failure mode")
def test_compute_fails_loudly_without_gpu(model_files):
    with pytest.raises(capi.B2Error, match="no CUDA device|no sm_100"):
        capi.Engine(model_files["micro_q8_0"])
    with pytest.raises(capi.B2Error, match="no sm_100|no CPU fallback"):
        capi.dequantize(capi.B2_T_Q4_K, np.zeros(144, dtype=np.uint8), 256)
    with pytest.raises(capi.B2Error):
        capi.mul_mat(capi.B2_T_Q8_0, np.zeros(34 * 8 * 4, dtype=np.uint8), 4, 256, np.zeros((1, 256), dtype=np.float32))


def test_gguf_writer_round_trips_through_gguf_py(model_files):
    import gguf
    r = gguf.GGUFReader(model_files["mini_q4_k_m"])
    names = [t.name for t in r.tensors]
    assert names[0] == "token_embd.weight" and names[-1] == "output.weight" and len(names) == 3 * 9 + 3
    types = {t.name: t.tensor_type.name for t in r.tensors}
    assert types["output.weight"] == "Q6_K" and types["blk.0.attn_q.weight"] == "Q4_K"
    assert types["blk.2.ffn_down.weight"] == "Q6_K" and types["blk.0.ffn_down.weight"] == "Q4_K"  # use_more_bits(i, 3)
    w = gguf.quants.dequantize(r.tensors[2].data, r.tensors[2].tensor_type)
    assert np.isfinite(w).all() and 0.5 / np.sqrt(512) < w.std() < 2 / np.sqrt(512)


def test_q4_k_m_type_mix_reproduces_known_file_sizes():
    import make_gguf as mg
    for arch, preset, gb in [("llama3-8b", "q4_k_m", 4.91), ("llama3-70b", "q4_k_m", 42.5), ("tinyllama-1.1b", "q8_0", 1.17)]:
        plan = mg.tensor_plan(mg.ARCHS[arch], preset, 4.0)
        total = sum(mg.tensor_nbytes(r, c, t) for _, r, c, t, _, _ in plan)
        assert abs(total / 1e9 - gb) < 0.01 * gb + 0.01


# ---- the decode kernel's weight-stream planner (host arithmetic, csrc/kernels_mega.cu: mega_plan_streams) ----
Q4_K, Q5_K, Q6_K, Q8_0 = 12, 13, 14, 8


def _llama_phases(n_layer, d, ff, kvd, vocab, t_main, t_v, t_down, t_out):
    """9 int64 per phase {kind, K, n_jobs, type0, rows0, type1, rows1, type2, rows2}: QKV, attention, wo, gate|up, down per
    layer, then the LM head -- the list Model::build_mega makes."""
    ph = []
    for l in range(n_layer):
        ph += [[0, d, 3, t_main, d, t_main, kvd, t_v[l % len(t_v)], kvd], [1, 0, 0, 0, 0, 0, 0, 0, 0], [2, d, 1, t_main, d, 0, 0, 0, 0],
               [3, d, 2, t_main, ff, t_main, ff, 0, 0], [2, ff, 1, t_down[l % len(t_down)], d, 0, 0, 0, 0]]
    ph.append([2, d, 1, t_out, vocab, 0, 0, 0, 0])
    return np.array(ph, dtype=np.int64)


@pytest.mark.parametrize("name,args,expect_bytes", [
    ("llama3-8b q4_k_m", (32, 4096, 14336, 1024, 128256, Q4_K, [Q6_K, Q4_K], [Q6_K, Q4_K], Q6_K), 4.6e9),
    ("tinyllama q8_0", (22, 2048, 5632, 256, 32000, Q8_0, [Q8_0], [Q8_0], Q8_0), 1.1e9),
    ("llama3-70b q4_k_m, 4 layers", (4, 8192, 28672, 1024, 128256, Q4_K, [Q6_K, Q5_K], [Q6_K, Q4_K], Q6_K), 2.9e9),
    ("mini", (3, 512, 1024, 128, 4608, Q4_K, [Q6_K], [Q6_K], Q6_K), None),
])
@pytest.mark.parametrize("n_cta", [148, 132, 16])
def test_mega_stream_plan_is_consistent(name, args, expect_bytes, n_cta):
    """Every (row, K-split) of every phase appears in exactly one tile of exactly one warp's stream, the tiles of a stream
    chain through their headers (size of the tile two ahead, last-of-phase flag, next phase) the way the kernel walks
    them, and the streams hold the weights' bytes plus 16 per tile.  The check itself is mega_plan_verify in the library."""
    L = capi.lib()
    L.b2_mega_plan_check.restype = ctypes.c_int64
    L.b2_mega_plan_check.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_void_p]
    d = _llama_phases(*args)
    n_tiles = ctypes.c_int64(0)
    r = L.b2_mega_plan_check(d.ctypes.data, len(d), n_cta, ctypes.byref(n_tiles))
    if r == capi.B2_EUNSUPPORTED:  # e.g. more rows x K-splits per CTA than the partial-sum buffer holds at a small grid
        pytest.skip((L.b2_last_error() or b"").decode())
    assert r > 0, (L.b2_last_error() or b"").decode()
    assert n_tiles.value > 0
    if expect_bytes and n_cta == 148:
        assert abs(r - expect_bytes) < 0.02 * expect_bytes


@pytest.mark.parametrize("n_tiles,nkb", [
    (48, 16), (32, 16), (224, 16), (32, 56), (1002, 16),      # llama3-8b: wq|wk|wv, attn_output, gate|up, ffn_down, output
    (80, 32), (64, 32), (448, 32), (64, 112),                  # llama3-70b
    (1, 1), (3, 7), (148, 1), (149, 1), (2, 300), (7, 4)])
@pytest.mark.parametrize("n_sm", [148, 132, 8])
def test_qgemm_work_split_is_consistent(n_tiles, nkb, n_sm):
    """The batched-decode GEMM cuts its (panel, super-block) units into equal contiguous CTA ranges; every unit belongs to
    exactly one range, the epilogue's arithmetic for "first CTA / number of segments of this panel" matches the ranges,
    and no CTA needs more than its two partial-tile slots.  The check itself is qgemm_plan_check in the library."""
    L = capi.lib()
    L.b2_qgemm_plan_check.restype = ctypes.c_int64
    L.b2_qgemm_plan_check.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_int]
    r = L.b2_qgemm_plan_check(n_tiles, nkb, n_sm)
    assert r >= 0, (L.b2_last_error() or b"").decode()
    assert r <= n_tiles


def _stream(pieces, stops):
    L = capi.lib()
    buf = b"".join(pieces)
    lens = (ctypes.c_int * len(pieces))(*[len(p) for p in pieces])
    arr = (ctypes.c_char_p * max(1, len(stops)))(*stops) if stops else (ctypes.c_char_p * 1)()
    s1, s2, used = ctypes.create_string_buffer(4096), ctypes.create_string_buffer(4096), ctypes.c_int(0)
    L.b2_stream_text.argtypes = [ctypes.c_char_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_int, ctypes.c_char_p, ctypes.c_int,
                                 ctypes.c_char_p, ctypes.c_int, ctypes.c_void_p]
    r = L.b2_stream_text(buf, lens, len(pieces), arr, len(stops), s1, 4096, s2, 4096, ctypes.byref(used))
    assert r >= 0, (L.b2_last_error() or b"").decode()
    return s1.value, s2.value, bool(r), used.value


def test_stream_text_holds_back_utf8_and_stop_words():
    """The text rules of the request path (stream_text.h; behaviour of grpc-server.cpp:1010-1123), no device needed: a
    character split over tokens is released whole; the beginning of a possible stop word is held back and released if it
    turns out not to be one; a completed stop word ends the generation, and neither it nor what followed reaches the client
    -- nor, as in the reference, do bytes of the completing token that precede the stop word in a stream."""
    e = "\u20ac".encode()  # 3 bytes
    assert _stream([b"a", e[:1], e[1:2], e[2:], b"b"], []) == (b"a" + e + b"b", b"a" + e + b"b", False, 5)
    assert _stream([b"x", e[:2]], [])[0] == b"x"                                        # still incomplete: nothing of it released
    assert _stream([b"Hello", b" wor", b"ld", b"!"], [b"world"]) == (b"Hello", b"Hello ", True, 3)   # " wor" is held back whole once "wor" may begin the word
    assert _stream([b"Hi", b" ##", b"x", b"y"], [b"###"]) == (b"Hi ##xy", b"Hi ##xy", False, 4)     # held back, then released
    assert _stream([b"Hi", b" ##", b"#", b"z"], [b"###"]) == (b"Hi", b"Hi ", True, 3)
    assert _stream([b"Hel", b"lo STOP rest"], [b"STOP"]) == (b"Hel", b"Hello ", True, 2)            # stream loses "lo ", the final text keeps it
    assert _stream([b"ab", b"cd"], [b"zz", b"cd", b"bc"]) == (b"", b"a", True, 2)                   # "b" may begin "bc": held; the earliest stop position wins


def test_partial_stop_rule_equals_the_references_own_code():
    """Which tail of the text is held back as the possible beginning of a stop word: the product's rule against
    find_partial_stop_string of the reference's server, compiled from the reference's own source (oracle/Makefile builds
    oracle/_ref/libstop_ref.so where /root/reference exists; the .so travels, the source does not)."""
    path = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle", "_ref", "libstop_ref.so")
    if not os.path.exists(path):
        pytest.skip("oracle/_ref/libstop_ref.so was not built (no /root/reference on this machine)")
    R = ctypes.CDLL(path)
    R.ref_find_partial_stop_string.restype = ctypes.c_int64
    R.ref_find_partial_stop_string.argtypes = [ctypes.c_char_p, ctypes.c_size_t, ctypes.c_char_p, ctypes.c_size_t]
    L = capi.lib()
    L.b2_partial_stop.restype = ctypes.c_int64
    L.b2_partial_stop.argtypes = [ctypes.c_char_p, ctypes.c_size_t, ctypes.c_char_p, ctypes.c_size_t]
    rng = np.random.default_rng(31)
    cases = [(b"", b""), (b"abc", b""), (b"", b"abc"), (b"Hello ##", b"###"), (b"say STO", b"STOP"), (b"STOP", b"STOP"),
             (b"aaa", b"aa"), (b"xab", b"abab"), (b"\xc3\xa9", b"\xa9\xc3")]
    for _ in range(4000):
        alphabet = b"ab#\n" if rng.integers(0, 2) else b"abcdefgh \xc3\xa9"
        text = bytes(alphabet[i] for i in rng.integers(0, len(alphabet), int(rng.integers(0, 12))))
        word = bytes(alphabet[i] for i in rng.integers(0, len(alphabet), int(rng.integers(0, 6))))
        cases.append((text, word))
    hits = 0
    for text, word in cases:
        want = R.ref_find_partial_stop_string(word, len(word), text, len(text))
        got = L.b2_partial_stop(text, len(text), word, len(word))
        assert got == want, (text, word, got, want)
        hits += want >= 0
    assert 500 < hits < len(cases) - 500


def test_prefix_reuse_equals_the_references_own_code():
    """How much of a new prompt a slot's cache already holds: common_part of the reference's server (compiled from its source,
    oracle/_ref) followed by its caller's rule that the cache's last token is evaluated again (grpc-server.cpp:1740-1749)."""
    path = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle", "_ref", "libstop_ref.so")
    if not os.path.exists(path):
        pytest.skip("oracle/_ref/libstop_ref.so was not built (no /root/reference on this machine)")
    R = ctypes.CDLL(path)
    R.ref_common_part.restype = ctypes.c_int64
    R.ref_common_part.argtypes = [ctypes.c_void_p, ctypes.c_size_t, ctypes.c_void_p, ctypes.c_size_t]
    L = capi.lib()
    L.b2_reusable_prefix.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_int]
    rng = np.random.default_rng(37)
    seen = set()
    for _ in range(2000):
        base = rng.integers(0, 5, int(rng.integers(0, 24))).astype(np.int32)
        a = np.ascontiguousarray(base[:int(rng.integers(0, len(base) + 1))])
        b = base.copy()
        if len(b) and rng.integers(0, 2):
            b[int(rng.integers(0, len(b)))] ^= 1
        common = R.ref_common_part(a.ctypes.data, len(a), b.ctypes.data, len(b))
        want = common - 1 if (common > 0 and common == len(a)) else common
        assert L.b2_reusable_prefix(a.ctypes.data, len(a), b.ctypes.data, len(b)) == want, (a, b)
        seen.add((common == len(a), common == len(b), common == 0))
    assert len(seen) >= 5


def test_streamed_text_equals_the_references_own_code():
    """The whole text half of process_token (grpc-server.cpp:1027-1096 with find_stopping_strings :976-1008), compiled from
    the reference's source and driven piece by piece, against StreamText on the same pieces and stop words: the bytes a
    streaming client receives, the final text, whether a stop word ended it and after how many tokens -- 5000 random
    sequences of pieces that split UTF-8 characters and stop words every possible way."""
    path = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle", "_ref", "libstop_ref.so")
    if not os.path.exists(path):
        pytest.skip("oracle/_ref/libstop_ref.so was not built (no /root/reference on this machine)")
    R = ctypes.CDLL(path)
    L = capi.lib()
    sig = [ctypes.c_char_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_int, ctypes.c_char_p, ctypes.c_int,
           ctypes.c_char_p, ctypes.c_int, ctypes.c_void_p]
    R.ref_stream_text.argtypes = sig
    L.b2_stream_text.argtypes = sig

    def run(fn, pieces, stops):
        lens = (ctypes.c_int * len(pieces))(*[len(p) for p in pieces])
        arr = (ctypes.c_char_p * max(1, len(stops)))(*stops) if stops else (ctypes.c_char_p * 1)()
        s1, s2, used = ctypes.create_string_buffer(4096), ctypes.create_string_buffer(4096), ctypes.c_int(0)
        r = fn(b"".join(pieces), lens, len(pieces), arr, len(stops), s1, 4096, s2, 4096, ctypes.byref(used))
        assert r >= 0
        return s1.value, s2.value, r, used.value

    rng = np.random.default_rng(47)
    alpha = [b"a", b"b", b"#", b" ", b"ST", b"OP", b"\xc3", b"\xa9", b"\xe6", b"\x97", b"\xa5", b"ab", b"b#", b"##"]
    n_stopped = n_held = 0
    for _ in range(5000):
        pieces = [b"".join(alpha[i] for i in rng.integers(0, len(alpha), int(rng.integers(1, 4)))) for _ in range(int(rng.integers(1, 12)))]
        stops = [b"".join(alpha[i] for i in rng.integers(0, 6, int(rng.integers(1, 4)))) for _ in range(int(rng.integers(0, 4)))]
        want = run(R.ref_stream_text, pieces, stops)
        assert run(L.b2_stream_text, pieces, stops) == want, (pieces, stops, want)
        n_stopped += want[2]
        n_held += want[0] != want[1]
    assert n_stopped > 1000 and n_held > 1000


def test_utf8_hold_back_equals_the_references_own_code():
    """"Does the generated text end inside a multi-byte character" -- the product's test against the block of process_token
    (grpc-server.cpp:1027-1054) compiled from the reference's source: random byte strings biased towards lead and
    continuation bytes, valid text cut at every byte, and invalid sequences."""
    path = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle", "_ref", "libstop_ref.so")
    if not os.path.exists(path):
        pytest.skip("oracle/_ref/libstop_ref.so was not built (no /root/reference on this machine)")
    R = ctypes.CDLL(path)
    R.ref_utf8_incomplete.argtypes = [ctypes.c_char_p, ctypes.c_size_t]
    L = capi.lib()
    L.b2_utf8_incomplete.argtypes = [ctypes.c_char_p, ctypes.c_size_t]
    rng = np.random.default_rng(43)
    cases = [b"", b"a", b"\xc3", b"\xc3\xa9", b"\xe6\x97", b"\xe6\x97\xa5", b"\xf0\x9f\x98", b"\xf0\x9f\x98\x80", b"\x80\x80\x80\x80\x80", b"\xff\xc3"]
    text = "naïve café 日本語 🙂🚀 x".encode()
    cases += [text[:i] for i in range(len(text) + 1)]
    pool = [0x41, 0x20, 0x80, 0xA9, 0xBF, 0xC3, 0xE6, 0xF0, 0xF8, 0xFF]
    for _ in range(4000):
        cases.append(bytes(pool[i] for i in rng.integers(0, len(pool), int(rng.integers(0, 8)))))
    n_inc = 0
    for c in cases:
        want = R.ref_utf8_incomplete(c, len(c))
        assert L.b2_utf8_incomplete(c, len(c)) == want, c
        n_inc += want
    assert 300 < n_inc < len(cases) - 300


def test_prompt_truncation_equals_the_references_own_code():
    """The truncation block of update_slots (grpc-server.cpp:1692-1721), compiled from the reference's source inside a scope
    that provides the names it uses (oracle/stop_harness.cc), against what the scheduler does to a prompt: n_keep negative,
    zero, small, beyond the window; prompts shorter than, equal to and far beyond the context."""
    path = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle", "_ref", "libstop_ref.so")
    if not os.path.exists(path):
        pytest.skip("oracle/_ref/libstop_ref.so was not built (no /root/reference on this machine)")
    R = ctypes.CDLL(path)
    R.ref_truncate_prompt.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_void_p]
    L = capi.lib()
    L.b2_fit_request_prompt.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_int]
    rng = np.random.default_rng(41)
    n_trunc = 0
    for trial in range(3000):
        n_ctx = int(rng.integers(8, 300))
        n = int(rng.integers(1, 4 * n_ctx))
        n_keep = int(rng.choice([-1, 0, 1, 4, n_ctx // 2, n_ctx - 5, n_ctx - 4, n_ctx, 2 * n_ctx, int(rng.integers(0, n_ctx + 8))]))
        a = rng.integers(0, 1000, n).astype(np.int32)
        b = a.copy()
        flag = ctypes.c_int(0)
        want_n = R.ref_truncate_prompt(a.ctypes.data, n, n_ctx, n_keep, ctypes.byref(flag))
        got_n = L.b2_fit_request_prompt(b.ctypes.data, n, n_ctx, n_keep)
        assert got_n == want_n and np.array_equal(a[:want_n], b[:got_n]), (n, n_ctx, n_keep, want_n, got_n)
        assert bool(flag.value) == (want_n != n) and want_n < max(n_ctx, n + 1)
        n_trunc += flag.value
    assert 500 < n_trunc < 2900


def test_prompt_truncation_rule():
    """A prompt that does not fit keeps n_keep tokens and its tail, dropping whole half-window blocks from the middle
    (grpc-server.cpp:1700-1721); one that fits is untouched."""
    L = capi.lib()
    L.b2_fit_prompt.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_int]
    for n, n_ctx, n_keep in [(10, 16, 4), (16, 16, 4), (100, 16, 4), (1000, 256, 0), (257, 256, 252), (300, 64, 60)]:
        t = np.arange(n, dtype=np.int32)
        m = L.b2_fit_prompt(t.ctypes.data, n, n_ctx, n_keep)
        if n < n_ctx:
            assert m == n
            continue
        block = max(1, (n_ctx - n_keep) // 2)
        erased = (n - n_keep - block) // block * block
        want = np.concatenate([np.arange(n_keep), np.arange(n_keep + erased, n)])
        assert m == len(want) < n_ctx and np.array_equal(t[:m], want), (n, n_ctx, n_keep)


def test_corrupt_gguf_headers_are_rejected(tmp_path, model_files):
    """Header fields are not trusted: a zero / non-power-of-two alignment, tensor offsets that wrap, element counts that
    overflow and counts beyond the file are parse errors, not divisions by zero, wild reads or huge allocations."""
    import struct
    good = open(model_files["mini_q4_k_m"], "rb").read()
    key = b"general.alignment"
    k = good.index(key) + len(key)
    assert struct.unpack("<I", good[k:k + 4])[0] == 4  # GGUF type u32
    for bad_align in (0, 24):
        p = tmp_path / f"align{bad_align}.gguf"
        p.write_bytes(good[:k + 4] + struct.pack("<I", bad_align) + good[k + 8:])
        with pytest.raises(capi.B2Error, match="alignment"):
            capi.Engine(str(p), vocab_only=True)
    # the tensor count: far more tensors than the file has bytes
    p = tmp_path / "count.gguf"
    p.write_bytes(good[:8] + struct.pack("<Q", 1 << 60) + good[16:])
    with pytest.raises(capi.B2Error, match="tensor count|truncated"):
        capi.Engine(str(p), vocab_only=True)
    # the offset of the first tensor info pushed to 2^64 - 64: offset + size wraps around
    name = b"token_embd.weight"
    t = good.index(name) + len(name)
    nd = struct.unpack("<I", good[t:t + 4])[0]
    off_at = t + 4 + 8 * nd + 4
    p = tmp_path / "offset.gguf"
    p.write_bytes(good[:off_at] + struct.pack("<Q", (1 << 64) - 64) + good[off_at + 8:])
    with pytest.raises(capi.B2Error, match="exceeds file size"):
        capi.Engine(str(p), vocab_only=True)
    # a dimension so large that the element count overflows
    p = tmp_path / "dims.gguf"
    p.write_bytes(good[:t + 4] + struct.pack("<Q", 1 << 62) + good[t + 12:])
    with pytest.raises(capi.B2Error, match="invalid shape|block multiple|exceeds"):
        capi.Engine(str(p), vocab_only=True)
