"""ctypes binding of include/b200llm.h (libb200llm.so).

This is the same C ABI the Go backend binds through cgo (backend/go/llm/b200/llm.go); Python uses it for
the tests, bench.py and the gRPC shim.  There is no fallback: if the shared library is missing, importing
this module raises, and every compute call raises B2Error when no sm_100 device is present.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Iterator, Optional, Sequence

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("B200LLM_LIB", os.path.join(_HERE, "lib", "libb200llm.so"))

B2_T_F32, B2_T_F16, B2_T_Q8_0, B2_T_Q4_K, B2_T_Q5_K, B2_T_Q6_K = 0, 1, 8, 12, 13, 14
B2_T_Q4_0, B2_T_Q4_1, B2_T_Q5_0, B2_T_Q5_1, B2_T_Q2_K, B2_T_Q3_K = 2, 3, 6, 7, 10, 11  # tensor-core path only
B2_T_IQ4_NL, B2_T_IQ4_XS = 20, 23  # likewise
B2_ETIMEOUT = -8
B2_TP_NCCL, B2_TP_FUSED = 0, 1
B2_ROPE_FROM_GGUF, B2_ROPE_NONE, B2_ROPE_LINEAR, B2_ROPE_YARN = 0, 1, 2, 3
B2_EUNSUPPORTED = -9
EV_FINAL, EV_STOP_EOS, EV_STOP_WORD, EV_STOP_LIMIT, EV_ERROR, EV_EMBEDDING = 1, 2, 4, 8, 16, 32

EXPORTS = [
    "b2_tp_set_mode", "b2_tp_info", "b2_abi_version", "b2_last_error", "b2_device_count", "b2_create", "b2_destroy", "b2_model_info", "b2_tokenize",
    "b2_token_to_piece", "b2_submit", "b2_poll", "b2_cancel", "b2_embed_result", "b2_release", "b2_metrics", "b2_eval",
    "b2_decode_greedy", "b2_eval_step", "b2_debug_hidden", "b2_bench_decode", "b2_bench_prefill", "b2_bench_gemv", "b2_dequantize", "b2_quantize_act",
    "b2_mul_mat", "b2_mul_mat_noalloc", "b2_sample", "b2_grammar_match", "b2_grammar_token_mask",
    "b2_bert_load", "b2_bert_free", "b2_bert_info", "b2_bert_tokenize", "b2_bert_embed_text", "b2_bert_embed_tokens",
    "b2_bert_embed_batch", "b2_bert_hidden",
]


class B2Error(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"b200llm error {code}: {msg}")
        self.code = code


class LoadOpts(C.Structure):
    _fields_ = [
        ("struct_size", C.c_uint32), ("model_path", C.c_char_p), ("n_ctx", C.c_int32), ("n_batch", C.c_int32),
        ("n_parallel", C.c_int32), ("device", C.c_int32), ("embeddings", C.c_int32), ("rope_freq_base", C.c_float),
        ("rope_freq_scale", C.c_float), ("kv_page_size", C.c_int32), ("vocab_only", C.c_int32),
        ("tp_size", C.c_int32), ("rope_scaling", C.c_int32), ("yarn_ext_factor", C.c_float), ("yarn_attn_factor", C.c_float),
        ("yarn_beta_fast", C.c_float), ("yarn_beta_slow", C.c_float), ("reserved", C.c_int32 * 1),
    ]


class ModelInfo(C.Structure):
    _fields_ = [(n, C.c_int32) for n in (
        "n_layer", "n_embd", "n_head", "n_head_kv", "head_dim", "n_ff", "n_vocab", "n_ctx_train", "n_ctx", "n_ctx_slot",
        "n_parallel", "n_batch", "bos_id", "eos_id", "eot_id", "add_bos")] + [
        ("rms_eps", C.c_float), ("rope_freq_base", C.c_float), ("rope_freq_scale", C.c_float),
        ("weight_bytes_device", C.c_int64), ("decode_stream_bytes", C.c_int64), ("kv_bytes_per_pos", C.c_int64),
        ("arch", C.c_char * 32), ("tokenizer_model", C.c_char * 32), ("name", C.c_char * 128)]


class Request(C.Structure):
    _fields_ = [
        ("struct_size", C.c_uint32), ("prompt", C.c_char_p), ("prompt_len", C.c_size_t),
        ("prompt_ids", C.POINTER(C.c_int32)), ("n_prompt_ids", C.c_int32), ("n_predict", C.c_int32),
        ("temperature", C.c_float), ("top_k", C.c_int32), ("top_p", C.c_float), ("min_p", C.c_float),
        ("repeat_penalty", C.c_float), ("repeat_last_n", C.c_int32), ("frequency_penalty", C.c_float),
        ("presence_penalty", C.c_float), ("seed", C.c_int32), ("ignore_eos", C.c_int32), ("cache_prompt", C.c_int32),
        ("n_keep", C.c_int32), ("embedding_mode", C.c_int32), ("stop", C.POINTER(C.c_char_p)),
        ("stop_len", C.POINTER(C.c_size_t)), ("n_stop", C.c_int32), ("stream", C.c_int32),
        ("mirostat", C.c_int32), ("mirostat_tau", C.c_float), ("mirostat_eta", C.c_float), ("typical_p", C.c_float),
        ("tfs_z", C.c_float), ("penalize_nl", C.c_int32), ("reserved", C.c_int32 * 2), ("grammar", C.c_char_p),
    ]


class SampleParams(C.Structure):
    _fields_ = [
        ("struct_size", C.c_uint32), ("temperature", C.c_float), ("top_k", C.c_int32), ("top_p", C.c_float), ("min_p", C.c_float),
        ("repeat_penalty", C.c_float), ("frequency_penalty", C.c_float), ("presence_penalty", C.c_float),
        ("mirostat", C.c_int32), ("mirostat_tau", C.c_float), ("mirostat_eta", C.c_float), ("mirostat_mu", C.c_float),
        ("ban_token", C.c_int32), ("nl_token", C.c_int32), ("seed", C.c_int64), ("prev", C.POINTER(C.c_int32)),
        ("n_prev", C.c_int32), ("tfs_z", C.c_float), ("typical_p", C.c_float),
    ]


class Event(C.Structure):
    _fields_ = [
        ("struct_size", C.c_uint32), ("flags", C.c_uint32), ("text", C.c_void_p), ("text_len", C.c_size_t),
        ("n_prompt_tokens", C.c_int32), ("n_decoded", C.c_int32), ("slot_id", C.c_int32), ("n_embd", C.c_int32),
        ("t_prompt_ms", C.c_double), ("t_gen_ms", C.c_double), ("reserved", C.c_int32 * 4),
    ]


class BertInfo(C.Structure):
    _fields_ = [("n_layer", C.c_int32), ("n_embd", C.c_int32), ("n_head", C.c_int32), ("n_ff", C.c_int32), ("n_ctx", C.c_int32),
                ("n_vocab", C.c_int32), ("weight_bytes_device", C.c_int64)]


class TPInfo(C.Structure):
    _fields_ = [("tp_size", C.c_int32), ("mode", C.c_int32), ("n_allreduce_per_step", C.c_int32), ("reserved", C.c_int32),
                ("allreduce_ms_per_step", C.c_double), ("shard_stream_bytes", C.c_int64)]


class Metrics(C.Structure):
    _fields_ = [
        ("slot_id", C.c_int32), ("tokens_per_second", C.c_float), ("tokens_generated", C.c_int32),
        ("prompt_tokens_processed", C.c_int32), ("n_slots_busy", C.c_int32), ("n_decode_steps", C.c_int64),
        ("n_tokens_decoded_total", C.c_int64), ("n_prompt_tokens_total", C.c_int64),
    ]


_lib = None


def lib() -> C.CDLL:
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} not found: build it with `make -C localai_b200/csrc` (or __graft_entry__.build()). "
            "There is no CPU fallback.")
    L = C.CDLL(LIB_PATH)
    vp, i32, i64 = C.c_void_p, C.c_int, C.c_int64
    L.b2_abi_version.restype = i32
    L.b2_last_error.restype = C.c_char_p
    L.b2_device_count.restype = i32
    L.b2_create.restype = vp
    L.b2_create.argtypes = [C.POINTER(LoadOpts)]
    L.b2_destroy.argtypes = [vp]
    L.b2_model_info.argtypes = [vp, C.POINTER(ModelInfo)]
    L.b2_tokenize.argtypes = [vp, C.c_char_p, C.c_size_t, i32, i32, vp, i32]
    L.b2_token_to_piece.argtypes = [vp, C.c_int32, vp, i32]
    L.b2_submit.restype = i64
    L.b2_submit.argtypes = [vp, C.POINTER(Request)]
    L.b2_poll.argtypes = [vp, i64, C.POINTER(Event), i32]
    L.b2_cancel.argtypes = [vp, i64]
    L.b2_embed_result.argtypes = [vp, i64, vp, i32]
    L.b2_release.argtypes = [vp, i64]
    L.b2_metrics.argtypes = [vp, C.POINTER(Metrics)]
    L.b2_tp_set_mode.argtypes = [vp, i32]
    L.b2_tp_info.argtypes = [vp, C.POINTER(TPInfo)]
    L.b2_eval.argtypes = [vp, i32, vp, i32, i32, vp, i32, vp]
    L.b2_decode_greedy.argtypes = [vp, i32, vp, i32, i32, vp, C.POINTER(C.c_double), C.POINTER(C.c_double)]
    L.b2_debug_hidden.argtypes = [vp, vp, i32]
    L.b2_bench_decode.argtypes = [vp, i32, i32, i32, i32, C.POINTER(C.c_double), C.POINTER(C.c_int64)]
    L.b2_bench_prefill.argtypes = [vp, i32, i32, i32, C.POINTER(C.c_double), C.POINTER(C.c_int64)]
    L.b2_bench_gemv.argtypes = [vp, i32, i32, i32, C.POINTER(C.c_double), C.POINTER(C.c_int64), C.POINTER(C.c_int)]
    L.b2_dequantize.argtypes = [i32, vp, i64, vp]
    L.b2_quantize_act.argtypes = [i32, vp, i64, vp, C.c_size_t]
    L.b2_mul_mat.argtypes = [i32, vp, i64, i64, vp, i32, vp, i32]
    L.b2_mul_mat_noalloc.argtypes = [i32, vp, i64, i64, vp, i32, vp, vp]
    L.b2_sample.argtypes = [i32, vp, i32, C.POINTER(SampleParams), i32, vp, vp]
    L.b2_bert_load.restype = vp
    L.b2_bert_load.argtypes = [C.c_char_p, i32, i32]
    L.b2_bert_free.restype = None
    L.b2_bert_free.argtypes = [vp]
    L.b2_bert_info.argtypes = [vp, C.POINTER(BertInfo)]
    L.b2_bert_tokenize.argtypes = [vp, C.c_char_p, C.c_size_t, vp, i32]
    L.b2_bert_embed_text.argtypes = [vp, C.c_char_p, C.c_size_t, vp, i32]
    L.b2_bert_embed_tokens.argtypes = [vp, vp, i32, vp, i32]
    L.b2_bert_embed_batch.argtypes = [vp, vp, vp, i32, vp]
    L.b2_bert_hidden.argtypes = [vp, vp, i32, vp]
    L.b2_grammar_match.argtypes = [C.c_char_p, C.c_char_p, C.c_size_t, C.POINTER(C.c_size_t), C.POINTER(C.c_int)]
    L.b2_grammar_token_mask.argtypes = [C.c_char_p, C.c_char_p, C.c_size_t, C.POINTER(C.c_char_p), C.POINTER(C.c_size_t), vp, i32, vp]
    _lib = L
    return L


def grammar_match(gbnf: str | bytes, text: str | bytes):
    """b2_grammar_match -> (bytes of `text` accepted, grammar may stop there).  Raises on a grammar that does not parse."""
    g = gbnf.encode("utf-8") if isinstance(gbnf, str) else bytes(gbnf)
    t = text.encode("utf-8") if isinstance(text, str) else bytes(text)
    n, stop = C.c_size_t(0), C.c_int(0)
    _check(lib().b2_grammar_match(g, t, len(t), C.byref(n), C.byref(stop)))
    return int(n.value), bool(stop.value)


def grammar_token_mask(gbnf: str | bytes, accepted: str | bytes, pieces: Sequence[bytes], is_eog: Sequence[bool] | None = None):
    """b2_grammar_token_mask -> bool[n_vocab]: which pieces may follow `accepted`."""
    g = gbnf.encode("utf-8") if isinstance(gbnf, str) else bytes(gbnf)
    a = accepted.encode("utf-8") if isinstance(accepted, str) else bytes(accepted)
    n = len(pieces)
    arr = (C.c_char_p * n)(*[bytes(p) for p in pieces])
    lens = (C.c_size_t * n)(*[len(p) for p in pieces])
    eog = np.ascontiguousarray(is_eog if is_eog is not None else np.zeros(n), dtype=np.uint8)
    bits = np.zeros((n + 31) // 32, dtype=np.uint32)
    _check(lib().b2_grammar_token_mask(g, a, len(a), arr, lens, eog.ctypes.data, n, bits.ctypes.data))
    return ((bits[np.arange(n) >> 5] >> (np.arange(n) & 31).astype(np.uint32)) & 1).astype(bool)


def _check(rc: int) -> int:
    if rc < 0:
        raise B2Error(rc, (lib().b2_last_error() or b"").decode("utf-8", "replace"))
    return rc


def _ptr(a: np.ndarray):
    return a.ctypes.data_as(C.c_void_p)


def device_count() -> int:
    return lib().b2_device_count()


# ------------------------------------------------------------------------------- kernel-level hooks
def dequantize(gtype: int, raw: np.ndarray, n_elems: int) -> np.ndarray:
    raw = np.ascontiguousarray(raw).view(np.uint8).reshape(-1)
    out = np.empty(n_elems, dtype=np.float32)
    _check(lib().b2_dequantize(gtype, _ptr(raw), n_elems, _ptr(out)))
    return out


def quantize_act(wtype: int, x: np.ndarray) -> np.ndarray:
    x = np.ascontiguousarray(x, dtype=np.float32).reshape(-1)
    out = np.empty(x.size * 4 + 64, dtype=np.uint8)
    n = _check(lib().b2_quantize_act(wtype, _ptr(x), x.size, _ptr(out), out.size))
    return out[:n].copy()


def mul_mat(wtype: int, w_raw: np.ndarray, rows: int, cols: int, x: np.ndarray, path: int = 0) -> np.ndarray:
    w_raw = np.ascontiguousarray(w_raw).view(np.uint8).reshape(-1)
    x = np.ascontiguousarray(x, dtype=np.float32).reshape(-1, cols)
    out = np.empty((x.shape[0], rows), dtype=np.float32)
    _check(lib().b2_mul_mat(wtype, _ptr(w_raw), rows, cols, _ptr(x), x.shape[0], _ptr(out), path))
    return out


def sample(logits: np.ndarray, n_draws: int = 1, temperature: float = 0.0, top_k: int = 0, top_p: float = 1.0, min_p: float = 0.0,
           repeat_penalty: float = 1.0, frequency_penalty: float = 0.0, presence_penalty: float = 0.0, prev: Sequence[int] = (),
           mirostat: int = 0, mirostat_tau: float = 5.0, mirostat_eta: float = 0.1, mirostat_mu: Optional[float] = None,
           ban_token: int = -1, nl_token: int = -1, seed: int = 0, device: int = 0, tfs_z: float = 1.0, typical_p: float = 1.0):
    """b2_sample: the device sampler alone on one row of host logits -> (tokens[n_draws], mu_after_each_draw[n_draws])."""
    lg = np.ascontiguousarray(logits, dtype=np.float32).reshape(-1)
    pv = np.ascontiguousarray(prev, dtype=np.int32)
    p = SampleParams()
    p.struct_size = C.sizeof(SampleParams)
    p.temperature, p.top_k, p.top_p, p.min_p = temperature, top_k, top_p, min_p
    p.repeat_penalty, p.frequency_penalty, p.presence_penalty = repeat_penalty, frequency_penalty, presence_penalty
    p.mirostat, p.mirostat_tau, p.mirostat_eta = mirostat, mirostat_tau, mirostat_eta
    p.mirostat_mu = 2.0 * mirostat_tau if mirostat_mu is None else mirostat_mu
    p.ban_token, p.nl_token, p.seed = ban_token, nl_token, seed
    p.tfs_z, p.typical_p = tfs_z, typical_p
    p.prev = pv.ctypes.data_as(C.POINTER(C.c_int32))
    p.n_prev = pv.size
    toks = np.empty(n_draws, dtype=np.int32)
    mus = np.empty(n_draws, dtype=np.float32)
    _check(lib().b2_sample(device, lg.ctypes.data_as(C.POINTER(C.c_float)), lg.size, C.byref(p), n_draws,
                           toks.ctypes.data_as(C.POINTER(C.c_int32)), mus.ctypes.data_as(C.POINTER(C.c_float))))
    return toks, mus


def mul_mat_noalloc(wtype: int, w_dev: int, rows: int, cols: int, x_dev: int, n_tok: int, out_dev: int, stream: int = 0):
    """b2_mul_mat_noalloc on raw DEVICE addresses (ints); asynchronous on `stream`; allocates nothing."""
    _check(lib().b2_mul_mat_noalloc(wtype, C.c_void_p(w_dev), rows, cols, C.c_void_p(x_dev), n_tok, C.c_void_p(out_dev),
                                    C.c_void_p(stream)))


# ------------------------------------------------------------------------------- engine
class Engine:
    """One loaded model on one GPU (b2_create ... b2_destroy)."""

    def __init__(self, model_path: str, n_ctx: int = 1024, n_batch: int = 512, n_parallel: int = 1, device: int = 0,
                 embeddings: bool = False, rope_freq_base: float = 0.0, rope_freq_scale: float = 0.0,
                 kv_page_size: int = 0, vocab_only: bool = False, tp_size: int = 0, rope_scaling: str | int = 0,
                 yarn_ext_factor: float = 0.0, yarn_attn_factor: float = 0.0, yarn_beta_fast: float = 0.0, yarn_beta_slow: float = 0.0):
        o = LoadOpts()
        o.rope_scaling = {"": 0, "none": B2_ROPE_NONE, "linear": B2_ROPE_LINEAR, "yarn": B2_ROPE_YARN}.get(rope_scaling, rope_scaling)
        o.yarn_ext_factor, o.yarn_attn_factor, o.yarn_beta_fast, o.yarn_beta_slow = yarn_ext_factor, yarn_attn_factor, yarn_beta_fast, yarn_beta_slow
        o.struct_size = C.sizeof(LoadOpts)
        o.model_path = os.fsencode(model_path)
        o.n_ctx, o.n_batch, o.n_parallel, o.device = n_ctx, n_batch, n_parallel, device
        o.embeddings = int(embeddings)
        o.vocab_only = int(vocab_only)
        o.tp_size = int(tp_size)
        o.rope_freq_base, o.rope_freq_scale, o.kv_page_size = rope_freq_base, rope_freq_scale, kv_page_size
        self._h = lib().b2_create(C.byref(o))
        if not self._h:
            raise B2Error(-1, (lib().b2_last_error() or b"").decode("utf-8", "replace"))
        self.info = ModelInfo()
        _check(lib().b2_model_info(self._h, C.byref(self.info)))

    def close(self):
        if getattr(self, "_h", None):
            lib().b2_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    # -- tokenizer
    def tokenize(self, text: str | bytes, add_bos: bool = True, parse_special: bool = True) -> list[int]:
        b = text.encode("utf-8") if isinstance(text, str) else bytes(text)
        cap = len(b) + 16  # never more tokens than bytes (+ BOS/EOS/space prefix)
        out = np.empty(cap, dtype=np.int32)
        n = lib().b2_tokenize(self._h, b, len(b), int(add_bos), int(parse_special), _ptr(out), cap)
        _check(n)
        return out[:n].tolist()

    def token_to_piece(self, tok: int) -> bytes:
        buf = C.create_string_buffer(256)
        n = _check(lib().b2_token_to_piece(self._h, tok, buf, 256))
        return buf.raw[:n]

    # -- low level
    def eval(self, tokens: Sequence[int], n_past: int = 0, slot: int = 0, logits_all: bool = False,
             want_embd: bool = False):
        t = np.ascontiguousarray(tokens, dtype=np.int32)
        V, d = self.info.n_vocab, self.info.n_embd
        logits = np.empty((t.size if logits_all else 1, V), dtype=np.float32)
        embd = np.empty(d, dtype=np.float32) if want_embd else None
        _check(lib().b2_eval(self._h, slot, _ptr(t), t.size, n_past, _ptr(logits), int(logits_all),
                             _ptr(embd) if want_embd else None))
        return (logits, embd) if want_embd else logits

    def eval_step(self, tokens: Sequence[int], pos: Sequence[int], slots: Sequence[int]) -> np.ndarray:
        """one decode step over several slots: logits [n][n_vocab]"""
        t = np.ascontiguousarray(tokens, dtype=np.int32)
        p = np.ascontiguousarray(pos, dtype=np.int32)
        s = np.ascontiguousarray(slots, dtype=np.int32)
        assert t.size == p.size == s.size
        logits = np.empty((t.size, self.info.n_vocab), dtype=np.float32)
        _check(lib().b2_eval_step(self._h, _ptr(t), _ptr(p), _ptr(s), t.size, _ptr(logits)))
        return logits

    def decode_greedy(self, prompt: Sequence[int], n_gen: int, slot: int = 0):
        p = np.ascontiguousarray(prompt, dtype=np.int32)
        out = np.zeros(max(n_gen, 1), dtype=np.int32)
        a, b = C.c_double(0), C.c_double(0)
        n = _check(lib().b2_decode_greedy(self._h, slot, _ptr(p), p.size, n_gen, _ptr(out), C.byref(a), C.byref(b)))
        return out[:n], a.value, b.value

    def arm_debug_hidden(self):
        _check(lib().b2_debug_hidden(self._h, None, 0))

    def debug_hidden(self) -> np.ndarray:
        n = (self.info.n_layer + 1) * self.info.n_embd
        out = np.empty(n, dtype=np.float32)
        _check(lib().b2_debug_hidden(self._h, _ptr(out), n))
        return out.reshape(self.info.n_layer + 1, self.info.n_embd)

    def bench_decode(self, batch: int, ctx_len: int, n_steps: int, n_warmup: int = 3):
        ms, l = C.c_double(0), C.c_int64(0)
        _check(lib().b2_bench_decode(self._h, batch, ctx_len, n_steps, n_warmup, C.byref(ms), C.byref(l)))
        return ms.value, l.value

    def tp_set_mode(self, mode: int):
        """0 = per-layer kernels + ncclAllReduce, 1 = persistent kernels exchanging partial rows over peer memory"""
        _check(lib().b2_tp_set_mode(self._h, mode))

    def tp_info(self) -> dict:
        t = TPInfo()
        _check(lib().b2_tp_info(self._h, C.byref(t)))
        return {k: getattr(t, k) for k, _ in TPInfo._fields_ if k != "reserved"}

    def last_mega_ms(self) -> float:
        """mean decode_mega_kernel launch (ms) of the last bench_decode(1, ...): CUDA events around every launch"""
        f = lib().b2_bench_last_mega_ms
        f.restype, f.argtypes = C.c_double, [C.c_void_p]
        return float(f(self._h))

    def bench_prefill(self, n_tokens: int, n_iters: int, n_warmup: int = 1):
        ms, l = C.c_double(0), C.c_int64(0)
        _check(lib().b2_bench_prefill(self._h, n_tokens, n_iters, n_warmup, C.byref(ms), C.byref(l)))
        return ms.value, l.value

    def bench_gemv(self, which: int, n_tok: int = 1, n_rounds: int = 4):
        ms, nb, ty = C.c_double(0), C.c_int64(0), C.c_int(0)
        _check(lib().b2_bench_gemv(self._h, which, n_tok, n_rounds, C.byref(ms), C.byref(nb), C.byref(ty)))
        return ms.value, nb.value, ty.value

    # -- requests
    def submit(self, prompt: str | bytes | None = None, prompt_ids: Optional[Sequence[int]] = None, n_predict: int = 0,
               temperature: float = 0.0, top_k: int = 40, top_p: float = 0.95, min_p: float = 0.0, seed: int = -1,
               ignore_eos: bool = False, cache_prompt: bool = False, n_keep: int = 0, embedding_mode: bool = False,
               stop: Sequence[str | bytes] = (), stream: bool = True, repeat_penalty: float = 1.0, repeat_last_n: int = 64,
               frequency_penalty: float = 0.0, presence_penalty: float = 0.0, mirostat: int = 0, mirostat_tau: float = 5.0,
               mirostat_eta: float = 0.1, typical_p: float = 1.0, tfs_z: float = 1.0, grammar: str | bytes | None = None,
               penalize_nl: bool = False) -> int:
        r = Request()
        r.struct_size = C.sizeof(Request)
        keep = []
        if prompt is not None:
            b = prompt.encode("utf-8") if isinstance(prompt, str) else bytes(prompt)
            r.prompt, r.prompt_len = b, len(b)
            keep.append(b)
        if prompt_ids is not None:
            ids = np.ascontiguousarray(prompt_ids, dtype=np.int32)
            r.prompt_ids = ids.ctypes.data_as(C.POINTER(C.c_int32))
            r.n_prompt_ids = ids.size
            keep.append(ids)
        r.n_predict, r.temperature, r.top_k, r.top_p, r.min_p, r.seed = n_predict, temperature, top_k, top_p, min_p, seed
        r.repeat_penalty, r.repeat_last_n = repeat_penalty, repeat_last_n
        r.frequency_penalty, r.presence_penalty = frequency_penalty, presence_penalty
        r.mirostat, r.mirostat_tau, r.mirostat_eta, r.typical_p, r.tfs_z = mirostat, mirostat_tau, mirostat_eta, typical_p, tfs_z
        r.ignore_eos, r.cache_prompt, r.n_keep = int(ignore_eos), int(cache_prompt), n_keep
        r.embedding_mode, r.stream = int(embedding_mode), int(stream)
        r.penalize_nl = int(penalize_nl)
        if grammar:
            g = grammar.encode("utf-8") if isinstance(grammar, str) else bytes(grammar)
            r.grammar = g
            keep.append(g)
        if stop:
            bs = [s.encode("utf-8") if isinstance(s, str) else bytes(s) for s in stop]
            arr = (C.c_char_p * len(bs))(*bs)
            lens = (C.c_size_t * len(bs))(*[len(s) for s in bs])
            r.stop, r.stop_len, r.n_stop = arr, lens, len(bs)
            keep += [bs, arr, lens]
        rid = lib().b2_submit(self._h, C.byref(r))
        if rid < 0:
            _check(int(rid))
        return int(rid)

    def poll(self, rid: int, timeout_ms: int = -1):
        ev = Event()
        ev.struct_size = C.sizeof(Event)
        rc = lib().b2_poll(self._h, rid, C.byref(ev), timeout_ms)
        if rc == B2_ETIMEOUT:
            return None
        _check(rc)
        text = C.string_at(ev.text, ev.text_len) if ev.text and ev.text_len else b""
        return dict(flags=ev.flags, text=text, n_prompt_tokens=ev.n_prompt_tokens, n_decoded=ev.n_decoded,
                    slot_id=ev.slot_id, n_embd=ev.n_embd, t_prompt_ms=ev.t_prompt_ms, t_gen_ms=ev.t_gen_ms)

    def events(self, rid: int, timeout_ms: int = -1) -> Iterator[dict]:
        """every event of the request up to the final one; timeout_ms bounds the wait for each (-1: none)"""
        while True:
            ev = self.poll(rid, timeout_ms)
            if ev is None:
                raise TimeoutError(f"request {rid}: no event within {timeout_ms} ms")
            yield ev
            if ev["flags"] & EV_FINAL:
                return

    def cancel(self, rid: int):
        _check(lib().b2_cancel(self._h, rid))

    def release(self, rid: int):
        _check(lib().b2_release(self._h, rid))

    def embed_result(self, rid: int) -> np.ndarray:
        out = np.empty(self.info.n_embd, dtype=np.float32)
        n = _check(lib().b2_embed_result(self._h, rid, _ptr(out), out.size))
        return out[:n]

    def metrics(self) -> dict:
        m = Metrics()
        _check(lib().b2_metrics(self._h, C.byref(m)))
        return {k: getattr(m, k) for k, _ in Metrics._fields_}


class Bert:
    """The `bert-embeddings` backend over the C ABI: Load / Embeddings / TokenEmbeddings (backend/go/llm/bert/bert.go)."""

    def __init__(self, model_path: str, device: int = 0, vocab_only: bool = False):
        self._h = lib().b2_bert_load(model_path.encode(), device, int(vocab_only))
        if not self._h:
            raise B2Error(-1, (lib().b2_last_error() or b"").decode("utf-8", "replace"))
        info = BertInfo()
        _check(lib().b2_bert_info(self._h, C.byref(info)))
        self.info = {k: getattr(info, k) for k, _ in BertInfo._fields_}
        self.n_embd = info.n_embd

    def close(self):
        if self._h:
            lib().b2_bert_free(self._h)
            self._h = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def tokenize(self, text: str | bytes) -> list[int]:
        t = text.encode("utf-8") if isinstance(text, str) else bytes(text)
        buf = np.empty(max(16, self.info["n_ctx"]), dtype=np.int32)
        n = _check(lib().b2_bert_tokenize(self._h, t, len(t), buf.ctypes.data, buf.size))
        return buf[:n].tolist()

    def embed(self, text: str | bytes) -> np.ndarray:
        t = text.encode("utf-8") if isinstance(text, str) else bytes(text)
        out = np.empty(self.n_embd, dtype=np.float32)
        _check(lib().b2_bert_embed_text(self._h, t, len(t), out.ctypes.data, out.size))
        return out

    def embed_tokens(self, ids: Sequence[int]) -> np.ndarray:
        a = np.ascontiguousarray(ids, dtype=np.int32)
        out = np.empty(self.n_embd, dtype=np.float32)
        _check(lib().b2_bert_embed_tokens(self._h, a.ctypes.data, a.size, out.ctypes.data, out.size))
        return out

    def embed_batch(self, seqs: Sequence[Sequence[int]]) -> np.ndarray:
        offs = np.zeros(len(seqs) + 1, dtype=np.int32)
        offs[1:] = np.cumsum([len(s) for s in seqs])
        ids = np.ascontiguousarray(np.concatenate([np.asarray(s, dtype=np.int32) for s in seqs]), dtype=np.int32)
        out = np.empty((len(seqs), self.n_embd), dtype=np.float32)
        _check(lib().b2_bert_embed_batch(self._h, ids.ctypes.data, offs.ctypes.data, len(seqs), out.ctypes.data))
        return out

    def hidden(self, ids: Sequence[int]) -> np.ndarray:
        a = np.ascontiguousarray(ids, dtype=np.int32)
        out = np.empty((a.size, self.n_embd), dtype=np.float32)
        _check(lib().b2_bert_hidden(self._h, a.ctypes.data, a.size, out.ctypes.data))
        return out
