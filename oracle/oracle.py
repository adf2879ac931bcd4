"""ctypes face of oracle/ggml_ref.c + GGUF loading through gguf-py.  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import
this module; the product package (localai_b200) never does.

The GGUF container is parsed with gguf-py (llama.cpp's own Python package, gguf.GGUFReader), so the
oracle's view of a file is independent of the product's C++ parser (localai_b200/csrc/gguf.cc).
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "_build", "libggml_ref.so")

T_F32, T_F16, T_Q8_0, T_Q4_K, T_Q5_K, T_Q6_K, T_Q8_K = 0, 1, 8, 12, 13, 14, 15
T_Q4_0, T_Q4_1, T_Q5_0, T_Q5_1 = 2, 3, 6, 7
T_Q2_K, T_Q3_K = 10, 11
T_IQ4_NL, T_IQ4_XS = 20, 23
(TK_TOK_EMBD, TK_OUTPUT_NORM, TK_OUTPUT, TK_ATTN_NORM, TK_WQ, TK_WK, TK_WV, TK_WO, TK_FFN_NORM, TK_GATE, TK_UP,
 TK_DOWN, TK_BQ, TK_BK, TK_BV) = range(15)

_LAYER_TENSORS = {
    "attn_norm": TK_ATTN_NORM, "attn_q": TK_WQ, "attn_k": TK_WK, "attn_v": TK_WV, "attn_output": TK_WO,
    "ffn_norm": TK_FFN_NORM, "ffn_gate": TK_GATE, "ffn_up": TK_UP, "ffn_down": TK_DOWN,
}
_LAYER_BIASES = {"attn_q": TK_BQ, "attn_k": TK_BK, "attn_v": TK_BV}  # blk.N.attn_q.bias ... (qwen2)
_GLOBAL_TENSORS = {"token_embd": TK_TOK_EMBD, "output_norm": TK_OUTPUT_NORM, "output": TK_OUTPUT}


def build(force: bool = False) -> str:
    if force or not os.path.exists(_LIB_PATH) or os.path.getmtime(_LIB_PATH) < os.path.getmtime(
            os.path.join(_HERE, "ggml_ref.c")):
        subprocess.check_call(["make", "-C", _HERE, "-s"])
    return _LIB_PATH


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_LIB_PATH)
        vp, i32, i64, f32 = C.c_void_p, C.c_int, C.c_int64, C.c_float
        L.ref_row_bytes.restype = i64
        L.ref_row_bytes.argtypes = [i32, i64]
        L.ref_dequantize_row.argtypes = [i32, vp, vp, i64]
        L.ref_quantize_row_q8_0.argtypes = [vp, vp, i64]
        L.ref_quantize_row_q8_K.argtypes = [vp, vp, i64]
        L.ref_vec_dot_type.restype = i32
        L.ref_vec_dot_type.argtypes = [i32]
        L.ref_quantize_activations.argtypes = [i32, vp, vp, i64]
        L.ref_vec_dot.restype = f32
        L.ref_vec_dot.argtypes = [i32, i32, vp, vp]
        L.ref_mul_mat.argtypes = [i32, vp, i64, i64, vp, i32, vp]
        L.ref_rope_cache.argtypes = [f32, f32, f32, i32, vp]
        L.ref_ctx_new.restype = vp
        L.ref_ctx_new.argtypes = [i32] * 7 + [f32] * 3
        L.ref_ctx_free.argtypes = [vp]
        L.ref_ctx_set_tensor.argtypes = [vp, i32, i32, vp, i32, i64, i64]
        L.ref_ctx_set_debug.argtypes = [vp, vp]
        L.ref_eval.restype = i32
        L.ref_eval.argtypes = [vp, vp, i32, i32, vp, i32, vp]
        L.ref_argmax.restype = i32
        L.ref_argmax.argtypes = [vp, i32]
        L.ref_generate_greedy.restype = i32
        L.ref_generate_greedy.argtypes = [vp, vp, i32, i32, vp, vp, vp]
        L.ref_num_threads.restype = i32
        _lib = L
    return _lib


def _ptr(a: np.ndarray):
    return a.ctypes.data_as(C.c_void_p)


# ----------------------------------------------------------------------------- block-level helpers
def dequantize(gtype: int, raw: np.ndarray, n_elems: int) -> np.ndarray:
    raw = np.ascontiguousarray(raw).view(np.uint8).reshape(-1)
    out = np.empty(n_elems, dtype=np.float32)
    lib().ref_dequantize_row(gtype, _ptr(raw), _ptr(out), n_elems)
    return out


def quantize_q8_0(x: np.ndarray) -> np.ndarray:
    x = np.ascontiguousarray(x, dtype=np.float32).reshape(-1)
    out = np.empty(x.size // 32 * 34, dtype=np.uint8)
    lib().ref_quantize_row_q8_0(_ptr(x), _ptr(out), x.size)
    return out


def quantize_q8_K(x: np.ndarray) -> np.ndarray:
    x = np.ascontiguousarray(x, dtype=np.float32).reshape(-1)
    out = np.empty(x.size // 256 * 292, dtype=np.uint8)
    lib().ref_quantize_row_q8_K(_ptr(x), _ptr(out), x.size)
    return out


def mul_mat(gtype: int, w_raw: np.ndarray, rows: int, cols: int, x: np.ndarray) -> np.ndarray:
    """out[t, r] = dot(W[r], quantise(x[t])) with the reference's activation quantisation."""
    w_raw = np.ascontiguousarray(w_raw).view(np.uint8).reshape(-1)
    x = np.ascontiguousarray(x, dtype=np.float32).reshape(-1, cols)
    out = np.empty((x.shape[0], rows), dtype=np.float32)
    lib().ref_mul_mat(gtype, _ptr(w_raw), rows, cols, _ptr(x), x.shape[0], _ptr(out))
    return out


def rope_cache(pos: int, freq_scale: float, freq_base: float, n_dims: int) -> np.ndarray:
    out = np.empty(n_dims, dtype=np.float32)
    lib().ref_rope_cache(float(pos), freq_scale, freq_base, n_dims, _ptr(out))
    return out


# ----------------------------------------------------------------------------- model
class OracleModel:
    """llama-arch GGUF evaluated by the C restatement (grpc-server.cpp:1910 llama_decode call site)."""

    def __init__(self, path: str, n_ctx: int = 512, rope_scaling: str | None = None, yarn_ext_factor: float = 0.0,
                 yarn_attn_factor: float = 0.0, yarn_beta_fast: float = 0.0, yarn_beta_slow: float = 0.0):
        """rope_scaling / yarn_*: the overrides ModelOptions carries (grpc-server.cpp:2279-2293); None / 0 = what the GGUF says"""
        import gguf  # llama.cpp's own reader

        self.reader = gguf.GGUFReader(path)  # keeps the mmap alive
        f = self.reader.fields

        def scalar(key, default=None):
            if key not in f:
                return default
            fld = f[key]
            return fld.parts[fld.data[0]][0]

        def string(key, default):
            if key not in f:
                return default
            fld = f[key]
            return bytes(fld.parts[fld.data[0]]).decode()

        A = string("general.architecture", "llama") + "."
        self.arch = A[:-1]
        self.n_layer = int(scalar(A + "block_count"))
        self.n_embd = int(scalar(A + "embedding_length"))
        self.n_head = int(scalar(A + "attention.head_count"))
        self.n_head_kv = int(scalar(A + "attention.head_count_kv", self.n_head))
        self.n_ff = int(scalar(A + "feed_forward_length"))
        self.rms_eps = float(scalar(A + "attention.layer_norm_rms_epsilon", 1e-5))
        self.rope_theta = float(scalar(A + "rope.freq_base", 10000.0))
        self.n_ctx = n_ctx
        self.tensors = {t.name: t for t in self.reader.tensors}
        self.n_vocab = int(self.tensors["token_embd.weight"].shape[1])
        L = lib()
        self.ctx = L.ref_ctx_new(self.n_layer, self.n_embd, self.n_head, self.n_head_kv, self.n_ff, self.n_vocab, n_ctx,
                                 self.rms_eps, self.rope_theta, 1.0)
        self._keep = []
        # RoPE scaling as llama.cpp derives it from the metadata (llama.cpp: llm_load_hparams / llama_new_context_with_model)
        stype = rope_scaling or string(A + "rope.scaling.type", "none")
        factor = float(scalar(A + "rope.scaling.factor", 1.0))
        self.rope_freq_scale = 1.0 / factor if stype in ("linear", "yarn") and factor > 0 else 1.0
        self.rope_ext_factor = yarn_ext_factor or (1.0 if stype == "yarn" else 0.0)
        overridden = bool(rope_scaling or yarn_ext_factor or yarn_attn_factor or yarn_beta_fast or yarn_beta_slow)
        n_ctx_orig = int(scalar(A + "rope.scaling.original_context_length", scalar(A + "context_length", 2048)))
        ff = None
        if "rope_freqs.weight" in self.tensors:
            ff = np.ascontiguousarray(np.array(self.tensors["rope_freqs.weight"].data, dtype=np.float32))
            self._keep.append(ff)
        if stype != "none" or ff is not None or overridden:
            L.ref_ctx_set_rope_scaling.argtypes = [C.c_void_p, C.c_float, C.c_float, C.c_float, C.c_int,
                                                   C.c_float, C.c_float, C.c_void_p]
            L.ref_ctx_set_rope_scaling(self.ctx, self.rope_freq_scale, self.rope_ext_factor, yarn_attn_factor or 1.0, n_ctx_orig,
                                       yarn_beta_fast or 32.0, yarn_beta_slow or 1.0, ff.ctypes.data if ff is not None else None)
        if self.arch in ("qwen2", "phi3"):  # LLAMA_ROPE_TYPE_NEOX (llama.cpp llama_rope_type)
            L.ref_ctx_set_rope_neox.argtypes = [C.c_void_p, C.c_int]
            L.ref_ctx_set_rope_neox(self.ctx, 1)
        for name, t in self.tensors.items():
            if name.endswith(".bias"):  # blk.N.attn_{q,k,v}.bias: one f32 row
                _, il, nm = name[:-len(".bias")].split(".", 2)
                if nm in _LAYER_BIASES:
                    data = np.ascontiguousarray(np.array(t.data, dtype=np.float32)).view(np.uint8).reshape(-1)
                    self._keep.append(data)
                    L.ref_ctx_set_tensor(self.ctx, _LAYER_BIASES[nm], int(il), _ptr(data), 0, 1, int(t.shape[0]))
                continue
            base = name.rsplit(".weight", 1)[0]
            data = np.ascontiguousarray(t.data).view(np.uint8).reshape(-1)
            self._keep.append(data)
            cols = int(t.shape[0])
            rows = int(t.shape[1]) if len(t.shape) > 1 else 1
            if base in _GLOBAL_TENSORS:
                L.ref_ctx_set_tensor(self.ctx, _GLOBAL_TENSORS[base], 0, _ptr(data), int(t.tensor_type), rows, cols)
            elif base.startswith("blk."):
                _, il, nm = base.split(".", 2)
                if self.arch == "phi3" and nm in ("attn_qkv", "ffn_up"):
                    # phi3 stores [q; k; v] and [gate; up] as the rows of one tensor each (llama.cpp build_phi3 takes
                    # views of the product; rows are contiguous, so the slices are byte ranges)
                    rb = data.size // rows
                    kvd = self.n_head_kv * (self.n_embd // self.n_head)
                    parts = ([(TK_WQ, 0, self.n_embd), (TK_WK, self.n_embd, kvd), (TK_WV, self.n_embd + kvd, kvd)] if nm == "attn_qkv"
                             else [(TK_GATE, 0, rows // 2), (TK_UP, rows // 2, rows // 2)])
                    for kind, r0, nr in parts:
                        L.ref_ctx_set_tensor(self.ctx, kind, int(il), _ptr(data[r0 * rb:]), int(t.tensor_type), nr, cols)
                    continue
                if nm in _LAYER_TENSORS:
                    L.ref_ctx_set_tensor(self.ctx, _LAYER_TENSORS[nm], int(il), _ptr(data), int(t.tensor_type), rows,
                                         cols)
        self._dbg = None

    def close(self):
        if self.ctx:
            lib().ref_ctx_free(self.ctx)
            self.ctx = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def enable_debug(self) -> np.ndarray:
        self._dbg = np.zeros((self.n_layer + 1, self.n_embd), dtype=np.float32)
        lib().ref_ctx_set_debug(self.ctx, _ptr(self._dbg))
        return self._dbg

    def eval(self, tokens, n_past: int = 0, logits_all: bool = False, want_embd: bool = False):
        toks = np.ascontiguousarray(tokens, dtype=np.int32)
        n = toks.size
        logits = np.empty((n if logits_all else 1, self.n_vocab), dtype=np.float32)
        embd = np.empty(self.n_embd, dtype=np.float32) if want_embd else None
        rc = lib().ref_eval(self.ctx, _ptr(toks), n, n_past, _ptr(logits), int(logits_all),
                            _ptr(embd) if want_embd else None)
        if rc != 0:
            raise RuntimeError("context exhausted")
        return (logits, embd) if want_embd else logits

    def generate_greedy(self, prompt, n_gen: int):
        toks = np.ascontiguousarray(prompt, dtype=np.int32)
        out = np.zeros(n_gen, dtype=np.int32)
        last = np.empty(self.n_vocab, dtype=np.float32)
        margins = np.zeros(n_gen, dtype=np.float32)
        n = lib().ref_generate_greedy(self.ctx, _ptr(toks), toks.size, n_gen, _ptr(out), _ptr(last), _ptr(margins))
        return out[:n], last, margins[:n]


def num_threads() -> int:
    return lib().ref_num_threads()


def set_num_threads(n: int) -> None:
    """Use n OpenMP threads regardless of OMP_NUM_THREADS (torchrun exports OMP_NUM_THREADS=1)."""
    lib().ref_set_num_threads(int(n))


# ---------------------------------------------------------------------------------------------- sampler chain
# Restatement of the two llama.cpp samplers the engine runs on the device beyond top-k/top-p/min-p (llama.cpp
# src/llama-sampling.cpp @ d5cb868, un-vendored: llama_sampler_penalties_apply, llama_sampler_mirostat_v2_apply;
# called through gpt_sampler_sample at backend/cpp/llama/grpc-server.cpp:1946).  Test infrastructure only.
def penalties_ref(logits, prev, repeat_penalty=1.0, frequency_penalty=0.0, presence_penalty=0.0, nl_token=-1):
    """logits after the repeat / frequency / presence penalties over the window `prev` (penalize_nl = false)."""
    x = np.array(logits, dtype=np.float32, copy=True)
    counts = {}
    for t in prev:
        counts[int(t)] = counts.get(int(t), 0) + 1
    rp, fp, pp = np.float32(repeat_penalty), np.float32(frequency_penalty), np.float32(presence_penalty)
    for t, c in counts.items():
        if t < 0 or t >= x.size or t == nl_token:
            continue
        v = x[t]
        v = v * rp if v <= 0 else v / rp
        x[t] = np.float32(v - (np.float32(c) * fp + pp))
    return x


def mirostat_v2_ref(logits, temperature, mu, tau, eta, token):
    """One mirostat-v2 step given WHICH token was drawn: returns (kept mask, p of `token` after truncation and
    renormalisation, mu after the update).  The draw itself depends on the random stream and is not restated."""
    z = (np.asarray(logits, dtype=np.float64) - np.max(logits)) / temperature
    p = np.exp(z)
    p /= p.sum()
    surprise = -np.log2(p)
    kept = surprise <= mu
    kept[int(np.argmax(p))] = True  # the most probable token always stays
    p2 = np.where(kept, p, 0.0)
    p2 /= p2.sum()
    obs = -np.log2(p2[token]) if p2[token] > 0 else np.inf
    return kept, p2[token], mu - eta * (obs - tau)


def _softmax_sorted(logits):
    order = np.argsort(-np.asarray(logits, dtype=np.float64), kind="stable")
    z = np.asarray(logits, dtype=np.float64)[order]
    p = np.exp(z - z[0])
    return order, p / p.sum()


def tail_free_ref(logits, z):
    """token ids kept by llama_sampler_tail_free_apply (min_keep = 1) on the given candidates"""
    order, p = _softmax_sorted(logits)
    n = p.size
    if z >= 1.0 or n <= 2:
        return set(order.tolist())
    d2 = np.abs(np.diff(np.diff(-p)))  # |(p_i - p_{i+1}) - (p_{i+1} - p_{i+2})|
    d2 = d2 / d2.sum() if d2.sum() > 1e-6 else np.full(n - 2, 1.0 / (n - 2))
    last, cum = n, 0.0
    for i in range(n - 2):
        cum += d2[i]
        if cum > z and i >= 1:
            last = i
            break
    return set(order[:last].tolist())


def typical_ref(logits, p_keep):
    """token ids kept by llama_sampler_typical_apply (min_keep = 1)"""
    order, p = _softmax_sorted(logits)
    if p_keep >= 1.0:
        return set(order.tolist())
    entropy = -(p * np.log(p)).sum()
    shifted = np.abs(-np.log(p) - entropy)
    idx = np.argsort(shifted, kind="stable")
    cum, last = 0.0, p.size
    for i, k in enumerate(idx):
        cum += p[k]
        if cum > p_keep:
            last = i + 1
            break
    return set(order[idx[:last]].tolist())


def mirostat_v1_ref(logits, temperature, mu, tau, eta, token, m=100):
    """One mirostat-v1 step given WHICH token was drawn (llama_sampler_mirostat_apply): returns (k, p of `token` among
    the k most probable, mu after the update)."""
    order, p = _softmax_sorted(np.asarray(logits, dtype=np.float64) / temperature)
    n = p.size
    mm = min(m, n)
    i = np.arange(mm - 1)
    t = np.log((i + 2) / (i + 1))
    b = np.log(p[:mm - 1] / p[1:mm])
    s_hat = (t * b).sum() / (t * t).sum()
    eps = s_hat - 1.0
    k = ((eps * 2.0 ** mu) / (1.0 - float(n) ** (-eps))) ** (1.0 / s_hat)
    k = n if not np.isfinite(k) or k >= n else max(int(k), 1)
    pk = p[:k] / p[:k].sum()
    where = np.nonzero(order[:k] == token)[0]
    pt = pk[where[0]] if where.size else 0.0
    obs = -np.log2(pt) if pt > 0 else np.inf
    return k, pt, mu - eta * (obs - tau)
