"""Parity tests proper: the CUDA path (through the C ABI) against the oracle and the committed goldens.

Bars (BASELINE.json north_star):
  * dequantised weight blocks          bit-exact
  * activation quantisation (Q8_0/Q8_K) bit-exact (integer work)
  * mat-mul / logits                    relative tolerance stated per test (fp32 summation order differs)
  * greedy token ids                    identical for the first 64 tokens
"""
import os

import numpy as np
import pytest

from localai_b200 import capi
from oracle import oracle as O

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden")

MATMUL_RTOL = 2e-5   # of max|out|: same integer block dots, only the fp32 accumulation order differs
# Through the whole network the comparison is chaotic at the 1e-2 level even though every kernel agrees with the
# oracle to ~2e-7: activations are re-quantised to int8 before each mat-mul (as llama.cpp does), so a 1-ulp
# difference in an input (fp32 summation order) occasionally moves one element across a rounding boundary, and
# that +-1 int8 step is amplified by the following layers (measured: isolated tokens at 6e-3..1e-2 of max|logit|,
# all others at 2e-7; once a flip has happened the difference sits at the int8 quantisation-noise floor, ~1-2 % of
# sigma(logits); see DESIGN.md "Parity").  Stated tolerance: max |dlogit| <= 0.25 sigma(logits) and
# median |dlogit| <= 0.05 sigma(logits); the greedy margin of the fixtures is >= 5 sigma.
LOGITS_MAX_SIGMA = 0.25
LOGITS_MEDIAN_SIGMA = 0.05


def assert_logits_close(got, want, what=""):
    sigma = float(want.std())
    err = np.abs(got - want)
    assert np.isfinite(got).all(), what
    assert err.max() <= LOGITS_MAX_SIGMA * sigma, f"{what}: max err {err.max():.3e} vs sigma {sigma:.3e}"
    assert np.median(err) <= LOGITS_MEDIAN_SIGMA * sigma, f"{what}: median err {np.median(err):.3e} vs sigma {sigma:.3e}"


@pytest.mark.parametrize("name", ["q8_0", "q4_k", "q5_k", "q6_k", "f16", "q4_0", "q4_1", "q5_0", "q5_1", "q2_k", "q3_k", "iq4_nl", "iq4_xs"])
def test_dequant_bit_exact_vs_golden_and_oracle(name):
    g = np.load(os.path.join(GOLD, f"dequant_{name}.npz"))
    n = g["values"].size // 256 * 256
    gt = int(g["ggml_type"])
    got = capi.dequantize(gt, g["blocks"], n)
    assert np.array_equal(got.view(np.uint32), g["values"][:n].view(np.uint32))
    assert np.array_equal(got.view(np.uint32), O.dequantize(gt, g["blocks"], n).view(np.uint32))


@pytest.mark.parametrize("wtype", [capi.B2_T_Q4_K, capi.B2_T_Q5_K, capi.B2_T_Q6_K, capi.B2_T_Q8_0, capi.B2_T_Q4_0, capi.B2_T_Q4_1,
                                   capi.B2_T_Q5_0, capi.B2_T_Q5_1, capi.B2_T_Q2_K, capi.B2_T_Q3_K, capi.B2_T_IQ4_NL,
                                   capi.B2_T_IQ4_XS])
def test_dequant_bit_exact_large_random(wtype):
    import make_gguf as mg
    rng = np.random.default_rng(wtype)
    n_blocks = 4096 * 256 // mg.BLOCK[wtype][0]
    raw = mg.random_blocks(rng, wtype, n_blocks, 0.03)
    n = 4096 * 256
    got = capi.dequantize(wtype, raw, n)
    assert np.array_equal(got.view(np.uint32), O.dequantize(wtype, raw, n).view(np.uint32))


def _act_inputs():
    rng = np.random.default_rng(5)
    x = (rng.standard_normal(4096) * np.repeat(rng.uniform(1e-3, 40, 16), 256)).astype(np.float32)
    x[256:512] = 0.0                       # all-zero block
    x[512:768] = np.float32(0.37)          # every |x| equal: first-max tie-break
    x[768] = -x[769:1024].max() * 3        # negative extreme
    x[1024:1056] = np.linspace(-127, 127, 32, dtype=np.float32) * 0.5  # .5 ties
    return x


def test_activation_quantisation_bit_exact():
    x = _act_inputs()
    assert np.array_equal(capi.quantize_act(capi.B2_T_Q8_0, x), O.quantize_q8_0(x))
    g = np.load(os.path.join(GOLD, "quant_q8_0.npz"))
    assert np.array_equal(capi.quantize_act(capi.B2_T_Q8_0, g["x"]), g["blocks"])
    for wt in (capi.B2_T_Q4_K, capi.B2_T_Q5_K, capi.B2_T_Q6_K):
        assert np.array_equal(capi.quantize_act(wt, x), O.quantize_q8_K(x))
    h = capi.quantize_act(capi.B2_T_F16, x).view(np.float16)
    assert np.array_equal(h.view(np.uint16), x.astype(np.float16).view(np.uint16))


@pytest.mark.parametrize("wtype", [capi.B2_T_Q8_0, capi.B2_T_Q4_K, capi.B2_T_Q5_K, capi.B2_T_Q6_K, capi.B2_T_F16])
@pytest.mark.parametrize("rows,cols,n_tok", [(64, 256, 1), (257, 1024, 3), (1000, 4096, 1), (96, 5632 - 5632 % 256, 5),
                                             (33, 14336, 2)])
def test_mul_mat_matches_oracle(wtype, rows, cols, n_tok):
    import make_gguf as mg
    rng = np.random.default_rng(rows * 7 + cols + n_tok)
    if wtype == capi.B2_T_F16:
        raw = (rng.standard_normal((rows, cols)) / np.sqrt(cols)).astype(np.float16).view(np.uint8)
    else:
        raw = mg.random_blocks(rng, wtype, rows * cols // mg.BLOCK[wtype][0], 1 / np.sqrt(cols))
    x = rng.standard_normal((n_tok, cols)).astype(np.float32)
    x[0, :256] = 0
    got = capi.mul_mat(wtype, raw, rows, cols, x, path=1)
    ref = O.mul_mat(wtype, raw, rows, cols, x)
    assert np.isfinite(got).all()
    assert np.abs(got - ref).max() <= MATMUL_RTOL * np.abs(ref).max()


@pytest.mark.parametrize("path", [3, 4, 5])
@pytest.mark.parametrize("wtype", [capi.B2_T_Q8_0, capi.B2_T_Q4_K, capi.B2_T_Q5_K, capi.B2_T_Q6_K, capi.B2_T_F16])
@pytest.mark.parametrize("rows,cols,n_tok", [(64, 256, 1), (258, 1024, 3), (1000, 4096, 1), (96, 5632, 4), (34, 14336, 2),
                                             (4096, 4096, 1)])
def test_fused_and_mega_kernels_match_oracle(wtype, rows, cols, n_tok, path):
    """the decode-path kernels (fused prologue/epilogue kernel, persistent mega-kernel) against the oracle mat-mul"""
    import make_gguf as mg
    if path == 4 and (n_tok != 1 or wtype == capi.B2_T_F16):
        pytest.skip("mega-kernel: one activation row, quantised weights")
    rng = np.random.default_rng(rows * 3 + cols + n_tok + path)
    if wtype == capi.B2_T_F16:
        raw = (rng.standard_normal((rows, cols)) / np.sqrt(cols)).astype(np.float16).view(np.uint8)
    else:
        raw = mg.random_blocks(rng, wtype, rows * cols // mg.BLOCK[wtype][0], 1 / np.sqrt(cols))
    x = rng.standard_normal((n_tok, cols)).astype(np.float32)
    try:
        got = capi.mul_mat(wtype, raw, rows, cols, x, path=path)
    except capi.B2Error as ex:
        # the mega-kernel declines what it has no lane mapping for (Q5_K; Q8_0 rows that are not a multiple of 512):
        # the engine then runs the per-operator kernels (path 3).  Everything else must be supported.
        mega_ok = wtype in (capi.B2_T_Q4_K, capi.B2_T_Q6_K) or (wtype == capi.B2_T_Q8_0 and cols % 512 == 0)
        if path == 4 and ex.code == capi.B2_EUNSUPPORTED and not mega_ok:
            pytest.skip("shape/type outside the mega-kernel's lane mappings")
        raise
    ref = O.mul_mat(wtype, raw, rows, cols, x)
    assert np.isfinite(got).all()
    assert np.abs(got - ref).max() <= MATMUL_RTOL * np.abs(ref).max()


# fp16 operands (one rounding of the dequantised fp32 weight and of the activation, 2^-11 relative each), fp32
# accumulation: the errors of a row's K products add like a random walk.  (Round 1 used bf16 operands: 1.5e-2.)
GEMM_TC_RTOL = 2e-3


@pytest.mark.parametrize("wtype,rows,cols,n_tok", [
    (capi.B2_T_Q4_K, 256, 256, 128), (capi.B2_T_Q4_K, 512, 1024, 70), (capi.B2_T_Q6_K, 300, 512, 130),
    (capi.B2_T_Q8_0, 256, 2048, 5), (capi.B2_T_Q5_K, 1024, 4096, 512), (capi.B2_T_F16, 256, 512, 33),
    # batched-decode shapes: one token tile, fewer output tiles than SMs -> the K range is split over the SMs
    (capi.B2_T_Q4_K, 4096, 4096, 64), (capi.B2_T_Q6_K, 1024, 14336, 64), (capi.B2_T_Q4_K, 512, 8192, 17),
    # legacy 32-element blocks: this path is the only one they have, from one token on
    (capi.B2_T_Q4_0, 512, 1024, 1), (capi.B2_T_Q4_1, 300, 576, 9), (capi.B2_T_Q5_0, 1024, 4096, 64), (capi.B2_T_Q5_1, 256, 2048, 130),
    (capi.B2_T_Q2_K, 512, 1024, 3), (capi.B2_T_Q3_K, 300, 2048, 70), (capi.B2_T_IQ4_NL, 300, 576, 5), (capi.B2_T_IQ4_XS, 512, 2048, 40)])
def test_gemm_tcgen05_matches_float_reference(wtype, rows, cols, n_tok):
    """the tensor-core path (prefill, batched decode): tcgen05 GEMM on fp16 tiles vs dequantised weights in float64"""
    import make_gguf as mg
    rng = np.random.default_rng(rows + cols + n_tok)
    if wtype == capi.B2_T_F16:
        raw = (rng.standard_normal((rows, cols)) / np.sqrt(cols)).astype(np.float16).view(np.uint8).reshape(-1)
    else:
        raw = mg.random_blocks(rng, wtype, rows * cols // mg.BLOCK[wtype][0], 1 / np.sqrt(cols)).reshape(-1)
    w = O.dequantize(wtype, raw, rows * cols).reshape(rows, cols).astype(np.float64)
    x = rng.standard_normal((n_tok, cols)).astype(np.float32)
    ref = x.astype(np.float64) @ w.T
    got = capi.mul_mat(wtype, raw, rows, cols, x, path=2)
    assert np.isfinite(got).all()
    assert np.abs(got - ref).max() <= GEMM_TC_RTOL * np.abs(ref).max()


# fp16 operands (q exact, d*sc and dmin*m rounded to fp16, one fused multiply-add), fp32 accumulation: each product is
# off by about 3 x 2^-11 relative, the errors of a row's K products add like a random walk
QGEMM_RTOL = 2e-3


@pytest.mark.parametrize("wtype,rows,cols,n_tok", [
    (capi.B2_T_Q4_K, 128, 256, 16), (capi.B2_T_Q4_K, 4096, 4096, 64), (capi.B2_T_Q4_K, 28672, 4096, 64), (capi.B2_T_Q4_K, 512, 8192, 17),
    (capi.B2_T_Q6_K, 1024, 14336, 64), (capi.B2_T_Q6_K, 300, 512, 2), (capi.B2_T_Q6_K, 16384, 2048, 33),
    (capi.B2_T_Q5_K, 1024, 4096, 48), (capi.B2_T_Q5_K, 200, 1024, 7), (capi.B2_T_Q8_0, 256, 2048, 5), (capi.B2_T_Q8_0, 2048, 2048, 64)])
def test_qgemm_matches_float_reference(wtype, rows, cols, n_tok):
    """batched decode (2..64 tokens): the GEMM that dequantises inside the kernel (fp16 operand in TMEM) vs float64"""
    import make_gguf as mg
    rng = np.random.default_rng(rows + cols + n_tok)
    raw = mg.random_blocks(rng, wtype, rows * cols // mg.BLOCK[wtype][0], 1 / np.sqrt(cols)).reshape(-1)
    w = O.dequantize(wtype, raw, rows * cols).reshape(rows, cols).astype(np.float64)
    x = rng.standard_normal((n_tok, cols)).astype(np.float32)
    ref = x.astype(np.float64) @ w.T
    got = capi.mul_mat(wtype, raw, rows, cols, x, path=6)
    assert np.isfinite(got).all()
    err = np.abs(got - ref).max() / np.abs(ref).max()
    assert err <= QGEMM_RTOL, err


def test_mul_mat_linearity_in_rows():
    """size-independent property: rows are independent, so any row subset gives the same values."""
    import make_gguf as mg
    rng = np.random.default_rng(11)
    rows, cols = 2048, 4096
    raw = mg.random_blocks(rng, capi.B2_T_Q4_K, rows * cols // 256, 0.02).reshape(rows, -1)
    x = rng.standard_normal((1, cols)).astype(np.float32)
    full = capi.mul_mat(capi.B2_T_Q4_K, raw, rows, cols, x)
    part = capi.mul_mat(capi.B2_T_Q4_K, raw[512:1024], 512, cols, x)
    assert np.array_equal(full[:, 512:1024], part)


MODELS = ["micro_mixed", "micro_q8_0", "mini_q4_k_m", "mini_q5_k", "mini_f16", "mini_q4_0", "mini_q5_1", "mini_q2_k", "mini_q3_k", "mini_iq4_nl", "mini_iq4_xs",
          "micro_linear", "micro_yarn", "micro_l31",  # RoPE scaling (linear, YaRN, rope_freqs.weight divisors)
          "micro_g8",                                 # 8 query heads per KV head
          "mini_qwen2",                               # qwen2 architecture (QKV biases, NEOX rotary pairs)
          "mini_qwen2_g7",                            # ... with 7 query heads per KV head (Qwen2.5-7B's grouping)
          "mini_phi3"]                                # phi3 architecture (fused attn_qkv / gate|up tensors, NEOX rotary pairs)


@pytest.mark.parametrize("name", MODELS)
def test_logits_match_oracle_prefill_and_decode(model_files, name):
    path = model_files[name]
    ref = O.OracleModel(path, n_ctx=128)
    rng = np.random.default_rng(3)
    toks = rng.integers(0, ref.n_vocab, 40).astype(np.int32)
    with capi.Engine(path, n_ctx=128, n_batch=16) as e:
        # prefill (3 chunks of <= 16 tokens), all logits
        got = e.eval(toks[:37], logits_all=True)
        want = ref.eval(toks[:37], logits_all=True)
        assert_logits_close(got, want, "prefill")
        # then token-by-token decode on top of the same KV cache
        for i in range(37, 40):
            g1 = e.eval(toks[i:i + 1], n_past=i)
            w1 = ref.eval(toks[i:i + 1], n_past=i)
            assert_logits_close(g1, w1, f"decode pos {i}")
        # embedding = final-norm hidden state of the last token
        _, embd = e.eval(toks[:8], want_embd=True)
        _, wembd = O.OracleModel(path, n_ctx=128).eval(toks[:8], want_embd=True)
        assert_logits_close(embd[None], wembd[None], "embedding")  # 8 tokens: tensor-core (fp16 operand) path


def test_hidden_states_per_layer(model_files):
    path = model_files["mini_q4_k_m"]
    ref = O.OracleModel(path, n_ctx=64)
    dbg = ref.enable_debug()
    toks = np.arange(100, 120, dtype=np.int32)
    ref.eval(toks)
    with capi.Engine(path, n_ctx=64) as e:
        e.arm_debug_hidden()
        e.eval(toks)
        got = e.debug_hidden()
    for l in range(dbg.shape[0]):
        assert np.abs(got[l] - dbg[l]).max() <= 0.1 * dbg[l].std(), f"layer {l}"
        assert np.median(np.abs(got[l] - dbg[l])) <= 0.02 * dbg[l].std(), f"layer {l}"


@pytest.mark.parametrize("name", MODELS)
def test_greedy_64_tokens_identical(model_files, name):
    path = model_files[name]
    ref = O.OracleModel(path, n_ctx=160)
    prompt = np.random.default_rng(9).integers(0, ref.n_vocab, 48).astype(np.int32)
    want, _, margins = ref.generate_greedy(prompt, 64)
    with capi.Engine(path, n_ctx=160, n_batch=32) as e:
        got, _, _ = e.decode_greedy(prompt, 64)
    print(f"{name}: min top-1 margin {margins.min():.4f}, median {np.median(margins):.4f}")
    assert got.tolist() == want.tolist()


@pytest.mark.parametrize("n_prompt", [300, 700])
def test_greedy_long_context(model_files, n_prompt, monkeypatch):
    """decode at a context where the mega-kernel's attention runs several rounds per CTA (each of the 4 CTAs that share
    a head owns > 64 positions) -- greedy tokens identical to the oracle.  The prompt goes through the integer kernels
    (B200LLM_GEMM_MIN_TOKENS) so that the KV cache holds the oracle's numerics, not the fp16 tensor-core path's."""
    monkeypatch.setenv("B200LLM_GEMM_MIN_TOKENS", "1000000")
    path = model_files["mini_q4_k_m"]
    ref = O.OracleModel(path, n_ctx=n_prompt + 40)
    prompt = np.random.default_rng(n_prompt).integers(0, ref.n_vocab, n_prompt).astype(np.int32)
    want, _, margins = ref.generate_greedy(prompt, 24)
    with capi.Engine(path, n_ctx=n_prompt + 40, n_batch=64) as e:
        got, _, _ = e.decode_greedy(prompt, 24)
    print(f"ctx {n_prompt}: min top-1 margin {margins.min():.4f}")
    assert got.tolist() == want.tolist()


def test_mega_tags_wrap_around(model_files, monkeypatch):
    """the decode kernel's hand-over tags (phase number of the launch sequence, 16 per decode step for this model) and its
    phase counter only grow and are compared modulo 2^32: start them just below the wrap and decode across it, in both
    waiting modes (every warp polls its slice / one thread polls the phase counter first)"""
    path = model_files["mini_q4_k_m"]
    ref = O.OracleModel(path, n_ctx=160)
    prompt = np.random.default_rng(21).integers(0, ref.n_vocab, 40).astype(np.int32)
    want, _, _ = ref.generate_greedy(prompt, 48)
    for probe, below in (("0", 300), ("0", 7), ("2", 20000), ("2", 300)):
        monkeypatch.setenv("B200LLM_LL_PROBE", probe)
        monkeypatch.setenv("B200LLM_MEGA_COUNTER0", str(2 ** 32 - below))
        with capi.Engine(path, n_ctx=160, n_batch=32) as e:
            got, _, _ = e.decode_greedy(prompt, 48)
        assert got.tolist() == want.tolist(), f"tags started {below} below the wrap (waiting mode {probe})"


@pytest.mark.parametrize("name", ["mini_q4_k_m", "micro_q8_0"])
def test_grouped_query_attention_is_bit_identical(model_files, name, monkeypatch):
    """the decode attention that serves all query heads of a KV group from one CTA gives exactly the per-head kernel's
    numbers (same arithmetic, same order per head)"""
    monkeypatch.setenv("B200LLM_NO_MEGA", "1")   # batch-1 decode through the per-operator path, i.e. launch_attention
    monkeypatch.setenv("B200LLM_NO_GRAPH", "1")
    path = model_files[name]
    prompt = np.random.default_rng(3).integers(0, 4000, 150).astype(np.int32)
    outs = []
    for no_gqa in ("0", "1"):
        monkeypatch.setenv("B200LLM_NO_GQA_ATT", no_gqa)
        with capi.Engine(path, n_ctx=256, n_batch=64) as e:
            e.eval(prompt[:-3])
            outs.append([e.eval(prompt[-3 + i:len(prompt) - 2 + i], n_past=len(prompt) - 3 + i) for i in range(3)])
    for a, b in zip(*outs):
        assert np.array_equal(a, b)


def test_batched_decode_equals_single(model_files):
    """4 sequences decoded in one batch give the logits each gives alone (slots are independent)."""
    path = model_files["mini_q4_k_m"]
    rng = np.random.default_rng(4)
    prompts = [rng.integers(0, 4000, n).astype(np.int32) for n in (5, 9, 17, 12)]
    with capi.Engine(path, n_ctx=4 * 64, n_parallel=4, n_batch=64) as e:
        singles = [e.eval(p, slot=i) for i, p in enumerate(prompts)]
        rids = [e.submit(prompt_ids=p, n_predict=8, temperature=0.0, ignore_eos=True, stream=False) for p in prompts]
        outs = []
        for r in rids:
            evs = list(e.events(r))
            assert evs[-1]["flags"] & capi.EV_FINAL and evs[-1]["n_decoded"] == 8
            outs.append(evs[-1])
        for i, p in enumerate(prompts):  # the same requests one by one give the same text
            r = e.submit(prompt_ids=p, n_predict=8, temperature=0.0, ignore_eos=True, stream=False)
            ev = list(e.events(r))[-1]
            assert ev["text"] == outs[i]["text"]
        assert all(np.isfinite(s).all() for s in singles)


def test_tinyllama_q8_0_greedy_64_full_size(gguf_dir):
    """BASELINE configs[0]: TinyLlama-1.1B-arch Q8_0, 128-token prompt, greedy; first 64 tokens identical."""
    import make_gguf as mg
    path = os.path.join(gguf_dir, "tinyllama_q8_0.gguf")
    mg.write_gguf(path, mg.ARCHS["tinyllama-1.1b"], "q8_0", seed=0, quiet=True)
    ref = O.OracleModel(path, n_ctx=256)
    prompt = np.random.default_rng(0).integers(3, 4000, 128).astype(np.int32)
    want, last, margins = ref.generate_greedy(prompt, 64)
    with capi.Engine(path, n_ctx=256, n_batch=128) as e:
        got, t_pre, t_dec = e.decode_greedy(prompt, 64)
        logits = e.eval(prompt)  # prompt logits vs oracle
    wl = O.OracleModel(path, n_ctx=256).eval(prompt)
    print(f"tinyllama: min margin {margins.min():.4f} (sigma {wl.std():.3f}); prefill {t_pre:.1f} ms, decode {t_dec / 64:.3f} "
          f"ms/token; max logit err {np.abs(logits - wl).max():.3e}, median {np.median(np.abs(logits - wl)):.3e}")
    assert_logits_close(logits, wl, "tinyllama prompt logits")
    assert got.tolist() == want.tolist()
    os.remove(path)


@pytest.mark.parametrize("n_tok", [2, 5, 16, 40, 64])
def test_batched_decode_paths_agree(model_files, n_tok, monkeypatch):
    """A step of 2..64 tokens that all want logits (the shape of a decode step over many streams; the engine runs it as one
    step) through (a) the GEMM that dequantises inside the kernel (the default), (b) the same launched with programmatic
    dependent launch, (c) the fp16-tile GEMM.  (a) and (b) are bit-identical --
    the launch attribute changes scheduling, not arithmetic -- and stay so when the graph is replayed; all three hold the
    oracle's logits within the stated tolerance."""
    path = model_files["mini_q4_k_m"]
    toks = np.random.default_rng(n_tok).integers(0, 4000, n_tok + 8).astype(np.int32)

    def run():
        with capi.Engine(path, n_ctx=128, n_batch=64) as e:
            e.eval(toks[:8])                                             # a short prompt first: the step attends to a non-empty cache
            first = e.eval(toks[8:], n_past=8, logits_all=True).copy()   # builds the graph
            again = e.eval(toks[8:], n_past=8, logits_all=True)          # replays it
            assert np.array_equal(first, again)
            return first

    a = run()
    monkeypatch.setenv("B200LLM_PDL", "1")
    b = run()
    monkeypatch.delenv("B200LLM_PDL")
    monkeypatch.setenv("B200LLM_NO_QGEMM", "1")
    c = run()
    assert np.array_equal(a, b)
    ref = O.OracleModel(path, n_ctx=128).eval(toks, logits_all=True)[8:]
    assert_logits_close(a, ref, "qgemm")
    assert_logits_close(c, ref, "fp16 tiles")


@pytest.mark.parametrize("n_streams,model", [(2, "mini_q4_k_m"), (7, "mini_q4_k_m"), (16, "mini_q4_k_m"), (33, "mini_q4_k_m"), (64, "mini_q4_k_m"),
                                             (5, "micro_g8"), (40, "micro_g8")])
def test_decode_step_over_many_slots(model_files, n_streams, model, monkeypatch):
    """ONE decode step over n_streams slots with prompts of different lengths (b2_eval_step: the batch shape the scheduler
    issues for concurrent streams).  The default path is a CUDA graph of: RMSNorm -> fp16 tiles, the GEMM that dequantises
    inside the kernel, the fused RoPE + KV-store + attention + tile-conversion kernel, ....  Its logits equal, bit for bit,
    those of the same step (a) replayed and (b) launched with programmatic dependent launch (the attribute changes
    scheduling, not arithmetic) and (c) run with RoPE / attention / conversion as three launches (the same arithmetic in
    the same order); they hold the oracle's logits of each stream within the stated tolerance, as does (d) the fp16-tile
    GEMM path."""
    path = model_files[model]
    rng = np.random.default_rng(100 + n_streams)
    prompts = [rng.integers(0, 4000, 5 + (3 * i) % 17).astype(np.int32) for i in range(n_streams)]
    toks = np.array([int(rng.integers(0, 4000)) for _ in prompts], np.int32)
    pos = np.array([len(p) for p in prompts], np.int32)
    slots = np.arange(n_streams, dtype=np.int32)

    def run(replay=False):
        with capi.Engine(path, n_ctx=32 * n_streams, n_batch=64, n_parallel=n_streams) as e:
            for s_, p in enumerate(prompts):
                e.eval(p, slot=s_)
            out = e.eval_step(toks, pos, slots).copy()
            if replay:  # the same step again overwrites the same cache cells with the same values
                assert np.array_equal(out, e.eval_step(toks, pos, slots))
            return out

    a = run(replay=True)
    monkeypatch.setenv("B200LLM_PDL", "1")
    assert np.array_equal(a, run())
    monkeypatch.delenv("B200LLM_PDL")
    monkeypatch.setenv("B200LLM_NO_FUSED_ATT", "1")
    assert np.array_equal(a, run())
    monkeypatch.delenv("B200LLM_NO_FUSED_ATT")
    monkeypatch.setenv("B200LLM_NO_QGEMM", "1")
    d = run()
    ref = np.stack([O.OracleModel(path, n_ctx=64).eval(np.append(p, t))[0] for p, t in zip(prompts, toks)])
    assert_logits_close(a, ref, "default path")
    assert_logits_close(d, ref, "fp16 tiles")


@pytest.mark.parametrize("name", ["mini_q4_k_m", "mini_f16", "micro_mixed"])
def test_prefill_fused_swiglu_epilogue(model_files, name, monkeypatch):
    """Prompt chunks above 64 tokens run the tcgen05 GEMM on fp16 tiles; ffn_gate | ffn_up are one interleaved tile set and
    the GEMM's epilogue applies SwiGLU and writes ffn_down's operand tiles.  Against the separate SwiGLU kernel only the fp32
    summation order differs (at 128 tokens the unfused gate / up GEMM splits K over the idle SMs, the fused one does not):
    the logits agree to a few 1e-3 sigma; both hold the oracle's within the stated tolerance."""
    path = model_files[name]
    toks = np.random.default_rng(5).integers(0, 4000, 150).astype(np.int32)

    def run():
        with capi.Engine(path, n_ctx=256, n_batch=128) as e:
            return e.eval(toks, logits_all=True)  # chunks of 128 + 22 tokens

    a = run()
    monkeypatch.setenv("B200LLM_NO_FUSED_SWIGLU", "1")
    b = run()
    ref = O.OracleModel(path, n_ctx=256).eval(toks, logits_all=True)
    assert np.abs(a - b).max() <= 0.02 * ref.std()
    assert_logits_close(a, ref, "fused SwiGLU epilogue")
    assert_logits_close(b, ref, "separate SwiGLU kernel")


@pytest.mark.parametrize("name,kw", [
    ("micro_yarn", dict(rope_scaling="linear")),   # what the reference does with an empty ModelOptions.RopeScaling on a YaRN file
    ("micro_yarn", dict(rope_scaling="none")),
    ("micro_linear", dict(rope_scaling="yarn")),
    ("micro_yarn", dict(yarn_ext_factor=0.5, yarn_attn_factor=1.25, yarn_beta_fast=16.0, yarn_beta_slow=2.0)),
])
def test_rope_scaling_overrides_of_model_options(model_files, name, kw):
    """ModelOptions.RopeScaling / Yarn* override what the GGUF says (grpc-server.cpp:2279-2293): the engine with
    b2_load_opts.rope_scaling / yarn_* holds the logits of the oracle configured the same way, and they differ from the
    file's own configuration."""
    path = model_files[name]
    toks = np.arange(7, 47, dtype=np.int32)
    ref = O.OracleModel(path, n_ctx=64, **kw).eval(toks, logits_all=True)
    plain = O.OracleModel(path, n_ctx=64).eval(toks, logits_all=True)
    assert np.abs(ref[8:] - plain[8:]).max() > 1e-3
    with capi.Engine(path, n_ctx=64, n_batch=16, **kw) as e:
        got = e.eval(toks, logits_all=True)
    assert_logits_close(got, ref, f"{name} {kw}")


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["mini_qwen2", "mini_phi3"])
def test_neox_architectures_on_the_llama_kernels(model_files, name, monkeypatch):
    """qwen2 / phi3 rotate the pairs (i, i + head_dim / 2).  LoadModel interleaves the rows of wq / wk (and their biases) per
    head instead, after which every llama kernel -- the persistent decode kernel included, for phi3 -- applies the
    architecture's own rotation.  Against the oracle (which rotates the un-interleaved rows), and against the engine with the
    rotation done by the per-operator kernel on un-interleaved rows (B200LLM_NEOX_KERNEL=1): prompt logits, then the decode
    path token by token."""
    path = model_files[name]
    toks = np.arange(11, 43, dtype=np.int32)
    ref = O.OracleModel(path, n_ctx=128)
    want = ref.eval(toks, logits_all=True)
    outs = {}
    for mode in ("interleaved", "kernel"):
        monkeypatch.setenv("B200LLM_NEOX_KERNEL", "1" if mode == "kernel" else "0")
        with capi.Engine(path, n_ctx=128, n_batch=64) as e:
            lg = e.eval(toks, logits_all=True)
            assert_logits_close(lg, want, f"{name} {mode} prompt")
            step = [lg[-1]]
            n_past, tok = len(toks), int(lg[-1].argmax())
            for _ in range(6):  # one token at a time: the decode path of this architecture
                l1 = e.eval(np.array([tok], dtype=np.int32), n_past=n_past)[-1]
                step.append(l1)
                n_past, tok = n_past + 1, int(l1.argmax())
            outs[mode] = (lg, np.stack(step))
    sig = float(want.std())
    assert np.abs(outs["interleaved"][0] - outs["kernel"][0]).max() <= 0.02 * sig
    assert np.array_equal(outs["interleaved"][1].argmax(-1), outs["kernel"][1].argmax(-1))
    # the oracle, fed the same tokens, agrees with the decode path step by step
    n_past = len(toks)
    for i in range(6):
        tok = int(outs["interleaved"][1][i].argmax())
        w1 = ref.eval(np.array([tok], dtype=np.int32), n_past=n_past)[-1]
        assert_logits_close(outs["interleaved"][1][i + 1][None], w1[None], f"{name} decode step {i}")
        n_past += 1
