"""Parity at the sizes BASELINE.json quotes its numbers on (the configurations bench.py times), and on a fixture whose
greedy continuation is NOT built into the weights.

  * configs[1]  Llama-3-8B-arch Q4_K_M (the bench file: tools/make_gguf.py, seed 0): 128-token prompt logits against the
                oracle within the stated tolerance, the first 64 greedy tokens through the decode mega-kernel identical;
  * configs[2]  64 concurrent streams on the same file give, stream by stream, the text each gives alone;
  * configs[3]  BERT-base-arch F16, one 512-token text against oracle/bert_ref.py;
  * unstructured fixtures (write_gguf(structured=False): purely random output matrix, near-flat logits): the oracle's
    greedy continuation is teacher-forced through the engine's decode path and the arg-max must agree wherever the
    oracle's top-1 margin exceeds twice the logits tolerance; the agreement count and the margin histogram are printed.
"""
import os

import numpy as np
import pytest

from localai_b200 import capi
from oracle import oracle as O
from test_parity_gpu import LOGITS_MAX_SIGMA, assert_logits_close

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def llama3_8b():
    import bench
    path, info = bench.ensure_model("llama3-8b", "q4_k_m", 0, 0, lambda: None)
    return path


def test_llama3_8b_q4_k_m_prompt_logits_and_greedy_64(llama3_8b):
    """The headline configuration at its own size: K = 4096 / 14336 splits, Q6_K ffn_down / output rows, the 128 256-row
    LM head, 32 layers of error growth."""
    path = llama3_8b
    prompt = np.random.default_rng(0).integers(3, 100000, 128).astype(np.int32)
    ref = O.OracleModel(path, n_ctx=256)
    want_logits = ref.eval(prompt)[-1]
    want, margins = [], []
    lg, n_past = want_logits, len(prompt)
    for _ in range(64):  # the oracle's greedy loop, continuing on its own KV cache
        tok = int(lg.argmax())
        top2 = np.partition(lg, -2)[-2:]
        margins.append(float(top2[1] - top2[0]))
        want.append(tok)
        lg = ref.eval(np.array([tok], dtype=np.int32), n_past=n_past)[-1]
        n_past += 1
    with capi.Engine(path, n_ctx=256, n_batch=128) as e:
        logits = e.eval(prompt)[-1]      # the prompt goes through the tensor-core path, as in production
        got, t_pre, t_dec = e.decode_greedy(prompt, 64)  # decode: the mega-kernel
    sigma = float(want_logits.std())
    err = np.abs(logits - want_logits)
    print(f"llama3-8b q4_k_m: prompt logits max err {err.max() / sigma:.4f} sigma, median {np.median(err) / sigma:.4f} sigma; "
          f"min greedy margin {min(margins) / sigma:.2f} sigma; prefill {t_pre:.1f} ms, decode {t_dec / 64:.3f} ms/token")
    assert_logits_close(logits[None], want_logits[None], "llama3-8b prompt logits")
    assert got.tolist() == want


def test_llama3_8b_64_streams_equal_single(llama3_8b):
    """configs[2]: 64 concurrent streams (continuous batching, tensor-core decode) against the same 64 requests one by
    one (mega-kernel decode)."""
    rng = np.random.default_rng(64)
    prompts = [rng.integers(3, 100000, int(n)).astype(np.int32) for n in rng.integers(4, 24, 64)]
    with capi.Engine(llama3_8b, n_ctx=64 * 64, n_parallel=64, n_batch=512) as e:
        rids = [e.submit(prompt_ids=p, n_predict=12, temperature=0.0, ignore_eos=True, stream=False) for p in prompts]
        batched = []
        for r in rids:
            ev = list(e.events(r))[-1]
            e.release(r)
            assert ev["flags"] & capi.EV_FINAL and ev["n_decoded"] == 12
            batched.append(ev["text"])
        m = e.metrics()
        for i, p in enumerate(prompts):
            r = e.submit(prompt_ids=p, n_predict=12, temperature=0.0, ignore_eos=True, stream=False)
            ev = list(e.events(r))[-1]
            e.release(r)
            assert ev["text"] == batched[i], f"stream {i}"
    assert m["n_tokens_decoded_total"] >= 64 * 12


def test_bert_base_f16_512_tokens(gguf_dir):
    """configs[3] at full size: BERT-base-arch F16, one text of 512 WordPiece ids."""
    import make_bert_gguf as mb
    from oracle import bert_ref as B

    path = os.path.join(gguf_dir, "bert_base_f16.gguf")
    mb.write_bert_gguf(path, mb.BERT_ARCHS["bert-base"], seed=0)
    ref = B.BertRef(path)
    rng = np.random.default_rng(512)
    ids = np.concatenate([[ref.cls], rng.integers(5, 30000, 510), [ref.sep]]).astype(np.int32)
    want_h = ref.embed_tokens(ids, return_hidden=True)
    want_e = ref.embed_tokens(ids)
    with capi.Bert(path) as b:
        got_h = b.hidden(ids)
        got_e = b.embed_tokens(ids)
        again = b.embed_tokens(ids)
    sigma = float(want_h.std())
    err = np.abs(got_h - want_h)
    cos = float((got_e * want_e).sum() / np.sqrt((got_e * got_e).sum() * (want_e * want_e).sum()))
    print(f"bert-base f16, 512 tokens: hidden max err {err.max() / sigma:.4f} sigma, median {np.median(err) / sigma:.5f} sigma, "
          f"cosine(embedding) {cos:.6f}")
    assert err.max() <= 0.15 * sigma and np.median(err) <= 0.012 * sigma  # bf16 tensor-core operands (tests/test_bert.py)
    assert cos >= 0.9995
    assert np.array_equal(got_e, again)
    os.remove(path)


@pytest.mark.parametrize("preset,arch", [("q4_k_m", "mini"), ("q8_0", "mini"), ("q4_k_m", "small")])
def test_unstructured_fixture_teacher_forced_argmax(gguf_dir, preset, arch, monkeypatch):
    """Random output matrix (no built-in continuation): real, near-flat logits.  The prompt runs through the integer
    kernels so that the KV cache holds the CPU path's numerics; then the oracle's greedy ids are fed one by one through
    the decode path (the mega-kernel) and every step's arg-max is compared with the oracle's."""
    import make_gguf as mg

    if arch not in mg.ARCHS:
        pytest.skip(f"no {arch} architecture in tools/make_gguf.py")
    monkeypatch.setenv("B200LLM_GEMM_MIN_TOKENS", "1000000")
    path = os.path.join(gguf_dir, f"{arch}_{preset}_unstructured.gguf")
    mg.write_gguf(path, mg.ARCHS[arch], preset, seed=11, gain_out=1.0, quiet=True, structured=False)
    ref = O.OracleModel(path, n_ctx=256)
    prompt = np.random.default_rng(17).integers(0, 4000, 96).astype(np.int32)
    lg = ref.eval(prompt)[-1]
    n_past = len(prompt)
    toks, margins, sigmas, oracle_logits = [], [], [], []
    for _ in range(64):
        tok = int(lg.argmax())
        top2 = np.partition(lg, -2)[-2:]
        toks.append(tok); margins.append(float(top2[1] - top2[0])); sigmas.append(float(lg.std())); oracle_logits.append(lg)
        lg = ref.eval(np.array([tok], dtype=np.int32), n_past=n_past)[-1]
        n_past += 1
    agree, must, worst = 0, 0, 0.0
    with capi.Engine(path, n_ctx=256, n_batch=128) as e:
        g = e.eval(prompt)[-1]
        n_past = len(prompt)
        for i in range(64):
            sig = sigmas[i]
            worst = max(worst, float(np.abs(g - oracle_logits[i]).max() / sig))
            ok = int(g.argmax()) == toks[i]
            agree += ok
            if margins[i] > 2 * LOGITS_MAX_SIGMA * sig:  # the tolerance cannot flip this one
                must += 1
                assert ok, f"position {i}: margin {margins[i] / sig:.3f} sigma, engine picked {int(g.argmax())}, oracle {toks[i]}"
            g = e.eval(np.array([toks[i]], dtype=np.int32), n_past=n_past)[-1]  # teacher forcing: the oracle's token
            n_past += 1
    ms = np.array(margins) / np.array(sigmas)
    hist = np.histogram(ms, bins=[0, 0.01, 0.03, 0.1, 0.3, 1, 3, 100])[0]
    print(f"{arch} {preset} unstructured: arg-max agreement {agree}/64 ({must} positions with margin > {2 * LOGITS_MAX_SIGMA} sigma); "
          f"worst logit error {worst:.4f} sigma; margins (sigma) histogram [<.01,.03,.1,.3,1,3,inf): {hist.tolist()}, min {ms.min():.4f}")
    assert worst <= LOGITS_MAX_SIGMA
    # the rest sit at margins below the int8 re-quantisation noise, where any last-bit change of a kernel flips a few: the 8-layer
    # fixture gave 60 and 59 with two builds whose prefill attention differs only in how exp() is rounded (worst error 0.17 / 0.16 sigma)
    assert agree >= 56
    os.remove(path)
