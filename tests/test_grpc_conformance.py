"""Wire-level conformance of the backend.proto `Backend` service (reference: backend/backend.proto:10-31), in the shape of
the reference's own backend test (backend/python/transformers/test.py:13-84): spawn the backend with `--addr`, then
Health -> "OK", LoadModel, the RPCs.  The server is pyshim/server.py over the C ABI -- the same C calls the Go backend
(backend/go/llm/b200) makes through cgo; Go and protoc are absent from this image.

CPU part: the declared messages equal backend.proto field for field (checked against the reference's file when it is
mounted), Health / LoadModel(VocabOnly) / TokenizeString / Status work without a GPU and a failed load is an error.
GPU part: Predict, PredictStream (pieces concatenate to Predict's text, the last Reply carries the counts as
grpc-server.cpp:2355-2358), Embedding, a request with a Grammar, GetMetrics.
"""
import os
import re
import socket
import subprocess
import sys
import time

import grpc
import numpy as np
import pytest

from pyshim import backend_pb as pb
from pyshim.server import Client

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_PROTO = "/root/reference/backend/backend.proto"


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.fixture(scope="module")
def backend():
    """`<backend> --addr host:port`, as pkg/model/process.go:83-95 launches it; ready when Health answers OK
    (pkg/model/initializers.go:362-383)."""
    addr = f"127.0.0.1:{_free_port()}"
    env = dict(os.environ, LLAMACPP_PARALLEL="2")
    proc = subprocess.Popen([sys.executable, "-m", "pyshim.server", "--addr", addr], cwd=ROOT, env=env)
    cl = Client(addr)
    ok = False
    for _ in range(200):
        try:
            if cl.Health(pb.HealthMessage(), timeout=2).message == b"OK":
                ok = True
                break
        except grpc.RpcError:
            time.sleep(0.1)
    assert ok, "backend did not answer Health"
    yield cl
    cl.close()
    proc.kill()
    proc.wait()


def test_messages_match_backend_proto():
    if not os.path.exists(REF_PROTO):
        pytest.skip("reference tree not mounted")
    text = re.sub(r"//[^\n]*", "", open(REF_PROTO).read())
    for name, fields in pb.MESSAGES.items():
        body = re.search(r"message\s+%s\s*\{([^}]*)\}" % name, text).group(1)
        want = {}
        for rep, typ, fname, num in re.findall(r"(repeated\s+|optional\s+)?([\w.]+)\s+(\w+)\s*=\s*(\d+)\s*;", body):
            want[fname] = (typ, int(num), rep.strip() == "repeated")
        got = {f[0]: (f[1], f[2], len(f) > 3) for f in fields}
        assert got == want, name
    svc = re.search(r"service\s+Backend\s*\{(.*?)\n\}", text, re.S).group(1)
    rpcs = {m[0]: (m[1], m[3], bool(m[2])) for m in re.findall(r"rpc\s+(\w+)\((\w+)\)\s+returns\s+\((stream\s+)?(\w+)\)", svc)}
    for rpc, sig in pb.RPCS.items():
        assert rpcs[rpc] == sig, rpc


def test_health_status_and_failed_load(backend):
    assert backend.Health(pb.HealthMessage()).message == b"OK"
    assert backend.Status(pb.HealthMessage()).state == 0  # UNINITIALIZED before LoadModel
    with pytest.raises(grpc.RpcError) as ei:  # pkg/grpc/server.go:50-53: a failed Load is a gRPC error
        backend.LoadModel(pb.ModelOptions(ModelFile="/nonexistent/model.gguf", VocabOnly=True))
    assert "Error loading model" in ei.value.details()


def test_vocab_only_load_and_tokenize(backend, model_files):
    r = backend.LoadModel(pb.ModelOptions(ModelFile=model_files["mini_q4_k_m"], VocabOnly=True, ContextSize=256))
    assert r.success and r.message == "Loading succeeded"
    t = backend.TokenizeString(pb.PredictOptions(Prompt="LocalAI is a drop-in replacement REST API"))
    from localai_b200 import capi
    with capi.Engine(model_files["mini_q4_k_m"], vocab_only=True) as e:
        want = e.tokenize("LocalAI is a drop-in replacement REST API")
    assert t.length == len(want) and list(t.tokens) == want
    assert backend.Status(pb.HealthMessage()).state == 2  # READY


@pytest.mark.gpu
def test_predict_stream_embedding_grammar_metrics(backend, model_files):
    from oracle import oracle as O

    r = backend.LoadModel(pb.ModelOptions(ModelFile=model_files["mini_q4_k_m"], ContextSize=512, NBatch=64, Embeddings=True))
    assert r.success, r.message
    prompt = "LocalAI is a drop-in replacement REST API"
    opts = dict(Prompt=prompt, Tokens=24, Temperature=0.0, IgnoreEOS=True, TopK=40, TopP=0.95)
    one = backend.Predict(pb.PredictOptions(**opts))
    ids = list(backend.TokenizeString(pb.PredictOptions(Prompt=prompt)).tokens)
    assert one.tokens == 24 and one.prompt_tokens == len(ids) and len(one.message) > 0
    parts = list(backend.PredictStream(pb.PredictOptions(**opts)))
    assert b"".join(p.message for p in parts) == one.message
    assert parts[-1].tokens == 24 and parts[-1].prompt_tokens == len(ids)      # grpc-server.cpp:2355-2358
    assert all(p.tokens == 0 for p in parts[:-1])
    # the bytes on the wire are the pieces of the oracle's greedy continuation
    want, _, _ = O.OracleModel(model_files["mini_q4_k_m"], n_ctx=256).generate_greedy(ids, 24)
    from localai_b200 import capi
    with capi.Engine(model_files["mini_q4_k_m"], vocab_only=True) as e:
        assert one.message == b"".join(e.token_to_piece(int(t)) for t in want)
    # stop words end the request and are not part of the text
    cut = one.message[6:10].decode("utf-8", "ignore")
    if cut and cut in one.message.decode("utf-8", "ignore"):
        st = backend.Predict(pb.PredictOptions(StopPrompts=[cut], **opts))
        assert cut.encode() not in st.message and len(st.message) < len(one.message)
    # Embedding: one text -> one vector of n_embd floats, the same for the same input (core/http/app_test.go:856-881)
    e1 = backend.Embedding(pb.PredictOptions(Embeddings="hello world"))
    e2 = backend.Embedding(pb.PredictOptions(Embeddings="hello world"))
    assert len(e1.embeddings) == 512 and list(e1.embeddings) == list(e2.embeddings)  # n_embd of the mini model
    assert np.isfinite(np.asarray(e1.embeddings)).all()
    # Grammar (backend.proto:98; every tools / function-call request carries one): the output fits the grammar
    g = 'root ::= ("yes" | "no") " " [0-9]+'
    gr = backend.Predict(pb.PredictOptions(Prompt=prompt, Tokens=12, Temperature=0.8, Seed=3, TopK=40, TopP=0.95, Grammar=g))
    n_ok, _ = capi.grammar_match(g, gr.message)
    assert len(gr.message) > 0 and n_ok == len(gr.message), gr.message  # every byte sent fits the grammar
    with pytest.raises(grpc.RpcError):  # a grammar that does not parse fails the request instead of running unconstrained
        backend.Predict(pb.PredictOptions(Prompt=prompt, Tokens=4, Grammar="root ::= ("))
    m = backend.GetMetrics(pb.MetricsRequest())
    assert m.tokens_generated > 0 and m.prompt_tokens_processed > 0
    assert backend.Status(pb.HealthMessage()).memory.total > 0


def test_option_mapping_equals_the_references_own_parse_options():
    """PredictOptions -> request fields: pyshim's request_kwargs (the Python twin of the Go shim's mapping) against the
    json that parse_options of the reference's server makes of the same message -- the function compiled from the
    reference's source with a stand-in for the generated protobuf class (oracle/options_harness.cc, oracle/_ref).
    Deliberate differences, all on values LocalAI's core never sends because it fills in its defaults first
    (core/config/backend_config.go:286-320): a zero repeat penalty / typical-p / tail-free z means "off" here, where the
    reference would divide by zero or truncate to nothing."""
    import ctypes
    import json

    from pyshim.server import request_kwargs

    path = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle", "_ref", "liboptions_ref.so")
    if not os.path.exists(path):
        pytest.skip("oracle/_ref/liboptions_ref.so was not built (no /root/reference on this machine)")

    class RefOpts(ctypes.Structure):
        _fields_ = ([(n, ctypes.c_int32) for n in ("tokens", "topk", "repeat", "mirostat", "nkeep", "seed")] +
                    [(n, ctypes.c_float) for n in ("topp", "tailfreesamplingz", "typicalp", "temperature", "penalty", "frequencypenalty",
                                                   "presencepenalty", "mirostattau", "mirostateta")] +
                    [(n, ctypes.c_int32) for n in ("promptcacheall", "penalizenl", "ignoreeos", "embeddings")] +
                    [(n, ctypes.c_char_p) for n in ("grammar", "prompt", "correlationid")] +
                    [("stopprompts", ctypes.POINTER(ctypes.c_char_p)), ("n_stopprompts", ctypes.c_int32)])

    R = ctypes.CDLL(path)
    R.ref_parse_options.argtypes = [ctypes.c_int, ctypes.POINTER(RefOpts), ctypes.c_char_p, ctypes.c_int]
    rng = np.random.default_rng(53)
    f32 = lambda x: float(np.float32(x))  # noqa: E731
    for trial in range(300):
        zeros = trial % 5 == 0  # every fifth message leaves the float knobs at the proto default
        o = pb.PredictOptions(
            Prompt="p%d" % trial, Tokens=int(rng.integers(0, 3)) * 64, TopK=int(rng.integers(0, 100)), Seed=int(rng.integers(-1, 1000)),
            TopP=0 if zeros else f32(rng.uniform(0.1, 1)), Temperature=f32(rng.uniform(0, 2)), Repeat=int(rng.integers(-1, 128)),
            Penalty=0 if zeros else f32(rng.uniform(1, 1.5)), FrequencyPenalty=f32(rng.uniform(0, 1)), PresencePenalty=f32(rng.uniform(0, 1)),
            Mirostat=int(rng.integers(0, 3)), MirostatTAU=f32(rng.uniform(1, 8)), MirostatETA=f32(rng.uniform(0.05, 0.5)),
            TypicalP=0 if zeros else f32(rng.uniform(0.2, 1)), TailFreeSamplingZ=0 if zeros else f32(rng.uniform(0.2, 1)),
            PenalizeNL=bool(rng.integers(0, 2)), NKeep=int(rng.integers(-1, 64)), IgnoreEOS=bool(rng.integers(0, 2)),
            PromptCacheAll=bool(rng.integers(0, 2)), Grammar="" if trial % 3 else 'root ::= "a"', StopPrompts=["</s>", "###"][:trial % 3])
        stops = [s.encode() for s in o.StopPrompts]
        arr = (ctypes.c_char_p * max(1, len(stops)))(*stops)
        r = RefOpts(o.Tokens, o.TopK, o.Repeat, o.Mirostat, o.NKeep, o.Seed, o.TopP, o.TailFreeSamplingZ, o.TypicalP, o.Temperature, o.Penalty,
                    o.FrequencyPenalty, o.PresencePenalty, o.MirostatTAU, o.MirostatETA, int(o.PromptCacheAll), int(o.PenalizeNL),
                    int(o.IgnoreEOS), 0, o.Grammar.encode(), o.Prompt.encode(), b"", arr, len(stops))
        buf = ctypes.create_string_buffer(1 << 16)
        assert R.ref_parse_options(1, ctypes.byref(r), buf, len(buf)) > 0
        want = json.loads(buf.value)
        got = request_kwargs(o)
        assert want["stream"] is True and want["prompt"] == o.Prompt
        assert want["n_predict"] == (got["n_predict"] if got["n_predict"] != 0 else -1)  # 0 = until EOS / context in both, spelled -1 there
        for jk, kk in (("top_k", "top_k"), ("seed", "seed"), ("repeat_last_n", "repeat_last_n"), ("mirostat", "mirostat"), ("n_keep", "n_keep"),
                       ("ignore_eos", "ignore_eos"), ("cache_prompt", "cache_prompt"), ("penalize_nl", "penalize_nl"), ("stop", "stop")):
            assert want[jk] == got[kk], (jk, want[jk], got[kk])
        for jk, kk in (("top_p", "top_p"), ("temperature", "temperature"), ("frequency_penalty", "frequency_penalty"),
                       ("presence_penalty", "presence_penalty"), ("mirostat_tau", "mirostat_tau"), ("mirostat_eta", "mirostat_eta")):
            assert f32(want[jk]) == f32(got[kk]), (jk, want[jk], got[kk])
        assert want["grammar"] == (got["grammar"] or "")
        for jk, kk in (("repeat_penalty", "repeat_penalty"), ("typical_p", "typical_p"), ("tfs_z", "tfs_z")):
            if f32(want[jk]) != 0.0:
                assert f32(want[jk]) == f32(got[kk]), (jk, want[jk], got[kk])
            else:
                assert got[kk] == 1.0, (jk, got[kk])  # the deliberate difference: zero means "off"


def test_load_option_mapping_equals_the_references_own_params_parse(monkeypatch):
    """ModelOptions -> load options: pyshim's load_kwargs (the twin of the Go shim's Load) against the gpt_params that
    params_parse of the reference's server fills from the same message (compiled from the reference's source,
    oracle/options_harness.cc).  A zero ContextSize / NBatch -- which LocalAI's core never sends -- means this engine's
    default here and llama.cpp's there; the YaRN fields are "0 = default" here and "left at llama.cpp's default" there."""
    import ctypes
    import json

    from pyshim.server import load_kwargs

    path = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle", "_ref", "liboptions_ref.so")
    if not os.path.exists(path):
        pytest.skip("oracle/_ref/liboptions_ref.so was not built (no /root/reference on this machine)")

    class RefModel(ctypes.Structure):
        _fields_ = ([(n, ctypes.c_char_p) for n in ("modelfile", "mmproj", "tensorsplit", "maingpu", "loraadapter", "lorabase", "ropescaling")] +
                    [(n, ctypes.c_int32) for n in ("contextsize", "threads", "ngpulayers", "nbatch", "mlock", "mmap", "flashattention", "nokvoffload",
                                                   "embeddings")] +
                    [(n, ctypes.c_float) for n in ("lorascale", "yarnextfactor", "yarnattnfactor", "yarnbetafast", "yarnbetaslow", "ropefreqbase",
                                                   "ropefreqscale")])

    R = ctypes.CDLL(path)
    R.ref_params_parse.argtypes = [ctypes.POINTER(RefModel), ctypes.c_char_p, ctypes.c_int]
    rng = np.random.default_rng(59)
    f32 = lambda x: float(np.float32(x))  # noqa: E731
    defaults = {"yarn_ext_factor": -1.0, "yarn_attn_factor": 1.0, "yarn_beta_fast": 32.0, "yarn_beta_slow": 1.0}
    for trial in range(200):
        if trial % 2:
            monkeypatch.setenv("LLAMACPP_PARALLEL", str(1 + trial % 7))
        else:
            monkeypatch.delenv("LLAMACPP_PARALLEL", raising=False)
        z = trial % 4 == 0
        o = pb.ModelOptions(
            ModelFile="/models/m%d.gguf" % trial, ContextSize=0 if trial % 9 == 0 else int(rng.integers(1, 64)) * 128,
            NBatch=0 if trial % 11 == 0 else int(rng.integers(1, 16)) * 128, MainGPU=["", "0", "3", "7"][trial % 4],
            Embeddings=bool(trial % 3 == 0), RopeScaling=["", "none", "yarn", "linear", "something"][trial % 5],
            YarnExtFactor=0 if z else f32(rng.uniform(0.1, 1)), YarnAttnFactor=0 if z else f32(rng.uniform(0.5, 2)),
            YarnBetaFast=0 if z else f32(rng.uniform(8, 64)), YarnBetaSlow=0 if z else f32(rng.uniform(0.5, 4)),
            RopeFreqBase=0 if z else f32(rng.uniform(1e4, 1e6)), RopeFreqScale=0 if z else f32(rng.uniform(0.1, 1)),
            Threads=int(rng.integers(1, 64)), NGPULayers=99, MMap=True, FlashAttention=bool(trial % 2))
        r = RefModel(o.ModelFile.encode(), b"", b"", o.MainGPU.encode(), b"", b"", o.RopeScaling.encode(), o.ContextSize, o.Threads, o.NGPULayers,
                     o.NBatch, int(o.MLock), int(o.MMap), int(o.FlashAttention), int(o.NoKVOffload), int(o.Embeddings), 0.0, o.YarnExtFactor,
                     o.YarnAttnFactor, o.YarnBetaFast, o.YarnBetaSlow, o.RopeFreqBase, o.RopeFreqScale)
        buf = ctypes.create_string_buffer(1 << 14)
        assert R.ref_params_parse(ctypes.byref(r), buf, len(buf)) > 0
        want = json.loads(buf.value)
        got = load_kwargs(o)
        assert want["model"] == o.ModelFile
        if o.ContextSize:
            assert want["n_ctx"] == got["n_ctx"]
        if o.NBatch:
            assert want["n_batch"] == got["n_batch"]
        assert want["n_parallel"] == got["n_parallel"] and want["main_gpu"] == got["device"] and want["embedding"] == got["embeddings"]
        assert {0: "none", 1: "linear", 2: "yarn"}[want["rope_scaling_type"]] == got["rope_scaling"]
        for k in ("rope_freq_base", "rope_freq_scale"):
            assert f32(want[k]) == f32(got[k]), k
        for k, dflt in defaults.items():
            if got[k] == 0:
                assert f32(want[k]) == f32(dflt), k  # untouched there = "0: default" here
            else:
                assert f32(want[k]) == f32(got[k]), k
