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
