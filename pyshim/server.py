#!/usr/bin/env python3
"""backend.proto `Backend` service over the C ABI (include/b200llm.h), the shape of backend/cpp/llama/grpc-server.cpp.

Test infrastructure: the deliverable for LocalAI is the Go backend in backend/go/llm/b200 (same C calls through cgo); Go
and protoc are absent from this image, so this file is the face the wire-level conformance test talks to
(tests/test_grpc_conformance.py).  Launched like every LocalAI backend: `python -m pyshim.server --addr 127.0.0.1:PORT`
(pkg/model/process.go:83-95).

RPC -> C ABI:
  Health         -> "OK"                                               (grpc-server.cpp:2305-2309)
  LoadModel      -> b2_create                                          (:2312-2332; message texts as pkg/grpc/server.go:45-55)
  Predict        -> b2_submit(stream=0) + b2_poll                      (:2380-2407)  Reply{message, tokens, prompt_tokens}
  PredictStream  -> b2_submit(stream=1) + b2_poll per piece            (:2333-2377)  the last Reply carries the counts
  Embedding      -> b2_submit(embedding_mode) + b2_embed_result        (:2410-2432)
  TokenizeString -> b2_tokenize                                        (pkg/grpc/server.go:161-180)
  Status         -> READY / BUSY + device bytes                        (pkg/grpc/server.go:182-196)
  GetMetrics     -> b2_metrics                                         (:2434-2457)
"""
from __future__ import annotations

import argparse
import os
import sys
import threading
from concurrent import futures

import grpc

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from localai_b200 import capi  # noqa: E402
from pyshim import backend_pb as pb  # noqa: E402


def request_kwargs(o) -> dict:
    """PredictOptions -> b2_request fields, the mapping of parse_options (grpc-server.cpp:2068-2137) and of
    backend/go/llm/b200/llm.go."""
    return dict(
        n_predict=o.Tokens,                # 0 => until EOS / context (:2098)
        temperature=o.Temperature, top_k=o.TopK, top_p=o.TopP,
        min_p=0.05,                        # not in PredictOptions: the reference runs with llama.cpp's default (:639)
        seed=o.Seed, ignore_eos=o.IgnoreEOS, cache_prompt=o.PromptCacheAll, n_keep=o.NKeep,
        stop=list(o.StopPrompts), repeat_penalty=o.Penalty if o.Penalty else 1.0, repeat_last_n=o.Repeat,
        frequency_penalty=o.FrequencyPenalty, presence_penalty=o.PresencePenalty, mirostat=o.Mirostat,
        mirostat_tau=o.MirostatTAU, mirostat_eta=o.MirostatETA, typical_p=o.TypicalP or 1.0,
        tfs_z=o.TailFreeSamplingZ or 1.0, grammar=o.Grammar or None, penalize_nl=o.PenalizeNL)


def load_kwargs(o, environ=os.environ) -> dict:
    """ModelOptions -> b2_load_opts fields, the mapping of params_parse (grpc-server.cpp:2205-2300) and of
    backend/go/llm/b200/llm.go.  0 / "" keep the engine's (= the reference's) defaults."""
    n_par = int(environ.get("LLAMACPP_PARALLEL", "1") or 1)  # :2225-2231
    return dict(
        n_ctx=o.ContextSize or 1024, n_batch=o.NBatch or 512, n_parallel=max(1, n_par),
        device=int(o.MainGPU) if o.MainGPU.strip().isdigit() else 0, embeddings=o.Embeddings,
        rope_freq_base=o.RopeFreqBase, rope_freq_scale=o.RopeFreqScale,
        # :2279-2281: "none" / "yarn" / anything else (the empty default too) = linear
        rope_scaling=o.RopeScaling if o.RopeScaling in ("none", "yarn") else "linear",
        yarn_ext_factor=o.YarnExtFactor, yarn_attn_factor=o.YarnAttnFactor, yarn_beta_fast=o.YarnBetaFast,
        yarn_beta_slow=o.YarnBetaSlow)


class Backend:
    def __init__(self):
        self.engine: capi.Engine | None = None
        self.lock = threading.Lock()
        self.inflight = 0

    # ---- helpers
    def _need_model(self, ctx):
        if self.engine is None:
            ctx.abort(grpc.StatusCode.FAILED_PRECONDITION, "no model loaded")
        return self.engine

    def _track(self, d):
        with self.lock:
            self.inflight += d

    # ---- RPCs
    def Health(self, req, ctx):
        return pb.Reply(message=b"OK")

    def LoadModel(self, req, ctx):
        try:
            eng = capi.Engine(req.ModelFile or req.Model, vocab_only=req.VocabOnly, **load_kwargs(req))
        except Exception as ex:  # pkg/grpc/server.go:50-53: Result{success:false} AND a gRPC error
            ctx.set_code(grpc.StatusCode.UNKNOWN)
            ctx.set_details(f"Error loading model: {ex}")
            return pb.Result(message=f"Error loading model: {ex}", success=False)
        old, self.engine = self.engine, eng
        if old is not None:
            old.close()
        return pb.Result(message="Loading succeeded", success=True)

    def _run(self, eng, req, stream):
        rid = eng.submit(prompt=req.Prompt, stream=stream, **request_kwargs(req))
        self._track(+1)
        try:
            for ev in eng.events(rid):
                yield ev
        finally:
            self._track(-1)
            eng.release(rid)

    def Predict(self, req, ctx):
        eng = self._need_model(ctx)
        try:
            text, last = b"", None
            for ev in self._run(eng, req, False):
                if ev["flags"] & capi.EV_ERROR:
                    return pb.Reply()  # the reference answers OK + empty on failure (:2401-2404)
                text += ev["text"]
                last = ev
            return pb.Reply(message=text, tokens=last["n_decoded"], prompt_tokens=last["n_prompt_tokens"])
        except capi.B2Error as ex:
            ctx.abort(grpc.StatusCode.INVALID_ARGUMENT, str(ex))

    def PredictStream(self, req, ctx):
        eng = self._need_model(ctx)
        try:
            for ev in self._run(eng, req, True):
                if ev["flags"] & capi.EV_ERROR:
                    return
                final = bool(ev["flags"] & capi.EV_FINAL)
                if ev["text"] or final:  # partial replies carry no counts (send_partial_response, :1200-1237)
                    yield pb.Reply(message=ev["text"], tokens=ev["n_decoded"] if final else 0,
                                   prompt_tokens=ev["n_prompt_tokens"] if final else 0)
        except capi.B2Error as ex:
            ctx.abort(grpc.StatusCode.INVALID_ARGUMENT, str(ex))

    def Embedding(self, req, ctx):
        eng = self._need_model(ctx)
        try:
            kw = dict(prompt=req.Embeddings, embedding_mode=True, stream=False)
            if len(req.EmbeddingTokens):
                kw["prompt_ids"] = list(req.EmbeddingTokens)
            rid = eng.submit(**kw)
            try:
                for ev in eng.events(rid):
                    if ev["flags"] & capi.EV_ERROR:
                        return pb.EmbeddingResult()
                return pb.EmbeddingResult(embeddings=eng.embed_result(rid).tolist())
            finally:
                eng.release(rid)
        except capi.B2Error as ex:
            ctx.abort(grpc.StatusCode.INVALID_ARGUMENT, str(ex))

    def TokenizeString(self, req, ctx):
        eng = self._need_model(ctx)
        ids = eng.tokenize(req.Prompt, add_bos=True, parse_special=True)
        return pb.TokenizationResponse(length=len(ids), tokens=ids)

    def Status(self, req, ctx):
        r = pb.StatusResponse()
        if self.engine is None:
            r.state = 0  # UNINITIALIZED
            return r
        r.state = 1 if self.inflight > 0 else 2  # BUSY / READY
        w = int(self.engine.info.weight_bytes_device)
        r.memory.total = w
        r.memory.breakdown["gpu_weights"] = w
        return r

    def GetMetrics(self, req, ctx):
        if self.engine is None:
            return pb.MetricsResponse()
        m = self.engine.metrics()
        return pb.MetricsResponse(slot_id=m["slot_id"], prompt_json_for_slot="", tokens_per_second=m["tokens_per_second"],
                                  tokens_generated=m["tokens_generated"], prompt_tokens_processed=m["prompt_tokens_processed"])


def make_server(addr: str, max_workers: int = 64):
    impl = Backend()
    handlers = {}
    for rpc, (req, resp, streaming) in pb.RPCS.items():
        mk = grpc.unary_stream_rpc_method_handler if streaming else grpc.unary_unary_rpc_method_handler
        handlers[rpc] = mk(getattr(impl, rpc), request_deserializer=pb.cls(req).FromString,
                           response_serializer=pb.cls(resp).SerializeToString)
    server = grpc.server(futures.ThreadPoolExecutor(max_workers=max_workers),
                         options=[("grpc.max_send_message_length", 50 << 20), ("grpc.max_receive_message_length", 50 << 20)])
    server.add_generic_rpc_handlers((grpc.method_handlers_generic_handler(pb.SERVICE, handlers),))
    port = server.add_insecure_port(addr)
    return server, port, impl


class Client:
    """What pkg/grpc/client.go does, for the tests: one callable per RPC."""

    def __init__(self, addr: str):
        self.channel = grpc.insecure_channel(addr)
        for rpc, (req, resp, streaming) in pb.RPCS.items():
            mk = self.channel.unary_stream if streaming else self.channel.unary_unary
            setattr(self, rpc, mk(f"/{pb.SERVICE}/{rpc}", request_serializer=pb.cls(req).SerializeToString,
                                  response_deserializer=pb.cls(resp).FromString))

    def close(self):
        self.channel.close()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--addr", default="127.0.0.1:50051")
    a = ap.parse_args()
    server, port, _ = make_server(a.addr)
    server.start()
    print(f"b200 backend listening on {a.addr}", file=sys.stderr, flush=True)
    server.wait_for_termination()


if __name__ == "__main__":
    main()
