"""backend.proto (reference: backend/backend.proto) as runtime-built protobuf descriptors.

protoc and grpc_tools are not on this image, so the messages the LLM path uses are declared here as field tables
(name, type, number[, repeated]) with the numbers of backend.proto -- the wire contract a LocalAI core speaks
(pkg/grpc/client.go).  Only the RPCs of the hot path are declared: Health, LoadModel, Predict, PredictStream, Embedding,
TokenizeString, Status, GetMetrics (backend.proto:10-31).
"""
from google.protobuf import descriptor_pb2, descriptor_pool, message_factory

_F = descriptor_pb2.FieldDescriptorProto
_TYPES = {"string": _F.TYPE_STRING, "bytes": _F.TYPE_BYTES, "int32": _F.TYPE_INT32, "int64": _F.TYPE_INT64,
          "uint64": _F.TYPE_UINT64, "float": _F.TYPE_FLOAT, "bool": _F.TYPE_BOOL}

# (name, type, number) or (name, type, number, "repeated"); a type that is not a scalar names another message
MESSAGES = {
    "HealthMessage": [],                                                      # backend.proto:66
    "MetricsRequest": [],                                                     # :34
    "MetricsResponse": [("slot_id", "int32", 1), ("prompt_json_for_slot", "string", 2), ("tokens_per_second", "float", 3),
                        ("tokens_generated", "int32", 4), ("prompt_tokens_processed", "int32", 5)],  # :36-42
    "Message": [("role", "string", 1), ("content", "string", 2)],             # :321-324
    "PredictOptions": [                                                       # :69-117
        ("Prompt", "string", 1), ("Seed", "int32", 2), ("Threads", "int32", 3), ("Tokens", "int32", 4), ("TopK", "int32", 5),
        ("Repeat", "int32", 6), ("Batch", "int32", 7), ("NKeep", "int32", 8), ("Temperature", "float", 9), ("Penalty", "float", 10),
        ("F16KV", "bool", 11), ("DebugMode", "bool", 12), ("StopPrompts", "string", 13, "repeated"), ("IgnoreEOS", "bool", 14),
        ("TailFreeSamplingZ", "float", 15), ("TypicalP", "float", 16), ("FrequencyPenalty", "float", 17),
        ("PresencePenalty", "float", 18), ("Mirostat", "int32", 19), ("MirostatETA", "float", 20), ("MirostatTAU", "float", 21),
        ("PenalizeNL", "bool", 22), ("LogitBias", "string", 23), ("MLock", "bool", 25), ("MMap", "bool", 26),
        ("PromptCacheAll", "bool", 27), ("PromptCacheRO", "bool", 28), ("Grammar", "string", 29), ("MainGPU", "string", 30),
        ("TensorSplit", "string", 31), ("TopP", "float", 32), ("PromptCachePath", "string", 33), ("Debug", "bool", 34),
        ("EmbeddingTokens", "int32", 35, "repeated"), ("Embeddings", "string", 36), ("RopeFreqBase", "float", 37),
        ("RopeFreqScale", "float", 38), ("NegativePromptScale", "float", 39), ("NegativePrompt", "string", 40),
        ("NDraft", "int32", 41), ("Images", "string", 42, "repeated"), ("UseTokenizerTemplate", "bool", 43),
        ("Messages", "Message", 44, "repeated"), ("Videos", "string", 45, "repeated"), ("Audios", "string", 46, "repeated"),
        ("CorrelationId", "string", 47)],
    "Reply": [("message", "bytes", 1), ("tokens", "int32", 2), ("prompt_tokens", "int32", 3)],  # :120-124
    "ModelOptions": [                                                         # :126-203
        ("Model", "string", 1), ("ContextSize", "int32", 2), ("Seed", "int32", 3), ("NBatch", "int32", 4), ("F16Memory", "bool", 5),
        ("MLock", "bool", 6), ("MMap", "bool", 7), ("VocabOnly", "bool", 8), ("LowVRAM", "bool", 9), ("Embeddings", "bool", 10),
        ("NUMA", "bool", 11), ("NGPULayers", "int32", 12), ("MainGPU", "string", 13), ("TensorSplit", "string", 14),
        ("Threads", "int32", 15), ("LibrarySearchPath", "string", 16), ("RopeFreqBase", "float", 17), ("RopeFreqScale", "float", 18),
        ("RMSNormEps", "float", 19), ("NGQA", "int32", 20), ("ModelFile", "string", 21), ("Device", "string", 22),
        ("UseTriton", "bool", 23), ("ModelBaseName", "string", 24), ("UseFastTokenizer", "bool", 25), ("PipelineType", "string", 26),
        ("SchedulerType", "string", 27), ("CUDA", "bool", 28), ("CFGScale", "float", 29), ("IMG2IMG", "bool", 30),
        ("CLIPModel", "string", 31), ("CLIPSubfolder", "string", 32), ("CLIPSkip", "int32", 33), ("Tokenizer", "string", 34),
        ("LoraBase", "string", 35), ("LoraAdapter", "string", 36), ("NoMulMatQ", "bool", 37), ("AudioPath", "string", 38),
        ("DraftModel", "string", 39), ("Quantization", "string", 40), ("MMProj", "string", 41), ("LoraScale", "float", 42),
        ("RopeScaling", "string", 43), ("YarnExtFactor", "float", 44), ("YarnAttnFactor", "float", 45), ("YarnBetaFast", "float", 46),
        ("YarnBetaSlow", "float", 47), ("ControlNet", "string", 48), ("Type", "string", 49), ("GPUMemoryUtilization", "float", 50),
        ("TrustRemoteCode", "bool", 51), ("EnforceEager", "bool", 52), ("SwapSpace", "int32", 53), ("MaxModelLen", "int32", 54),
        ("TensorParallelSize", "int32", 55), ("FlashAttention", "bool", 56), ("NoKVOffload", "bool", 57)],
    "Result": [("message", "string", 1), ("success", "bool", 2)],             # :205-208
    "EmbeddingResult": [("embeddings", "float", 1, "repeated")],              # :210-212
    "TokenizationResponse": [("length", "int32", 1), ("tokens", "int32", 2, "repeated")],  # :299-302
}
# rpc name -> (request, response, server-streaming)    backend.proto:10-31
RPCS = {
    "Health": ("HealthMessage", "Reply", False), "Predict": ("PredictOptions", "Reply", False),
    "LoadModel": ("ModelOptions", "Result", False), "PredictStream": ("PredictOptions", "Reply", True),
    "Embedding": ("PredictOptions", "EmbeddingResult", False), "TokenizeString": ("PredictOptions", "TokenizationResponse", False),
    "Status": ("HealthMessage", "StatusResponse", False), "GetMetrics": ("MetricsRequest", "MetricsResponse", False),
}
SERVICE = "backend.Backend"


def _build():
    fd = descriptor_pb2.FileDescriptorProto(name="b200_backend.proto", package="backend", syntax="proto3")
    for mname, fields in MESSAGES.items():
        m = fd.message_type.add(name=mname)
        for f in fields:
            fld = m.field.add(name=f[0], number=f[2], json_name=f[0],
                              label=_F.LABEL_REPEATED if len(f) > 3 else _F.LABEL_OPTIONAL)
            if f[1] in _TYPES:
                fld.type = _TYPES[f[1]]
            else:
                fld.type, fld.type_name = _F.TYPE_MESSAGE, ".backend." + f[1]
    # MemoryUsageData { uint64 total = 1; map<string, uint64> breakdown = 2; }      backend.proto:304-307
    mu = fd.message_type.add(name="MemoryUsageData")
    mu.field.add(name="total", number=1, type=_F.TYPE_UINT64, label=_F.LABEL_OPTIONAL)
    ent = mu.nested_type.add(name="BreakdownEntry")
    ent.options.map_entry = True
    ent.field.add(name="key", number=1, type=_F.TYPE_STRING, label=_F.LABEL_OPTIONAL)
    ent.field.add(name="value", number=2, type=_F.TYPE_UINT64, label=_F.LABEL_OPTIONAL)
    mu.field.add(name="breakdown", number=2, type=_F.TYPE_MESSAGE, label=_F.LABEL_REPEATED,
                 type_name=".backend.MemoryUsageData.BreakdownEntry")
    # StatusResponse { enum State {UNINITIALIZED=0; BUSY=1; READY=2; ERROR=-1;} State state = 1; MemoryUsageData memory = 2; }   :309-319
    sr = fd.message_type.add(name="StatusResponse")
    en = sr.enum_type.add(name="State")
    for n, v in (("UNINITIALIZED", 0), ("BUSY", 1), ("READY", 2), ("ERROR", -1)):
        en.value.add(name=n, number=v)
    sr.field.add(name="state", number=1, type=_F.TYPE_ENUM, label=_F.LABEL_OPTIONAL, type_name=".backend.StatusResponse.State")
    sr.field.add(name="memory", number=2, type=_F.TYPE_MESSAGE, label=_F.LABEL_OPTIONAL, type_name=".backend.MemoryUsageData")
    svc = fd.service.add(name="Backend")
    for rpc, (req, resp, stream) in RPCS.items():
        svc.method.add(name=rpc, input_type=".backend." + req, output_type=".backend." + resp, server_streaming=stream)
    pool = descriptor_pool.DescriptorPool()
    pool.Add(fd)
    names = list(MESSAGES) + ["MemoryUsageData", "StatusResponse"]
    return {n: message_factory.GetMessageClass(pool.FindMessageTypeByName("backend." + n)) for n in names}


_classes = _build()
globals().update(_classes)


def cls(name):
    return _classes[name]
