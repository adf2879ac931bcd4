"""localai_b200 -- B200-native (sm_100a) drop-in for the engine of LocalAI's llama-cpp backend.

The product is libb200llm.so (localai_b200/csrc -> localai_b200/lib), bound through the C ABI in
include/b200llm.h.  This package holds the ctypes binding (capi), the host-side mirror of LocalAI's
`grpc.LLM` backend interface (backend) and a grpcio server speaking backend.proto (grpc_server).
"""
from . import capi  # noqa: F401
from .capi import B2Error, Engine  # noqa: F401

__all__ = ["capi", "Engine", "B2Error"]
