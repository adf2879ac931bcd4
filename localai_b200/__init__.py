"""localai_b200 -- B200-native (sm_100a) drop-in for the engine of LocalAI's llama-cpp backend.

The product is libb200llm.so (localai_b200/csrc -> localai_b200/lib), bound through the C ABI in
include/b200llm.h.  This package holds the ctypes binding (capi) and the in-process replica pool (replicas).  The Go
backend that LocalAI loads lives in backend/go/llm/b200; the grpcio face of backend.proto used by the conformance test
is pyshim/ (test infrastructure, not part of the product).
"""
from . import capi  # noqa: F401
from .capi import B2Error, Engine  # noqa: F401

__all__ = ["capi", "Engine", "B2Error"]
