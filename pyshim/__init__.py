"""Python face of backend.proto over the C ABI (test infrastructure: Go / protoc are absent from this image)."""
