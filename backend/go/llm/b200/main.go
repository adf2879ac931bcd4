package main

// B200-native LLM backend for LocalAI.  Drop this directory into LocalAI's backend/go/llm/ and build it like the
// other Go backends (see INTEGRATION.md).  Mirrors backend/go/llm/bert/main.go:15-21 of the reference.
// NOTE: the Go toolchain is absent from the image this repository is built in, so this file is compiled and
// exercised only inside a LocalAI checkout; the same C ABI is exercised here through ctypes (tests/).

import (
	"flag"

	grpc "github.com/mudler/LocalAI/pkg/grpc"
)

var (
	addr = flag.String("addr", "localhost:50051", "the address to connect to")
)

func main() {
	flag.Parse()

	if err := grpc.StartServer(*addr, &LLM{}); err != nil {
		panic(err)
	}
}
