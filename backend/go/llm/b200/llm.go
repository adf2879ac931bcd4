package main

// LLM implements grpc.LLM (pkg/grpc/interface.go:7-27) on top of libb200llm.so (include/b200llm.h).
// It embeds base.Base, not base.SingleThread: Locking() is false, so concurrent RPCs reach the engine, whose own
// thread batches them (continuous batching); LocalAI must be started with --parallel-requests for the client side
// not to serialise (pkg/grpc/client.go:103-106).

/*
#cgo CFLAGS: -I${SRCDIR}/../../../../include
#cgo LDFLAGS: -L${SRCDIR} -lb200llm -Wl,-rpath,$ORIGIN
#include <stdlib.h>
#include <string.h>
#include "b200llm.h"
*/
import "C"

import (
	"fmt"
	"os"
	"strconv"
	"unsafe"

	"github.com/mudler/LocalAI/pkg/grpc/base"
	pb "github.com/mudler/LocalAI/pkg/grpc/proto"
)

type LLM struct {
	base.Base
	eng *C.b2_engine
}

func lastError() error { return fmt.Errorf("b200llm: %s", C.GoString(C.b2_last_error())) }

// Load <- ModelOptions (backend.proto:162-235); replaces grpc-server.cpp:2312-2332.
func (l *LLM) Load(opts *pb.ModelOptions) error {
	var o C.b2_load_opts
	o.struct_size = C.uint32_t(unsafe.Sizeof(o))
	path := C.CString(opts.ModelFile)
	defer C.free(unsafe.Pointer(path))
	o.model_path = path
	o.n_ctx = C.int32_t(opts.ContextSize)
	o.n_batch = C.int32_t(opts.NBatch)
	o.n_parallel = 1
	if v, err := strconv.Atoi(os.Getenv("LLAMACPP_PARALLEL")); err == nil && v > 0 { // grpc-server.cpp:2225-2231
		o.n_parallel = C.int32_t(v)
	}
	if v, err := strconv.Atoi(opts.MainGPU); err == nil {
		o.device = C.int32_t(v)
	}
	if opts.Embeddings {
		o.embeddings = 1
	}
	if opts.VocabOnly {
		o.vocab_only = 1
	}
	o.rope_freq_base = C.float(opts.RopeFreqBase)
	o.rope_freq_scale = C.float(opts.RopeFreqScale)
	l.eng = C.b2_create(&o)
	if l.eng == nil {
		return lastError()
	}
	return nil
}

func (l *LLM) request(opts *pb.PredictOptions, stream bool, embedding bool) (C.int64_t, func(), error) {
	var r C.b2_request
	r.struct_size = C.uint32_t(unsafe.Sizeof(r))
	text := opts.Prompt
	if embedding {
		text = opts.Embeddings
	}
	cprompt := C.CString(text)
	r.prompt = cprompt
	r.prompt_len = C.size_t(len(text))
	frees := []unsafe.Pointer{unsafe.Pointer(cprompt)}
	if embedding && len(opts.EmbeddingTokens) > 0 {
		ids := (*C.int32_t)(C.malloc(C.size_t(4 * len(opts.EmbeddingTokens))))
		frees = append(frees, unsafe.Pointer(ids))
		s := unsafe.Slice(ids, len(opts.EmbeddingTokens))
		for i, t := range opts.EmbeddingTokens {
			s[i] = C.int32_t(t)
		}
		r.prompt_ids, r.n_prompt_ids = ids, C.int32_t(len(opts.EmbeddingTokens))
	}
	r.n_predict = C.int32_t(opts.Tokens) // 0 => until EOS/context, as grpc-server.cpp:2098
	r.temperature = C.float(opts.Temperature)
	r.top_k, r.top_p = C.int32_t(opts.TopK), C.float(opts.TopP)
	r.repeat_penalty, r.repeat_last_n = C.float(opts.Penalty), C.int32_t(opts.Repeat)
	r.frequency_penalty, r.presence_penalty = C.float(opts.FrequencyPenalty), C.float(opts.PresencePenalty)
	// LocalAI's default config sends mirostat 2, tau 5, eta 0.1 (core/config/backend_config.go:286-289)
	r.mirostat, r.mirostat_tau, r.mirostat_eta = C.int32_t(opts.Mirostat), C.float(opts.MirostatTAU), C.float(opts.MirostatETA)
	r.typical_p, r.tfs_z = C.float(opts.TypicalP), C.float(opts.TailFreeSamplingZ)
	r.seed = C.int32_t(opts.Seed)
	if opts.IgnoreEOS {
		r.ignore_eos = 1
	}
	if opts.PromptCacheAll {
		r.cache_prompt = 1
	}
	r.n_keep = C.int32_t(opts.NKeep)
	if embedding {
		r.embedding_mode = 1
	}
	if stream {
		r.stream = 1
	}
	if n := len(opts.StopPrompts); n > 0 {
		arr := (**C.char)(C.malloc(C.size_t(n) * C.size_t(unsafe.Sizeof(uintptr(0)))))
		frees = append(frees, unsafe.Pointer(arr))
		s := unsafe.Slice(arr, n)
		for i, w := range opts.StopPrompts {
			s[i] = C.CString(w)
			frees = append(frees, unsafe.Pointer(s[i]))
		}
		r.stop, r.n_stop = arr, C.int32_t(n)
	}
	id := C.b2_submit(l.eng, &r)
	release := func() {
		for _, p := range frees {
			C.free(p)
		}
	}
	if id < 0 {
		release()
		return 0, nil, lastError()
	}
	return id, release, nil
}

// drain polls events until the final one; f receives every text piece.
func (l *LLM) drain(id C.int64_t, f func(string)) error {
	defer C.b2_release(l.eng, id)
	for {
		var ev C.b2_event
		ev.struct_size = C.uint32_t(unsafe.Sizeof(ev))
		if rc := C.b2_poll(l.eng, id, &ev, -1); rc != 0 {
			return lastError()
		}
		if ev.flags&C.B2_EV_ERROR != 0 {
			return fmt.Errorf("b200llm: %s", C.GoStringN(ev.text, C.int(ev.text_len)))
		}
		if ev.text_len > 0 {
			f(C.GoStringN(ev.text, C.int(ev.text_len)))
		}
		if ev.flags&C.B2_EV_FINAL != 0 {
			return nil
		}
	}
}

// Predict <- grpc-server.cpp:2380-2407
func (l *LLM) Predict(opts *pb.PredictOptions) (string, error) {
	id, release, err := l.request(opts, false, false)
	if err != nil {
		return "", err
	}
	defer release()
	out := ""
	err = l.drain(id, func(s string) { out += s })
	return out, err
}

// PredictStream <- grpc-server.cpp:2333-2377; the callee closes the channel (pkg/grpc/base/base.go:43-46).
func (l *LLM) PredictStream(opts *pb.PredictOptions, results chan string) error {
	id, release, err := l.request(opts, true, false)
	if err != nil {
		close(results)
		return err
	}
	go func() {
		defer release()
		defer close(results)
		_ = l.drain(id, func(s string) { results <- s })
	}()
	return nil
}

// Embeddings <- grpc-server.cpp:2410-2432
func (l *LLM) Embeddings(opts *pb.PredictOptions) ([]float32, error) {
	id, release, err := l.request(opts, false, true)
	if err != nil {
		return nil, err
	}
	defer release()
	var ev C.b2_event
	ev.struct_size = C.uint32_t(unsafe.Sizeof(ev))
	for ev.flags&C.B2_EV_FINAL == 0 {
		if rc := C.b2_poll(l.eng, id, &ev, -1); rc != 0 {
			return nil, lastError()
		}
	}
	defer C.b2_release(l.eng, id)
	out := make([]float32, int(ev.n_embd))
	if len(out) == 0 {
		return out, nil
	}
	if n := C.b2_embed_result(l.eng, id, (*C.float)(unsafe.Pointer(&out[0])), C.int(len(out))); n < 0 {
		return nil, lastError()
	}
	return out, nil
}

// TokenizeString: the reference C++ backend does not implement it (SURVEY.md 3.5); this one does.
func (l *LLM) TokenizeString(opts *pb.PredictOptions) (pb.TokenizationResponse, error) {
	text := C.CString(opts.Prompt)
	defer C.free(unsafe.Pointer(text))
	cap := len(opts.Prompt) + 16
	buf := make([]int32, cap)
	n := C.b2_tokenize(l.eng, text, C.size_t(len(opts.Prompt)), 1, 1, (*C.int32_t)(unsafe.Pointer(&buf[0])), C.int(cap))
	if n < 0 {
		return pb.TokenizationResponse{}, lastError()
	}
	return pb.TokenizationResponse{Length: int32(n), Tokens: buf[:n]}, nil
}
