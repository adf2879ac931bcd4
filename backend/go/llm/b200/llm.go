package main

// LLM implements grpc.LLM (pkg/grpc/interface.go:7-27) on top of libb200llm.so (include/b200llm.h).
// It embeds base.Base, not base.SingleThread: Locking() is false, so concurrent RPCs reach the engine, whose own
// thread batches them (continuous batching); LocalAI must be started with --parallel-requests for the client side
// not to serialise (pkg/grpc/client.go:103-106).
//
// One backend process serves one model (pkg/model/loader.go:117-141), so models that fit one GPU are replicated
// INSIDE this process: Load creates one engine per visible sm_100 device (B200_REPLICAS / MainGPU narrow that) and
// every request goes to the replica with the fewest requests in flight -- the policy of get_slot's LRU pick
// (grpc-server.cpp:609-628) one level up.  No data crosses GPUs.

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
	"runtime"
	"strconv"
	"sync"
	"unsafe"

	"github.com/mudler/LocalAI/pkg/grpc/base"
	pb "github.com/mudler/LocalAI/pkg/grpc/proto"
)

type replica struct {
	eng      *C.b2_engine
	inflight int
}

type LLM struct {
	base.Base
	mu       sync.Mutex
	replicas []*replica
	weights  uint64 // device bytes of one replica's weights (Status)
}

// b2_last_error() is thread-local in C, and a goroutine may move to another OS thread between a failing call and the
// read of its message: every call whose failure is reported pins the goroutine first (callPinned).
func callPinned(f func() bool) error {
	runtime.LockOSThread()
	defer runtime.UnlockOSThread()
	if f() {
		return nil
	}
	return fmt.Errorf("b200llm: %s", C.GoString(C.b2_last_error()))
}

// Load <- ModelOptions (backend.proto:126-203); replaces grpc-server.cpp:2312-2332.
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
	if opts.Embeddings {
		o.embeddings = 1
	}
	if opts.VocabOnly {
		o.vocab_only = 1
	}
	o.rope_freq_base = C.float(opts.RopeFreqBase)
	o.rope_freq_scale = C.float(opts.RopeFreqScale)
	// RopeScaling as the reference reads it (grpc-server.cpp:2279-2281): "none", "yarn", anything else (the empty
	// default included) = linear with the scale the GGUF was trained with
	switch opts.RopeScaling {
	case "none":
		o.rope_scaling = C.B2_ROPE_NONE
	case "yarn":
		o.rope_scaling = C.B2_ROPE_YARN
	default:
		o.rope_scaling = C.B2_ROPE_LINEAR
	}
	o.yarn_ext_factor = C.float(opts.YarnExtFactor)
	o.yarn_attn_factor = C.float(opts.YarnAttnFactor)
	o.yarn_beta_fast = C.float(opts.YarnBetaFast)
	o.yarn_beta_slow = C.float(opts.YarnBetaSlow)

	// which devices: MainGPU set => that one; else every sm_100 device, capped by B200_REPLICAS
	devices := []int{}
	if v, err := strconv.Atoi(opts.MainGPU); err == nil && opts.MainGPU != "" {
		devices = append(devices, v)
	} else {
		n := int(C.b2_device_count())
		if v, err := strconv.Atoi(os.Getenv("B200_REPLICAS")); err == nil && v > 0 && v < n {
			n = v
		}
		if n < 1 || opts.VocabOnly {
			n = 1 // vocab_only needs no device; without a device b2_create reports B2_ENODEV below
		}
		for d := 0; d < n; d++ {
			devices = append(devices, d)
		}
	}
	for _, d := range devices {
		o.device = C.int32_t(d)
		var eng *C.b2_engine
		if err := callPinned(func() bool { eng = C.b2_create(&o); return eng != nil }); err != nil {
			l.free()
			return err
		}
		l.replicas = append(l.replicas, &replica{eng: eng})
	}
	var info C.b2_model_info_t
	if C.b2_model_info(l.replicas[0].eng, &info) == 0 {
		l.weights = uint64(info.weight_bytes_device)
	}
	return nil
}

func (l *LLM) free() {
	for _, r := range l.replicas {
		C.b2_destroy(r.eng)
	}
	l.replicas = nil
}

// pick the replica with the fewest requests in flight; done() gives the slot back
func (l *LLM) pick() (*replica, func(), error) {
	l.mu.Lock()
	defer l.mu.Unlock()
	if len(l.replicas) == 0 {
		return nil, nil, fmt.Errorf("b200llm: no model loaded")
	}
	best := l.replicas[0]
	for _, r := range l.replicas[1:] {
		if r.inflight < best.inflight {
			best = r
		}
	}
	best.inflight++
	return best, func() { l.mu.Lock(); best.inflight--; l.mu.Unlock() }, nil
}

// request <- parse_options (grpc-server.cpp:2068-2137): PredictOptions -> b2_request.  The returned release() frees
// the C copies, releases the request id and gives the replica slot back; call it on every path.
func (l *LLM) request(opts *pb.PredictOptions, stream bool, embedding bool) (*replica, C.int64_t, func(), error) {
	rep, done, err := l.pick()
	if err != nil {
		return nil, 0, nil, err
	}
	var r C.b2_request
	r.struct_size = C.uint32_t(unsafe.Sizeof(r))
	text := opts.Prompt
	if embedding {
		text = opts.Embeddings
	}
	frees := []unsafe.Pointer{}
	cstr := func(s string) *C.char {
		p := C.CString(s)
		frees = append(frees, unsafe.Pointer(p))
		return p
	}
	r.prompt = cstr(text)
	r.prompt_len = C.size_t(len(text))
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
	r.min_p = 0.05 // PredictOptions has no MinP: the reference runs with llama.cpp's default (grpc-server.cpp:639)
	r.repeat_penalty, r.repeat_last_n = C.float(opts.Penalty), C.int32_t(opts.Repeat)
	r.frequency_penalty, r.presence_penalty = C.float(opts.FrequencyPenalty), C.float(opts.PresencePenalty)
	// LocalAI's default config sends mirostat 2, tau 5, eta 0.1 (core/config/backend_config.go:286-289)
	r.mirostat, r.mirostat_tau, r.mirostat_eta = C.int32_t(opts.Mirostat), C.float(opts.MirostatTAU), C.float(opts.MirostatETA)
	r.typical_p, r.tfs_z = C.float(opts.TypicalP), C.float(opts.TailFreeSamplingZ)
	r.seed = C.int32_t(opts.Seed)
	if opts.IgnoreEOS {
		r.ignore_eos = 1
	}
	if opts.PenalizeNL { // grpc-server.cpp:2111
		r.penalize_nl = 1
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
	if opts.Grammar != "" && !embedding { // grpc-server.cpp:2114; every tools / function-call request carries one
		r.grammar = cstr(opts.Grammar)
	}
	if n := len(opts.StopPrompts); n > 0 {
		arr := (**C.char)(C.malloc(C.size_t(n) * C.size_t(unsafe.Sizeof(uintptr(0)))))
		frees = append(frees, unsafe.Pointer(arr))
		s := unsafe.Slice(arr, n)
		for i, w := range opts.StopPrompts {
			s[i] = cstr(w)
		}
		r.stop, r.n_stop = arr, C.int32_t(n)
	}
	var id C.int64_t
	err = callPinned(func() bool { id = C.b2_submit(rep.eng, &r); return id >= 0 })
	for _, p := range frees { // b2_submit deep-copies the request
		C.free(p)
	}
	if err != nil {
		done()
		return nil, 0, nil, err
	}
	release := func() {
		C.b2_release(rep.eng, id)
		done()
	}
	return rep, id, release, nil
}

// drain polls events until the final one; f receives every text piece.
func drain(rep *replica, id C.int64_t, f func(string)) error {
	for {
		var ev C.b2_event
		ev.struct_size = C.uint32_t(unsafe.Sizeof(ev))
		if err := callPinned(func() bool { return C.b2_poll(rep.eng, id, &ev, -1) == 0 }); err != nil {
			return err
		}
		if ev.flags&C.B2_EV_ERROR != 0 {
			return fmt.Errorf("b200llm: %s", C.GoStringN(ev.text, C.int(ev.text_len)))
		}
		if ev.text_len > 0 && f != nil {
			f(C.GoStringN(ev.text, C.int(ev.text_len)))
		}
		if ev.flags&C.B2_EV_FINAL != 0 {
			return nil
		}
	}
}

// Predict <- grpc-server.cpp:2380-2407
func (l *LLM) Predict(opts *pb.PredictOptions) (string, error) {
	rep, id, release, err := l.request(opts, false, false)
	if err != nil {
		return "", err
	}
	defer release()
	out := ""
	err = drain(rep, id, func(s string) { out += s })
	return out, err
}

// PredictStream <- grpc-server.cpp:2333-2377; the callee closes the channel (pkg/grpc/base/base.go:43-46).
func (l *LLM) PredictStream(opts *pb.PredictOptions, results chan string) error {
	rep, id, release, err := l.request(opts, true, false)
	if err != nil {
		close(results)
		return err
	}
	go func() {
		defer release()
		defer close(results)
		_ = drain(rep, id, func(s string) { results <- s })
	}()
	return nil
}

// Embeddings <- grpc-server.cpp:2410-2432
func (l *LLM) Embeddings(opts *pb.PredictOptions) ([]float32, error) {
	rep, id, release, err := l.request(opts, false, true)
	if err != nil {
		return nil, err
	}
	defer release() // every return path below gives the request back
	if err := drain(rep, id, nil); err != nil {
		return nil, err // includes B2_EV_ERROR events
	}
	var info C.b2_model_info_t
	if err := callPinned(func() bool { return C.b2_model_info(rep.eng, &info) == 0 }); err != nil {
		return nil, err
	}
	out := make([]float32, int(info.n_embd))
	if len(out) == 0 {
		return out, nil
	}
	var n C.int
	if err := callPinned(func() bool {
		n = C.b2_embed_result(rep.eng, id, (*C.float)(unsafe.Pointer(&out[0])), C.int(len(out)))
		return n >= 0
	}); err != nil {
		return nil, err
	}
	return out[:int(n)], nil
}

// TokenizeString: the reference C++ backend does not implement it (SURVEY.md 3.5); this one does.
func (l *LLM) TokenizeString(opts *pb.PredictOptions) (pb.TokenizationResponse, error) {
	l.mu.Lock()
	if len(l.replicas) == 0 {
		l.mu.Unlock()
		return pb.TokenizationResponse{}, fmt.Errorf("b200llm: no model loaded")
	}
	eng := l.replicas[0].eng // the tokenizer is host code, identical in every replica
	l.mu.Unlock()
	text := C.CString(opts.Prompt)
	defer C.free(unsafe.Pointer(text))
	cap := len(opts.Prompt) + 16
	buf := make([]int32, cap)
	var n C.int
	if err := callPinned(func() bool {
		n = C.b2_tokenize(eng, text, C.size_t(len(opts.Prompt)), 1, 1, (*C.int32_t)(unsafe.Pointer(&buf[0])), C.int(cap))
		return n >= 0
	}); err != nil {
		return pb.TokenizationResponse{}, err
	}
	return pb.TokenizationResponse{Length: int32(n), Tokens: buf[:n]}, nil
}

// Busy / Status <- pkg/grpc/server.go:173-180, base.Base.Status (pkg/grpc/base/base.go): state + memory of the backend.
func (l *LLM) Busy() bool {
	l.mu.Lock()
	defer l.mu.Unlock()
	for _, r := range l.replicas {
		if r.inflight > 0 {
			return true
		}
	}
	return false
}

func (l *LLM) Status() (pb.StatusResponse, error) {
	state := pb.StatusResponse_READY
	if len(l.replicas) == 0 {
		state = pb.StatusResponse_UNINITIALIZED
	} else if l.Busy() {
		state = pb.StatusResponse_BUSY
	}
	total := l.weights * uint64(len(l.replicas))
	return pb.StatusResponse{State: state, Memory: &pb.MemoryUsageData{Total: total, Breakdown: map[string]uint64{"gpu_weights": total}}}, nil
}

// Metrics sums b2_metrics over the replicas: what GetMetrics (grpc-server.cpp:2434-2457) reports.  pkg/grpc's Go
// server has no GetMetrics hook in the LLM interface (interface.go:7-27); INTEGRATION.md shows the three lines a
// maintainer adds to pkg/grpc/server.go to serve it from here.
func (l *LLM) Metrics() *pb.MetricsResponse {
	out := &pb.MetricsResponse{}
	l.mu.Lock()
	reps := append([]*replica(nil), l.replicas...)
	l.mu.Unlock()
	for _, r := range reps {
		var m C.b2_metrics_t
		if C.b2_metrics(r.eng, &m) != 0 {
			continue
		}
		out.TokensPerSecond += float32(m.tokens_per_second)
		out.TokensGenerated += int32(m.tokens_generated)
		out.PromptTokensProcessed += int32(m.prompt_tokens_processed)
	}
	return out
}
