/*
 * b200llm.h -- C ABI of libb200llm.so: a B200-native (sm_100a) replacement for the engine inside
 * LocalAI's `llama-cpp` backend.  Everything the reference's gRPC server asks of llama.cpp on the
 * LoadModel / Predict / PredictStream / TokenizeString / Embedding path goes through the entry points
 * below; a Go (cgo), Python (ctypes) or C++ host binds exactly these.  Plain pointers and sizes only.
 *
 * Citations are relative to /root/reference (siddimore/LocalAI @ c8bfb72).
 *
 * Conventions: functions return 0 (or a count) on success and a negative B2_E* code on failure; the
 * message is available from b2_last_error() (thread-local).  All `char*` are UTF-8 / raw bytes and are
 * NOT required to be NUL-terminated when a length is given.  Pointers are HOST pointers unless the
 * name says `dev`.  There is no CPU fallback: every compute entry point fails with B2_ENODEV when no
 * sm_100 device is present.
 */
#ifndef B200LLM_H
#define B200LLM_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B2_ABI_VERSION 1

enum {
    B2_OK = 0,
    B2_EINVAL = -1,   /* bad argument */
    B2_ENODEV = -2,   /* no usable CUDA device */
    B2_EIO = -3,      /* file / GGUF parse error */
    B2_ENOMEM = -4,   /* device or host allocation failed */
    B2_ECUDA = -5,    /* CUDA runtime error */
    B2_ENOSPC = -6,   /* output buffer too small (count needed is returned negated where documented) */
    B2_ESTATE = -7,   /* request id unknown / engine shutting down */
    B2_ETIMEOUT = -8, /* b2_poll timed out (no event yet) */
    B2_EUNSUPPORTED = -9
};

/* ggml tensor type ids used by GGUF (llama.cpp ggml.h `enum ggml_type`, un-vendored: Makefile:11) */
enum { B2_T_F32 = 0, B2_T_F16 = 1, B2_T_Q4_0 = 2, B2_T_Q4_1 = 3, B2_T_Q5_0 = 6, B2_T_Q5_1 = 7, B2_T_Q8_0 = 8, B2_T_Q2_K = 10,
       B2_T_Q3_K = 11, B2_T_IQ4_NL = 20, B2_T_IQ4_XS = 23, B2_T_Q4_K = 12,
       B2_T_Q5_K = 13, B2_T_Q6_K = 14 };
/* Q4_0 / Q4_1 / Q5_0 / Q5_1 (legacy 32-element blocks) and Q2_K / Q3_K / IQ4_NL / IQ4_XS run on the tensor-core path only: dequantised once into resident
 * fp16 operand tiles; the integer decode kernels cover Q8_0 and the K-quants. */

typedef struct b2_engine b2_engine;

int b2_abi_version(void);
const char *b2_last_error(void);
/* number of sm_100 devices visible; 0 if none */
int b2_device_count(void);

/* ------------------------------------------------------------------------------------------------
 * LoadModel.  Replaces BackendServiceImpl::LoadModel -> params_parse -> llama_server_context::load_model
 * -> initialize (backend/cpp/llama/grpc-server.cpp:2312-2332, 2205-2300, 444-486, 508-557) and, for a
 * Go backend, LLM.Load(*pb.ModelOptions) (pkg/grpc/interface.go:14).  Fields mirror pb.ModelOptions
 * (backend/backend.proto:162-235).
 */
typedef struct {
    uint32_t struct_size;   /* sizeof(b2_load_opts), for ABI evolution */
    const char *model_path; /* ModelOptions.ModelFile (absolute path of a GGUF file) */
    int32_t n_ctx;          /* ContextSize; total KV positions, split evenly among slots (grpc-server.cpp:512). 0 = 1024 */
    int32_t n_batch;        /* NBatch: prompt chunk size per engine step (grpc-server.cpp:1864). 0 = 512 */
    int32_t n_parallel;     /* LLAMACPP_PARALLEL: number of slots (grpc-server.cpp:2225-2231). 0 = 1 */
    int32_t device;         /* CUDA ordinal (MainGPU) */
    int32_t embeddings;     /* ModelOptions.Embeddings (grpc-server.cpp:2275) */
    float rope_freq_base;   /* 0 = from GGUF (grpc-server.cpp:2279-2299) */
    float rope_freq_scale;  /* 0 = 1.0 */
    int32_t kv_page_size;   /* KV page length in tokens; 0 = 32 */
    int32_t vocab_only;     /* ModelOptions.VocabOnly: load metadata + tokenizer only; needs no GPU; compute calls fail */
    int32_t tp_size;        /* tensor-parallel degree (ModelOptions.TensorSplit names that many GPUs; backend.proto:175-176,
                               grpc-server.cpp:2240-2262): 0 / 1 = one GPU; 2, 4, 8 = the model is sharded over the devices
                               device .. device + tp_size - 1 of this process (70B-class models) */
    int32_t rope_scaling;   /* ModelOptions.RopeScaling (grpc-server.cpp:2279-2281): 0 = as the GGUF says, B2_ROPE_NONE / _LINEAR / _YARN
                               override it (the reference maps "none" / "yarn" / anything else to none / YaRN / linear) */
    float yarn_ext_factor;  /* YarnExtFactor, YarnAttnFactor, YarnBetaFast, YarnBetaSlow (grpc-server.cpp:2282-2293): 0 = llama.cpp's */
    float yarn_attn_factor; /* defaults (ext_factor 1 for YaRN else 0, attn_factor 1, beta_fast 32, beta_slow 1) */
    float yarn_beta_fast;
    float yarn_beta_slow;
    int32_t reserved[1];
} b2_load_opts;
enum { B2_ROPE_FROM_GGUF = 0, B2_ROPE_NONE = 1, B2_ROPE_LINEAR = 2, B2_ROPE_YARN = 3 };

b2_engine *b2_create(const b2_load_opts *opts);
void b2_destroy(b2_engine *e);

typedef struct {
    int32_t n_layer, n_embd, n_head, n_head_kv, head_dim, n_ff, n_vocab, n_ctx_train;
    int32_t n_ctx, n_ctx_slot, n_parallel, n_batch;
    int32_t bos_id, eos_id, eot_id, add_bos;
    float rms_eps, rope_freq_base, rope_freq_scale;
    int64_t weight_bytes_device; /* bytes of weights resident in HBM */
    int64_t decode_stream_bytes; /* weight bytes one decode step must read */
    int64_t kv_bytes_per_pos;    /* KV bytes appended per position (all layers) */
    char arch[32];
    char tokenizer_model[32];
    char name[128];
} b2_model_info_t;
int b2_model_info(b2_engine *e, b2_model_info_t *out);

/* ------------------------------------------------------------------------------------------------
 * TokenizeString.  Replaces llama_server_context::tokenize -> ::llama_tokenize (grpc-server.cpp:559-607)
 * and LLM.TokenizeString (pkg/grpc/interface.go:22; pkg/grpc/server.go:152-171).  Returns the number of
 * tokens, or -(needed) if cap is too small.
 */
int b2_tokenize(b2_engine *e, const char *text, size_t n, int add_bos, int parse_special, int32_t *out, int cap);
/* llama_token_to_piece (grpc-server.cpp:116,124,1012).  Returns bytes written or -(needed). */
int b2_token_to_piece(b2_engine *e, int32_t token, char *out, int cap);

/* ------------------------------------------------------------------------------------------------
 * Tensor parallelism (b2_load_opts.tp_size > 1).  Replaces llama.cpp's tensor_split / main_gpu, the reference's only
 * multi-GPU knobs (backend/backend.proto:175-176 -> backend/cpp/llama/grpc-server.cpp:2240-2262).  Every entry point of
 * this header works on such an engine unchanged; the two below select and report the exchange between the shards.
 *   mode 0 (B2_TP_NCCL)   per-layer kernels + ncclAllReduce after attn_output and ffn_down (BASELINE.json north_star)
 *   mode 1 (B2_TP_FUSED)  one persistent decode kernel per GPU; partial rows are stored into the peers' memory over
 *                         NVLink and summed by the consuming phase (the default when the model's shapes allow it)
 */
enum { B2_TP_NCCL = 0, B2_TP_FUSED = 1 };
typedef struct {
    int32_t tp_size;               /* 1: not tensor-parallel */
    int32_t mode;                  /* B2_TP_NCCL / B2_TP_FUSED */
    int32_t n_allreduce_per_step;  /* 2 x n_layer */
    int32_t reserved;
    double allreduce_ms_per_step;  /* B2_TP_NCCL: time inside the collectives per step of the last b2_bench_decode (CUDA events, shard 0) */
    int64_t shard_stream_bytes;    /* weight bytes ONE shard streams per decode step */
} b2_tp_info_t;
int b2_tp_set_mode(b2_engine *e, int mode);
int b2_tp_info(b2_engine *e, b2_tp_info_t *out);

/* ------------------------------------------------------------------------------------------------
 * Predict / PredictStream / Embedding.  Replaces parse_options + request_completion + the result queue
 * (grpc-server.cpp:2068-2137, 1323-1359, 2333-2432; utils.hpp:192-409) and LLM.Predict / PredictStream /
 * Embeddings (pkg/grpc/interface.go:12-15).  Fields mirror pb.PredictOptions (backend.proto:106-153).
 * Streaming is pull-based: b2_submit returns a request id; the caller loops on b2_poll until an event
 * has B2_EV_FINAL set.  Any thread may submit/poll; the engine thread owns the device.
 */
typedef struct {
    uint32_t struct_size;
    const char *prompt;        /* PredictOptions.Prompt (or .Embeddings in embedding mode) */
    size_t prompt_len;
    const int32_t *prompt_ids; /* optional pre-tokenised prompt (EmbeddingTokens); used when n_prompt_ids > 0 */
    int32_t n_prompt_ids;
    int32_t n_predict;         /* Tokens; 0 => -1 (until EOS / context full) as grpc-server.cpp:2098 */
    float temperature;         /* <= 0 => greedy */
    int32_t top_k;
    float top_p;
    float min_p;
    float repeat_penalty;      /* Penalty */
    int32_t repeat_last_n;     /* Repeat */
    float frequency_penalty, presence_penalty;
    int32_t seed;              /* Seed (-1 = random) */
    int32_t ignore_eos;        /* IgnoreEOS (grpc-server.cpp:697-699) */
    int32_t cache_prompt;      /* PromptCacheAll (grpc-server.cpp:2097) */
    int32_t n_keep;            /* NKeep */
    int32_t embedding_mode;    /* Embedding RPC: n_predict = 0, return the embedding vector */
    const char *const *stop;   /* StopPrompts */
    const size_t *stop_len;    /* optional lengths; NULL => NUL-terminated */
    int32_t n_stop;
    int32_t stream;            /* 1: an event per emitted piece; 0: one final event with the whole text */
    int32_t mirostat;          /* Mirostat: 0 off, 1 = v1, 2 = v2 (what LocalAI's default config sends,
                                  core/config/backend_config.go:286-289) */
    float mirostat_tau;        /* MirostatTAU (<= 0 => 5.0) */
    float mirostat_eta;        /* MirostatETA (<= 0 => 0.1) */
    float typical_p;           /* TypicalP: 0 or >= 1 = off (with typical-p on, at most the 256 most probable tokens are considered) */
    float tfs_z;               /* TailFreeSamplingZ: 0 or >= 1 = off */
    int32_t penalize_nl;       /* PenalizeNL (grpc-server.cpp:2111): 1 = the newline token takes the repeat / frequency /
                                  presence penalties like any other; 0 (the proto default) = it is exempt */
    int32_t reserved[2];
    const char *grammar;       /* Grammar: GBNF text, NUL-terminated (NULL or "" = none).  The sampled token is checked
                                  against it; when it does not fit, the row is re-sampled under a vocabulary mask
                                  (grpc-server.cpp:655,2114).  A text that does not parse fails b2_submit (B2_EINVAL). */
} b2_request;

enum {
    B2_EV_FINAL = 1,     /* last event of the request */
    B2_EV_STOP_EOS = 2,  /* stopped_eos */
    B2_EV_STOP_WORD = 4, /* stopped_word */
    B2_EV_STOP_LIMIT = 8,/* stopped_limit (n_predict reached or context full) */
    B2_EV_ERROR = 16,
    B2_EV_EMBEDDING = 32 /* result available from b2_embed_result */
};

typedef struct {
    uint32_t struct_size;
    uint32_t flags;
    const char *text;      /* piece bytes; valid until the next b2_poll for the same id */
    size_t text_len;
    int32_t n_prompt_tokens;  /* tokens_evaluated (Reply.prompt_tokens, grpc-server.cpp:2355-2358) */
    int32_t n_decoded;        /* tokens_predicted (Reply.tokens) */
    int32_t slot_id;
    int32_t n_embd;           /* >0 when B2_EV_EMBEDDING */
    double t_prompt_ms;       /* timings as get_formated_timings (grpc-server.cpp:305-318) */
    double t_gen_ms;
    int32_t reserved[4];
} b2_event;

int64_t b2_submit(b2_engine *e, const b2_request *req);
int b2_poll(b2_engine *e, int64_t id, b2_event *ev, int timeout_ms);
int b2_cancel(b2_engine *e, int64_t id);
/* copy the embedding of a finished embedding-mode request (send_embedding, grpc-server.cpp:1292-1321) */
int b2_embed_result(b2_engine *e, int64_t id, float *out, int cap);
/* forget a finished request (frees its result buffers) */
int b2_release(b2_engine *e, int64_t id);

/* GetMetrics (grpc-server.cpp:2434-2457; backend.proto:36-42) */
typedef struct {
    int32_t slot_id;
    float tokens_per_second;
    int32_t tokens_generated;
    int32_t prompt_tokens_processed;
    int32_t n_slots_busy;
    int64_t n_decode_steps;
    int64_t n_tokens_decoded_total;
    int64_t n_prompt_tokens_total;
} b2_metrics_t;
int b2_metrics(b2_engine *e, b2_metrics_t *out);

/* ------------------------------------------------------------------------------------------------
 * Synchronous low-level forward: the llama_decode call itself (grpc-server.cpp:1910) on one slot,
 * bypassing the scheduler.  Used by the parity tests, bench.py and embedders that drive their own loop.
 * Must not be mixed with in-flight b2_submit requests on the same slot.
 *
 * b2_eval: evaluate `n` tokens at positions [n_past, n_past+n) of `slot`.
 *   logits: NULL, or host [n_vocab] (last token) / [n][n_vocab] if logits_all.
 *   embd:   NULL or host [n_embd]: final-norm hidden state of the last token (llama_get_embeddings).
 * b2_eval_step: ONE decode step over several slots -- token i at position pos[i] of slot slots[i] (1 <= n <= n_batch) --
 *   the batch shape the scheduler issues for concurrent streams (llama_decode on a multi-sequence batch,
 *   grpc-server.cpp:1885-1915); logits: NULL or host [n][n_vocab].
 * b2_decode_greedy: prompt -> n_gen greedy tokens, sampling on device; only token ids cross PCIe.
 */
int b2_eval(b2_engine *e, int slot, const int32_t *tokens, int n, int n_past, float *logits, int logits_all, float *embd);
int b2_eval_step(b2_engine *e, const int32_t *tokens, const int32_t *pos, const int32_t *slots, int n, float *logits);
int b2_decode_greedy(b2_engine *e, int slot, const int32_t *prompt, int n_prompt, int n_gen, int32_t *out,
                     double *prefill_ms, double *decode_ms);
/* hidden state of the last evaluated token after each layer: host [n_layer+1][n_embd] (parity bisecting) */
int b2_debug_hidden(b2_engine *e, float *out, int cap_floats);

/* The sampler alone (gpt_sampler_sample, grpc-server.cpp:1946) on host logits: `n_draws` successive draws from the same
 * row, the random state and the mirostat mu carried from draw to draw; penalties are applied to a fresh copy of the
 * row for every draw.  out_mu (optional): mu after each draw.  Test hook. */
typedef struct {
    uint32_t struct_size;
    float temperature;
    int32_t top_k;
    float top_p, min_p;
    float repeat_penalty, frequency_penalty, presence_penalty;
    int32_t mirostat;
    float mirostat_tau, mirostat_eta, mirostat_mu;
    int32_t ban_token, nl_token;
    int64_t seed;
    const int32_t *prev; /* penalty window (the last n_prev tokens) */
    int32_t n_prev;
    float tfs_z, typical_p; /* 0 or >= 1 = off */
} b2_sample_params;
/* ------------------------------------------------------------------------------------------------
 * BERT embeddings: the `bert-embeddings` backend.  Replaces bert.New / Embeddings / TokenEmbeddings of go-bert.cpp as
 * bound by backend/go/llm/bert/bert.go:18,30,33 (Load, Embeddings with PredictOptions.Embeddings or .EmbeddingTokens),
 * for a BERT encoder stored as GGUF (general.architecture = "bert", WordPiece vocabulary).  One text -> one vector of
 * n_embd floats: mean over the tokens' last hidden states, L2-normalised (bert.cpp bert_eval).  Calls on one handle
 * are serialised (the reference backend is base.SingleThread).  vocab_only != 0 loads the tokenizer alone (no device).
 */
typedef struct b2_bert b2_bert;
typedef struct {
    int32_t n_layer, n_embd, n_head, n_ff, n_ctx, n_vocab;
    int64_t weight_bytes_device;
} b2_bert_info_t;
b2_bert *b2_bert_load(const char *model_path, int device, int vocab_only); /* NULL on failure: b2_last_error() */
void b2_bert_free(b2_bert *b);
int b2_bert_info(b2_bert *b, b2_bert_info_t *out);
/* [CLS] pieces... [SEP], cut to the model's context; returns the count, or -(count needed) if cap is too small */
int b2_bert_tokenize(b2_bert *b, const char *text, size_t text_len, int32_t *out, int cap);
/* Embeddings(text) / TokenEmbeddings(ids): writes n_embd floats, returns n_embd */
int b2_bert_embed_text(b2_bert *b, const char *text, size_t text_len, float *out, int cap);
int b2_bert_embed_tokens(b2_bert *b, const int32_t *ids, int n_ids, float *out, int cap);
/* several token sequences in one forward (no counterpart in the reference, whose API carries one text per call):
 * ids back to back, offsets[n_seq + 1]; out [n_seq][n_embd]; at most 32768 tokens per call */
int b2_bert_embed_batch(b2_bert *b, const int32_t *ids, const int32_t *offsets, int n_seq, float *out);
/* test hook: last hidden states [n_ids][n_embd] of one sequence */
int b2_bert_hidden(b2_bert *b, const int32_t *ids, int n_ids, float *out);

/* Grammar matcher alone (host only, no device needed).  b2_grammar_match parses `gbnf` and feeds `text` one code point
 * at a time: *n_accepted = bytes taken before the first code point that does not fit, *can_stop = 1 if the accepted
 * text is a complete sentence of the grammar.  b2_grammar_token_mask: after `accepted`, bit t of mask_bits
 * ((n_vocab + 31) / 32 words) = piece t may come next (end-of-generation pieces only where the grammar may stop).
 * A grammar that does not parse: B2_EINVAL, reason in b2_last_error.  Replaces llama_grammar_parse / _apply / _accept
 * behind gpt_sampler_init / _sample / _accept (grpc-server.cpp:890,1946,1948). */
int b2_grammar_match(const char *gbnf, const char *text, size_t text_len, size_t *n_accepted, int *can_stop);
int b2_grammar_token_mask(const char *gbnf, const char *accepted, size_t accepted_len, const char *const *pieces, const size_t *piece_len,
                          const uint8_t *is_eog, int n_vocab, uint32_t *mask_bits);
int b2_sample(int device, const float *logits, int n_vocab, const b2_sample_params *p, int n_draws, int32_t *out_tokens, float *out_mu);

/* Device-resident benchmark legs (inputs already in HBM, timed with CUDA events on the engine's stream).
 * b2_bench_decode: `n_steps` single-token decode steps for `batch` concurrent sequences whose KV already
 * holds `ctx_len` positions; returns average ms per step in *ms_per_step and kernel launches per step.
 * b2_bench_prefill: one prompt of n_tokens through the prefill path; average ms over n_iters. */
int b2_bench_decode(b2_engine *e, int batch, int ctx_len, int n_steps, int n_warmup, double *ms_per_step,
                    int64_t *launches_per_step);
int b2_bench_prefill(b2_engine *e, int n_tokens, int n_iters, int n_warmup, double *ms_per_iter,
                     int64_t *launches_per_iter);

/* b2_bench_gemv: times the streaming dequant-GEMV kernel alone (CUDA events on the engine's stream) over one
 * weight matrix of EVERY layer in turn (so each launch reads weights that are not in L2), n_rounds times.
 * which: 0 wq, 1 wk, 2 wv, 3 wo, 4 ffn_gate, 5 ffn_up, 6 ffn_down, 7 output (lm_head; single matrix).
 * n_tok 1..4: the decode kernels; 5..64: the batched-decode GEMM that dequantises inside the kernel, for which
 * which = 8 (wq|wk|wv) and 9 (ffn_gate|ffn_up) time the multi-matrix launches the engine makes.
 * Returns average ms per launch, algorithmic bytes per launch (rows x row bytes) and the ggml type id. */
int b2_bench_gemv(b2_engine *e, int which, int n_tok, int n_rounds, double *ms_per_launch, int64_t *bytes_per_launch,
                  int *ggml_type);

/* ------------------------------------------------------------------------------------------------
 * Kernel-level test hooks (host buffers in, host buffers out; each call uploads, runs the same kernels
 * the engine uses, downloads).  `blocks` are in GGUF on-disk block layout.
 *
 * b2_dequantize: on-device dequantisation of `n_elems` weights (dequantize_row_* semantics; bit-exact).
 * b2_quantize_act: the activation quantisation llama.cpp applies before a mat-mul with weights of
 *   `wtype` (Q8_0 -> Q8_0 blocks, K-quants -> Q8_K blocks, F16 -> f16); out receives the blocks in
 *   ggml's AoS layout (block_q8_0 34 B / block_q8_K 292 B) so they can be compared byte for byte.
 * b2_mul_mat: out[t][r] = dot(W[r,:], quantise(x[t,:])), W = rows x cols of `wtype`, x = n_tok x cols f32.
 *   path: 0 = auto, 1 = streaming dequant-GEMV (one op per kernel), 2 = tensor-core GEMM (prefill path),
 *         3 = fused decode kernel (quantises in its prologue), 5 = fused decode kernel on pre-quantised rows,
 *         4 = persistent decode mega-kernel with a single mat-vec phase (n_tok = 1, even rows),
 *         6 = batched-decode GEMM that dequantises inside the kernel (Q4_K / Q5_K / Q6_K / Q8_0, n_tok <= 64).
 */
int b2_dequantize(int type, const void *blocks, int64_t n_elems, float *out);
int b2_quantize_act(int wtype, const float *x, int64_t n_elems, void *out, size_t out_cap);
int b2_mul_mat(int wtype, const void *blocks, int64_t rows, int64_t cols, const float *x, int n_tok, float *out, int path);

/* ------------------------------------------------------------------------------------------------
 * b2_mul_mat_noalloc -- has no counterpart in the reference.
 *
 * The same quantised mat-mul as b2_mul_mat (out[t][r] = dot(W[r,:], quantise(x[t,:]))), for callers that must not
 * allocate device memory: frameworks that have already handed all of HBM to their own pool, latency-critical
 * loops that cannot afford cudaMalloc's locks and implicit synchronisation, embedders under a "no allocation
 * after start-up" rule.  Everything is passed in as DEVICE pointers owned by the caller (weights in GGUF block
 * layout, activations f32 [n_tok][cols], result f32 [n_tok][rows]); the scratch for the quantised activations is a
 * fixed 128 KiB array inside the library, which every GPU that loads the library holds until exit.
 *   - limits: wtype in {B2_T_Q4_K, B2_T_Q5_K, B2_T_F16}; cols % 256 == 0 and cols <= 16384; 1 <= n_tok <= 4.
 *     Anything larger is rejected with B2_EINVAL.
 *   - w_dev must be 16-byte aligned, x_dev 16-byte aligned; all pointers must belong to the CURRENT device.
 *   - runs asynchronously on `cuda_stream` (a cudaStream_t, NULL = default stream); performs no allocation.
 *   - the first call on each device looks up the scratch array's device address; make that first call before
 *     starting any stream capture.
 *   - all calls on a device share the one scratch array (it keeps the previous call's contents, and needs no
 *     clearing: each call rewrites what it reads), so calls that overlap on the GPU race -- serialise them. */
int b2_mul_mat_noalloc(int wtype, const void *w_dev, int64_t rows, int64_t cols, const float *x_dev, int n_tok,
                       float *out_dev, void *cuda_stream);

#ifdef __cplusplus
}
#endif
#endif /* B200LLM_H */
