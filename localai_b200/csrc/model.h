// model.h -- llama-architecture model resident in HBM + the forward pass (the llama_decode replacement).
#pragma once
#include <cuda_fp16.h>

#include <memory>
#include <string>
#include <unordered_map>
#include <vector>

#include "common.h"
#include "gguf.h"
#include "kernels.h"

namespace b2 {

struct SampleSpec;

struct HParams {
    int n_layer = 0, n_embd = 0, n_head = 0, n_head_kv = 0, head_dim = 0, n_ff = 0, n_vocab = 0, n_ctx_train = 0;
    float rms_eps = 1e-5f, rope_base = 10000.f, rope_scale = 1.f;
    bool rope_neox = false;      // the architecture rotates pairs (i, i + head_dim / 2) (qwen2, phi3) instead of (2i, 2i + 1)
    bool rope_neox_kernel = false; // ... and the kernels do it (A/B switch); normally the rows of wq / wk are interleaved at load instead
    float rope_ext_factor = 0.f; // 1 for YaRN (llama.rope.scaling.type)
    float rope_attn_factor = 1.f, rope_beta_fast = 32.f, rope_beta_slow = 1.f; // llama.cpp's defaults, b2_load_opts overrides
    int rope_n_ctx_orig = 0;     // llama.rope.scaling.original_context_length
    std::string arch, name;
};

struct LayerW {
    DTensor attn_norm, wq, wk, wv, wo, ffn_norm, gate, up, down;
    DTensor bq, bk, bv; // attn_{q,k,v}.bias (qwen2); data == nullptr when the model has none
};

// A batch for one forward pass: n tokens, each with its own position and slot (sequence).
struct Batch {
    int n = 0;
    const int32_t *tokens = nullptr; // host; nullptr => token ids are already in the device token buffer
    const int32_t *pos = nullptr;    // host
    const int32_t *slot = nullptr;   // host
    // rows (token indices) whose logits are wanted; nullptr + n_out == -1 => all rows
    const int32_t *out_rows = nullptr; // host
    int n_out = 0;
};

struct TPGroup; // tp.cu: the shards of one tensor-parallel model, their NCCL communicators and worker threads

class Model {
  public:
    // tp_size > 1: this object is shard `tp_rank` of a tensor-parallel model on device o.device + tp_rank (SURVEY.md 8e
    // row 2): column-parallel wq|wk|wv|gate|up|output (rows tp_rank * R/tp .. of each), row-parallel wo|down (the
    // K-slice that multiplies this shard's heads / ffn rows), n_head_kv / tp KV heads, everything else replicated.
    // Shard 0, created with o.tp_size, creates and drives the others.
    Model(const GGUFFile &f, const b2_load_opts &o, int tp_rank = 0, int tp_size = 0);
    ~Model();
    Model(const Model &) = delete;

    // Runs the graph on `stream` (asynchronous).  Logits land in logits_dev() [n_out][n_vocab];
    // the final-norm hidden states of the output rows land in embd_dev() [n_out][n_embd].
    void forward(const Batch &b);
    // greedy sample all output rows of the last forward into next_tok_dev(); if `feed`, also write them to
    // the device token buffer and advance the device positions (device-side decode loop, no host round trip)
    void sample_greedy(int n_rows, bool feed);
    // decode step whose inputs (tokens/pos/slot for `n` sequences) already live on the device
    void forward_device_inputs(int n);
    void set_device_inputs(const int32_t *tokens, const int32_t *pos, const int32_t *slot, int n);

    void sync() { B2_CUDA(cudaStreamSynchronize(stream)); }

    const float *logits_dev() const { return logits_; }
    float *logits_dev_mut() { return logits_; } // the sampler applies penalties in place
    float *sample_mu_dev() { return sample_mu_; }
    const float *embd_dev() const { return xn_out_; }
    const int32_t *next_tok_dev() const { return next_tok_; }
    float *debug_hidden_dev() { return dbg_hidden_; }
    SampleSpec *sample_specs_dev(int n);
    std::vector<std::pair<int32_t, float>> &sample_specs_held() { return sample_specs_held_; } // what sample_rows knows the device copy to hold
    uint32_t *token_mask_dev(); // (n_vocab + 31) / 32 words, allocated on first use (grammar resample)
    void enable_debug_hidden();
    int bench_gemv(int which, int n_tok, int n_rounds, double *ms_per_launch, int64_t *bytes_per_launch, int *ggml_type);

    HParams hp;          // of the whole model (global head counts and widths, also in a tensor-parallel shard)
    int tp_size = 1, tp_rank = 0;
    int nh_l = 0, nkv_l = 0, ff_l = 0, dq_l = 0, kvd_l = 0, vocab_l = 0; // this shard's heads, KV heads, ffn rows, q width, KV width, LM-head rows
    int n_ctx = 0, n_ctx_slot = 0, n_parallel = 1, n_batch = 512, page_size = 32, max_pages = 0;
    int device = 0;
    int64_t weight_bytes = 0, decode_stream_bytes = 0, kv_bytes_per_pos = 0;
    cudaStream_t stream = nullptr;

    std::vector<LayerW> layers;
    DTensor tok_embd, output_norm, output;

    // tensor-parallel hooks (tp.cu)
    enum TPMode { TP_NCCL = 0, TP_FUSED = 1 };
    void tp_set_mode(int mode);      // TP_NCCL: per-layer kernels + ncclAllReduce; TP_FUSED: persistent kernels that exchange partial rows over peer memory
    double tp_last_allreduce_ms = 0; // TP_NCCL: time inside the all-reduces of the last forward, measured with events on shard 0
    int tp_n_allreduce = 0;          // all-reduces per decode step
    void tp_timing(bool on);
    double tp_allreduce_ms();
    void tp_sync();
    int tp_mode() const { return tp_mode_; }

  private:
    friend struct TPGroup;
    TPGroup *tp_ = nullptr;          // shard 0 only
    void tp_destroy();
    Model *tp_root_ = nullptr;       // shards > 0: shard 0
    int tp_mode_ = TP_FUSED;
    float *xp_ = nullptr;            // [n_batch][n_embd] partial output of the row-parallel mat-muls (input of the all-reduce)
    void tp_init(const GGUFFile &f, const b2_load_opts &o);
    void tp_build_fused();                               // phase lists + weight streams of the fused decode kernels, all shards
    bool tp_fused_ready_ = false;
    uint2 *tp_part_[2] = {nullptr, nullptr};             // [tp_size][n_embd] LL words: the shards' partial rows (attn_output / ffn_down alternate)
    cudaEvent_t tp_done_ = nullptr, tp_in_ = nullptr;
    void forward_tp(const Batch &b);
    void forward_device_inputs_tp(int n);
    void tp_step(int n, int t0, int out_row0);           // forward of tokens [t0, t0 + n) on THIS shard (called on its worker thread)
    void run_layer_tp(const LayerW &L, const KVLayout &kv, int n, int t0);
    void run_decode_mega_tp(int t0, int out_row0);
    void tp_allreduce(int t0, int n);                    // xp_ -> x_ (sum over the shards)
    // column / K slices of a tensor: rows [row0, row0 + nrows) x columns [col0, col0 + ncols) (col ranges on block boundaries)
    // pair_heads > 0: the kept rows are `pair_heads`-row heads whose rows are interleaved on the way (i, i + h/2 -> 2i, 2i + 1)
    DTensor upload(const GGUFFile &f, const std::string &name, int64_t want_rows, int64_t want_cols, bool required,
                   int64_t row0 = 0, int64_t nrows = -1, int64_t col0 = 0, int64_t ncols = -1, int pair_heads = 0);
    void quantize_for(const DTensor &W, const float *x, int64_t x_stride, int n, uint8_t **out, ActFmt *fmt);
    void matmul(const DTensor &W, const float *x, int64_t x_stride, int n, float *out, int64_t out_stride,
                const float *residual, int64_t res_stride);
    void run_graph(int n, const int32_t *out_rows_dev, int n_out);
    void run_layer_unfused(const LayerW &L, const KVLayout &kv, int n);
    void run_layer_fused(const LayerW &L, const KVLayout &kv, int n);
    void run_decode_graphed(int n);
    struct Run { int t0, n, pos0, slot; };
    std::vector<Run> runs_; // host view of the current batch (empty: inputs only on the device)
    void attention(const KVLayout &kv, int n);
    bool attention_to_tiles(const KVLayout &kv, int n, void *x_tiles);
    void run_layer_gemm(const LayerW &L, const KVLayout &kv, int n);
    void gemm(const DTensor &W, int n, float *out, int64_t out_stride, const float *residual, int64_t res_stride, const void *x_tiles = nullptr);
    void *resident_pair_tiles(const LayerW &L);
    std::unordered_map<const uint8_t *, void *> pair_tiles_; // interleaved ffn_gate | ffn_up operand tiles, keyed by the gate pointer
    void *gemm_x2_ = nullptr;                                // operand tiles written by a GEMM epilogue (SwiGLU output -> ffn_down)
    size_t gemm_x2_bytes_ = 0;
    bool no_fused_swiglu_ = false;                           // B200LLM_NO_FUSED_SWIGLU=1 (A/B tests)
    void gemm_multi(const DTensor *const *Ws, float *const *outs, const int64_t *strides, int n_jobs, int n);
    void ensure_gemm_scratch(size_t w_bytes, size_t x_bytes, size_t x2_bytes = 0);
    void *resident_tiles(const DTensor &W);
    std::unordered_map<const uint8_t *, void *> w_tiles_; // resident fp16 operand tiles, keyed by the weight pointer
    bool no_shadow_ = false;
    size_t shadow_reserve_ = (size_t)6 << 30;            // leave this much HBM free
    void *gemm_w_ = nullptr, *gemm_x_ = nullptr;
    float *gemm_part_ = nullptr;
    size_t gemm_part_bytes_ = 0;
    size_t gemm_w_bytes_ = 0, gemm_x_bytes_ = 0;
    bool no_gemm_ = false;
    int gemm_min_tokens_ = -1; // B200LLM_GEMM_MIN_TOKENS; -1 = automatic
    int gemm_threshold() const;
    bool tiles_only_ = false; // some weight is in a format only the tensor-core path reads
    // batched decode (2..64 tokens): mat-muls that dequantise inside the GEMM (kernels_qgemm.cu) on a tile-major copy of the weights
    std::unordered_map<const uint8_t *, void *> q_tiles_;
    float *qg_part_ = nullptr;
    unsigned *qg_counters_ = nullptr;
    bool slots_distinct_ = false; // the current batch has one token per slot (set by forward / forward_device_inputs)
    bool qkv_bias_ = false;       // the model has attn_{q,k,v}.bias tensors
    bool no_fused_att_ = false;   // B200LLM_NO_FUSED_ATT=1: RoPE / attention / tile conversion as three launches (A/B tests)
    bool no_qgemm_ = false;   // B200LLM_NO_QGEMM=1: the fp16-tile path for every batch size (A/B tests)
    void *resident_qtiles(const DTensor &W);
    bool qgemm_multi(const DTensor *const *Ws, float *const *outs, const int64_t *strides, int n_jobs, int n, const float *residual,
                     int64_t res_stride);
    bool qgemm_layer_ready(const LayerW &L);
    void run_layer_qgemm(const LayerW &L, const KVLayout &kv, int n);
    void build_mega();
    void run_decode_mega();
    MegaPhase *mega_phases_ = nullptr;
    MegaGeo mega_geo_;         // shared-memory split of the decode kernel for this model (activation row vs weight rings)
    MegaStreams mega_streams_; // the decode kernel's copy of the weights: one contiguous stream per (CTA, warp)
    unsigned mega_tag_base_ = 0;   // LL tag base of the next launch (advances by the number of phases)
    unsigned *mega_hint_ = nullptr; // phase counter of the decode kernel (MegaArgs.hint)
    unsigned mega_hint_base_ = 0;
    // LL-word buffers ({fp32, tag} pairs) through which the phases of the decode kernel hand activations to each other
    uint2 *x_ll_ = nullptr, *q_ll_ = nullptr, *att_ll_ = nullptr, *gate_ll_ = nullptr, *kcur_ll_ = nullptr, *vcur_ll_ = nullptr,
          *xq_ll_ = nullptr, *mega_att_scores_ = nullptr, *mega_att_part_ = nullptr;
    int mega_att_parts_ = 1;
    long long *mega_dbg_ = nullptr;
    std::vector<cudaEvent_t> mega_ev_;
    size_t mega_ev_n_ = 0;
    bool mega_timing_ = false;
  public:
    long long *mega_dbg() { return mega_dbg_; }
    // bench hook: CUDA events around every decode_mega_kernel launch (the roofline figure is per launch of that kernel)
    void mega_timing_begin(int max_launches);
    double mega_timing_end(); // mean ms per launch since mega_timing_begin (0: the kernel did not run)
    bool mega_ready() const { return mega_ready_; }
    int mega_n_phases() const { return mega_n_phases_; }
    int n_sm() const { return n_sm_; }
  private:
    int mega_n_phases_ = 0, n_sm_ = 0;
    bool mega_ready_ = false, no_mega_ = false;
    bool unfused_ = false, no_graph_ = false;
    std::vector<cudaGraphExec_t> decode_graphs_;
    std::vector<int64_t> decode_graph_launches_;

    std::vector<void *> allocs_;
    uint8_t *staging_ = nullptr;
    size_t staging_bytes_ = 0;
    void *dalloc(size_t bytes);

    // KV cache
    __half *k_pool_ = nullptr, *v_pool_ = nullptr; // [n_layer][n_pages][page_size][kvd]
    int32_t *page_table_ = nullptr;                // [n_parallel][max_pages]
    int64_t pool_layer_stride_ = 0;
    float *rope_table_ = nullptr; // [n_ctx_slot][head_dim] (cos, sin) pairs

    // scratch (sized for n_batch tokens)
    float *x_ = nullptr, *xn_ = nullptr, *q_ = nullptr, *k_ = nullptr, *v_ = nullptr, *att_ = nullptr, *gate_ = nullptr,
          *up_ = nullptr, *logits_ = nullptr, *xn_out_ = nullptr, *dbg_hidden_ = nullptr;
    uint8_t *actq_[4] = {nullptr, nullptr, nullptr, nullptr}; // per ActFmt quantised activation rows
    bool actq_valid_[4] = {false, false, false, false};
    int32_t *tok_ = nullptr, *pos_ = nullptr, *slot_ = nullptr, *out_rows_ = nullptr, *next_tok_ = nullptr;
    std::vector<std::pair<int32_t, float>> sample_specs_held_;
    std::vector<int32_t> in_host_; // staging of a step's tokens | positions | slots (pageable: the copy returns once it is staged)
    int logits_rows_cap_ = 0;
    SampleSpec *sample_specs_ = nullptr;
    float *sample_mu_ = nullptr;
    uint32_t *token_mask_ = nullptr;
    void ensure_logits_rows(int n);
    void invalidate_acts();
};

} // namespace b2
