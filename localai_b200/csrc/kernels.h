// kernels.h -- launch wrappers for the hand-written sm_100a kernels (kernels_*.cu).
#pragma once
#include <cuda_fp16.h>

#include "common.h"

namespace b2 {

// Counts kernel launches made through these wrappers (bench.py reports it as gpu_launches).
extern thread_local int64_t g_launch_count;

// --- load time ---------------------------------------------------------------------------------------
// GGUF row layout -> device row layout (identity copy for Q4_K/Q5_K/F16/F32, SoA repack for Q8_0/Q6_K).
void launch_repack(int type, const uint8_t *src_gguf, uint8_t *dst, int64_t rows, int64_t cols, cudaStream_t s);

// --- dequantisation ----------------------------------------------------------------------------------
// out[i, :] = dequant(W[row_ids ? row_ids[i] : row0 + i, :]) ; bit-exact with dequantize_row_* (unfused mul/sub).
void launch_dequant_rows(const DTensor &W, const int32_t *row_ids_dev, int64_t row0, int n_rows, float *out,
                         int64_t out_stride, cudaStream_t s);

// --- activation quantisation (quantize_row_q8_0_ref / quantize_row_q8_K_ref semantics) ---------------
void launch_quantize_act(ActFmt f, const float *x, int64_t x_stride, int n_tok, int64_t K, uint8_t *out, cudaStream_t s);

// --- streaming dequant-GEMV: out[t*out_stride + r] = dot(W[r], xq[t]) (+ residual[t*res_stride + r]) ----
void launch_gemv(const DTensor &W, const uint8_t *xq, ActFmt f, int n_tok, float *out, int64_t out_stride,
                 const float *residual, int64_t res_stride, cudaStream_t s);

// --- element-wise / normalisation --------------------------------------------------------------------
// y[i] = rmsnorm(x[rows ? rows[i] : i]) * w
void launch_rmsnorm(const float *x, int64_t x_stride, const int32_t *rows_dev, const float *w, float *y, int64_t y_stride,
                    int n_tok, int d, float eps, cudaStream_t s);
void launch_swiglu(const float *gate, const float *up, float *out, int64_t n, cudaStream_t s);

// --- RoPE + paged KV store -----------------------------------------------------------------------------
struct KVLayout {
    __half *k_pool;            // [n_pages][page_size][kvd]
    __half *v_pool;
    const int32_t *page_table; // [n_slots][max_pages_per_slot]
    int page_size, max_pages, kvd;
};
// q: [n_tok][n_head*hd] rotated in place; k,v: [n_tok][kvd]; pos/slot: per-token position and slot id (device)
// bq / bk / bv: optional projection biases added first (qwen2); neox: rotate pairs (i, i + hd/2) instead of (2i, 2i+1)
void launch_rope_kv_store(float *q, float *k, const float *v, const int32_t *pos, const int32_t *slot, int n_tok,
                          int n_head, int n_head_kv, int hd, const float *rope_table, const KVLayout &kv, cudaStream_t s, const float *bq = nullptr,
                          const float *bk = nullptr, const float *bv = nullptr, bool neox = false);
// out[t][h*hd + i] = sum_j softmax_j(q_t . k_j * scale) v_j over positions j <= pos[t] of slot[t]
void launch_attention(const float *q, const int32_t *pos, const int32_t *slot, int n_tok, int n_head, int n_head_kv,
                      int hd, const KVLayout &kv, float *out, int max_ctx, cudaStream_t s);

// flash-style prefill attention (kernels_fattn.cu): tokens [t0, t0 + n_q) are one sequence (`slot`), token t0 + i at
// position pos0 + i; their K/V must already be in the cache
bool fattn_supported(int hd);
// x_tiles != nullptr: the result goes into the next GEMM's fp16 operand tiles (rows t0 .. of the batch) instead of `out`
void launch_fattn_prefill(const float *q, int t0, int n_q, int pos0, int slot, int n_head, int n_head_kv, int hd,
                          const KVLayout &kv, float *out, cudaStream_t s, void *x_tiles = nullptr);

// --- fused decode mat-vec (kernels_fused.cu) -----------------------------------------------------------------
enum { EPI_PLAIN = 0, EPI_SWIGLU = 1, EPI_QKV = 2 };
struct FusedJob {
    const uint8_t *W;
    int type;
    int64_t rows, row_stride;
};
struct FusedGemv {
    // activations: either already quantised rows (xq_pre) or fp32 rows x[(x_rows ? x_rows[t] : t)] that the kernel
    // normalises (if norm_w) and quantises itself
    const uint8_t *xq_pre = nullptr;
    const float *x = nullptr;
    int64_t x_stride = 0;
    const int32_t *x_rows = nullptr; // device
    const float *norm_w = nullptr;
    float eps = 0.f;
    float *xn_out = nullptr; // optional: normalised rows [n_tok][K]
    int64_t K = 0;
    int fmt = 0; // ActFmt of every job's weights
    int n_jobs = 0;
    FusedJob job[3];
    int epi = EPI_PLAIN;
    float *out = nullptr; // plain/swiglu: [n_tok][out_stride]; qkv: q rows
    int64_t out_stride = 0;
    const float *residual = nullptr;
    int64_t res_stride = 0;
    // EPI_QKV
    const int32_t *pos = nullptr, *slot = nullptr; // device
    const float *rope_table = nullptr;
    KVLayout kv{};
    int hd = 0;
    const float *qkv_bias[3] = {nullptr, nullptr, nullptr}; // added to q / k / v before the rotation (qwen2)
    // filled by the launcher
    int64_t xrb = 0;
    int n_tasks = 0;
};
void launch_fused_gemv(FusedGemv a, int n_tok, cudaStream_t s);

// --- persistent decode mega-kernel (kernels_mega.cu) ------------------------------------------------------------
enum { PH_QKV = 0, PH_ATT = 1, PH_PLAIN = 2, PH_SWIGLU = 3 };
// Activations that one phase hands to the next travel as 8-byte {fp32 value, 32-bit tag} words ("LL" words, the idea
// of NCCL's low-latency protocol): the producer stores value and tag with ONE 64-bit store, the consumer polls the
// word itself until it carries the tag of the producing phase.  Data and "ready" flag cannot be seen out of order,
// so there is no fence, no atomic and no separate barrier round trip between two phases.
struct MegaPhase {
    int kind = PH_PLAIN;
    int64_t K = 0;
    int fmt = 0;
    int n_jobs = 0;
    FusedJob job[3];
    // input row: either plain fp32 written before the launch (x, src_p = -1) or LL words written by phase src_p
    const float *x = nullptr;
    const uint2 *x_ll = nullptr;   // PH_ATT: q
    int src_p = -1;
    const float *norm_w = nullptr; // RMSNorm weights applied before quantisation (optional)
    // output row: LL words (consumed by a later phase) or plain fp32 (logits); QKV: q; ATT: attention output
    float *out = nullptr;
    uint2 *out_ll = nullptr;
    const float *residual = nullptr; // plain fp32 residual (written before the launch)
    const uint2 *res_ll = nullptr;   // residual in LL words (written by an earlier phase; no tag check needed)
    float *xn_out = nullptr;       // optional: the normalised row (LM head -> embeddings)
    __half *k_pool = nullptr, *v_pool = nullptr; // this layer's KV pools (QKV, ATT)
    // tensor parallelism (tp.cu): a row-parallel mat-mul's output is a PARTIAL row; it is stored, as LL words, into the
    // receive buffer of every shard (peer_part[j] = shard j's buffer for this shard's partial, peer memory over NVLink).
    // The phase that consumes the row first adds the reduce_n partial rows in rank order plus the residual and
    // publishes the sum in x_ll (each CTA a slice of it), then reads x_ll like any other input.
    int n_peers = 0;
    uint2 *peer_part[8] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    int reduce_n = 0;              // > 0: x_ll must first be assembled from part_ll[reduce_n][reduce_len]
    int reduce_len = 0;
    const uint2 *part_ll = nullptr;
    uint2 *x_ll_w = nullptr;       // x_ll, writable (the reducer's output)
    const uint2 *red_res_ll = nullptr; // the residual the reducer adds: LL words (no tag check) ...
    const float *red_residual = nullptr; // ... or plain fp32 written before the launch
    uint2 *xq_ll = nullptr;        // if set: the row is quantised once per GPU (block b by CTA b) into this buffer of
                                   // LL words holding the image of the shared-memory activation row, then copied
    int units = 0;                 // row pairs (PLAIN/QKV) or rows (SWIGLU) split across the CTAs
    // lane geometry, filled by mega_phase_prepare() so that the kernel needs no integer division:
    int S = 1, s_log = 0;          // K-splits (power of two <= warps per CTA): a row is spread over S warps
    int lpr = 32, lpr_log = 5;     // lanes that share one (row, K-split); rpp = 32 / lpr rows per warp pass when lpr < 32
    int rpp = 1, np = 0;           // np: Q8_0 16-byte pieces per lane
    int units_q = 0, units_r = 0;  // units / n_cta, units % n_cta: the first units_r CTAs take units_q + 1 units
    int j0[3] = {0, 0, 0};         // first row of each matrix in the concatenated row space
    int rb[3] = {0, 0, 0};         // per matrix: bytes of one row's K-split = row pitch inside a staged tile
    int R[3] = {0, 0, 0};          // per matrix: rows per tile
};
// Where a warp's weight stream starts (one entry per (CTA, warp), built by MegaStreams).
struct MegaStreamStart {
    const uint8_t *src = nullptr; // first tile of the stream (16-byte aligned)
    uint32_t size0 = 0, size1 = 0; // bytes of the first two tiles (0: the stream has no such tile)
    int32_t first_p = 0;           // first phase in which this warp has tiles (n_phases: never)
    uint32_t total = 0;            // bytes of the whole stream
    int32_t pad[2] = {0, 0};
};
// Shared-memory geometry of the decode kernel, chosen per model at load (mega_geometry): the activation row / attention
// scratch takes what the model's longest row and context need, the weight rings get the rest of the 227 KB.
struct MegaGeo {
    int stage_bytes = 9232; // one ring stage: 16-byte tile header + tile data (multiple of 16); 2 stages per warp
    int xs_bytes = 34816;   // quantised activation row (and, together with the partial sums, the attention scratch)
};
MegaGeo mega_geometry(const struct MegaPhase *phases, int n_phases, int n_ctx_slot, int hd); // declared below
struct MegaArgs {
    MegaGeo geo;
    int ll_poll_ns = 32;     // sleep between two polls of an LL word (B200LLM_LL_POLL_NS)
    int ll_probe = 0;        // how a CTA waits for its input row before loading it (B200LLM_LL_PROBE): 0 = every warp polls its slice,
                             // 2 = one thread polls the phase counter `hint` first
    int wait_prefetch = 0;   // bytes of its stream a warp pulls into L2 when it finishes a phase (B200LLM_MEGA_WAIT_PREFETCH; 0 = off)
    unsigned *hint = nullptr; // phase counter: every CTA adds 1 (relaxed, no fence: the tags stay the proof) when it has stored a phase's outputs
    unsigned hint_base = 0;   // its value when this launch starts; the host adds n_cta * n_phases per launch
    const MegaPhase *phases = nullptr; // device
    int n_phases = 0;
    const MegaStreamStart *starts = nullptr; // device, [n_cta][warps per CTA]
    const int32_t *pos = nullptr, *slot = nullptr; // device, one sequence
    const float *rope_table = nullptr;
    const int32_t *page_table = nullptr;
    int page_size = 0, max_pages = 0, kvd = 0, n_head = 0, n_head_kv = 0, hd = 0;
    float eps = 0.f, att_scale = 0.f;
    int page_shift = 0;      // log2(page_size)
    unsigned tag_base = 0;   // LL tag of phase p's outputs = tag_base + p + 1; the host adds n_phases per launch
    uint2 *kcur_ll = nullptr, *vcur_ll = nullptr; // [kvd] the current token's K / V row (f16-rounded values) as LL words
    // attention of one head split over att_parts CTAs (power of two, att_parts * n_head <= CTAs; 1 = one CTA per head)
    int att_parts = 1, att_parts_log = 0;
    uint2 *att_scores = nullptr;     // [n_head][att_scores_stride] scores of the current layer (LL words)
    int64_t att_scores_stride = 0;
    uint2 *att_part_out = nullptr;   // [n_head][att_parts][hd] partial outputs (LL words)
    float *out_override = nullptr;   // if set: the last phase's plain output row goes here instead of phase.out (tp.cu: logits row of this step)
    float *xn_override = nullptr;    // likewise for phase.xn_out
    long long *dbg = nullptr; // optional [n_cta][n_phases][16] clock64 stamps and per-phase cycle counts (scratch/mega_dbg.py)
};
void launch_decode_mega(const MegaArgs &a, int n_sm, cudaStream_t s);
bool mega_phase_prepare(MegaPhase &ph, int n_cta, const MegaGeo &geo = MegaGeo()); // fills the lane geometry; false = shape/type the kernel has no mapping for
bool mega_phase_supported(const MegaPhase &ph);
void mega_fill_knobs(MegaArgs &a); // the environment's tuning knobs
// The weights of a phase list, repacked into one contiguous byte stream per (CTA, warp) in the order that warp
// consumes them: tile = [16-byte header][n row-splits x rb bytes].  Built once per model (device-side scatter from the
// GGUF-layout rows); the decode kernel then needs no schedule arithmetic, only "copy the next tile of my stream".
struct MegaStreams {
    uint8_t *data = nullptr;           // device, all streams
    size_t bytes = 0;                  // of data
    MegaStreamStart *starts = nullptr; // device
    int64_t n_tiles = 0;
    // phases: prepared (mega_phase_prepare) host copies.  Allocates with cudaMalloc; free() releases.
    void build(const MegaPhase *phases, int n_phases, int n_cta, cudaStream_t s, const MegaGeo &geo = MegaGeo());
    void free();
};
// host-only: plan the streams of a phase list and replay every warp's walk (throws b2::Error on an inconsistency);
// returns the stream bytes.  Needs no device: the unit test of the planner (tests/test_capi_cpu.py).
int64_t mega_plan_check(const MegaPhase *phases, int n_phases, int n_cta, int64_t *n_tiles, const MegaGeo &geo = MegaGeo());
size_t mega_attention_ctx_limit(int hd, const MegaGeo &geo = MegaGeo());

// --- tensor-core GEMM for prefill / batched decode (kernels_gemm.cu) ----------------------------------------------
size_t gemm_w_tiles_bytes(int64_t rows, int64_t cols);
size_t gemm_x_tiles_bytes(int n_tok, int64_t cols);
// fp32 activations [n_tok][K] -> fp16 operand tiles (do once per activation, shared by several matrices)
void launch_gemm_prepare_x(const float *x, int64_t x_stride, int n_tok, int64_t K, void *x_tiles, cudaStream_t s);
// the same with the producer of the activations fused in: RMSNorm(x) * w (optionally also written as fp32 rows to
// xn_out), silu(gate) * up
void launch_gemm_prepare_x_rmsnorm(const float *x, int64_t x_stride, const float *w, int n_tok, int64_t K, float eps, void *x_tiles,
                                   float *xn_out, cudaStream_t s);
void launch_gemm_prepare_x_swiglu(const float *gate, const float *up, int n_tok, int64_t K, void *x_tiles, cudaStream_t s);
// x = LayerNorm(x + h + bias) * w + b, written back as fp32 rows and as operand tiles (BERT; K <= 2048)
void launch_gemm_prepare_x_bias_residual_ln(float *x, const float *h, const float *bias, const float *w, const float *b, int n_tok,
                                            int64_t K, float eps, void *x_tiles, cudaStream_t s);
// gelu(x + bias): the BERT feed-forward activation (bert.cu)
void launch_gemm_prepare_x_bias_gelu(const float *x, const float *bias, int n_tok, int64_t K, void *x_tiles, cudaStream_t s);
// quantised W -> fp16 operand tiles (gemm_w_tiles_bytes of them)
void launch_gemm_dequant_w(const DTensor &W, void *w_tiles, cudaStream_t s);
// out[t][r] = W[r] . x[t] (+ residual): tcgen05/TMEM GEMM on fp16 operand tiles; W is dequantised into w_tiles first
// unless w_tiles_ready (the caller keeps W's tiles resident)
// splitk_scratch (optional, fp32): with <= 128 tokens and fewer output tiles than SMs the K range is split over the
// idle SMs; the slices' partial tiles go through this buffer and are summed in slice order
void launch_gemm_tc(const DTensor &W, const void *x_tiles, int n_tok, float *out, int64_t out_stride, const float *residual,
                    int64_t res_stride, void *w_tiles, bool w_tiles_ready, float *splitk_scratch, size_t splitk_bytes, cudaStream_t s);
// GEMMs whose epilogue applies the activation and writes the next GEMM's operand tiles directly:
//   SwiGLU over an interleaved ffn_gate | ffn_up tile set (launch_gemm_dequant_w_pair; ff a multiple of 128),
//   gelu(W x + bias) (the BERT feed-forward; N a multiple of 256)
size_t gemm_w_pair_tiles_bytes(int64_t rows, int64_t cols);
void launch_gemm_dequant_w_pair(const DTensor &gate, const DTensor &up, void *w_tiles, cudaStream_t s);
void launch_gemm_tc_swiglu(int64_t ff, int64_t cols, const void *pair_tiles, const void *x_tiles, int n_tok, void *x2_tiles, cudaStream_t s);
void launch_gemm_tc_bias_gelu(int64_t N, int64_t cols, const void *w_tiles, const void *x_tiles, int n_tok, const float *bias, void *x2_tiles,
                              cudaStream_t s);
// up to three matrices that share the activation (wq|wk|wv, gate|up) in ONE launch; residual only with one matrix
struct GemmJob {
    const DTensor *W;
    void *w_tiles;
    bool w_tiles_ready;
    float *out;
    int64_t out_stride;
    const float *bias = nullptr; // added in the epilogue (such a launch is never split over K)
};
// out_f16: the outputs are rows of fp16 (out_stride in halves) with the matrices' biases added; never split over K
void launch_gemm_tc_multi(const GemmJob *jobs, int n_jobs, const void *x_tiles, int n_tok, const float *residual, int64_t res_stride,
                          float *splitk_scratch, size_t splitk_bytes, cudaStream_t s, bool out_f16 = false);

// --- batched decode mat-mul with the dequantisation inside the GEMM (kernels_qgemm.cu) --------------------------------
constexpr int QGEMM_MAX_TOKENS = 64;
bool qgemm_type_supported(int type);                          // Q4_K, Q5_K, Q6_K, Q8_0
size_t qgemm_w_bytes(int type, int64_t rows, int64_t cols);   // of the tile-major repacked copy of a matrix
size_t qgemm_x_bytes(int n_tok, int64_t cols);                // of the fp16 activation tiles
void launch_qgemm_repack(const DTensor &W, void *q_tiles, cudaStream_t s);
// fp32 activations -> fp16 operand tiles (plain, RMSNorm(x) * w, silu(gate) * up)
void launch_qgemm_prepare_x(const float *x, int64_t x_stride, int n_tok, int64_t K, void *x_tiles, cudaStream_t s);
void launch_qgemm_prepare_x_rmsnorm(const float *x, int64_t x_stride, const float *w, int n_tok, int64_t K, float eps, void *x_tiles,
                                    cudaStream_t s);
void launch_qgemm_prepare_x_swiglu(const float *gate, const float *up, int n_tok, int64_t K, void *x_tiles, cudaStream_t s);
// decode batches whose tokens sit in different slots: RoPE + KV store + attention + fp16 activation tiles in one launch;
// false = no mapping for this head geometry / context (the caller runs the three separate kernels)
bool launch_attention_decode_fused(float *q, const float *k, const float *v, const int32_t *pos, const int32_t *slot, int n_tok, int n_head,
                                   int n_head_kv, int hd, const float *rope_table, const KVLayout &kv, int max_ctx, void *x_tiles, int n_pad,
                                   cudaStream_t s);
struct QGemmJob {
    const DTensor *W;
    const void *q_tiles; // launch_qgemm_repack(W)
    float *out;
    int64_t out_stride;
};
struct QGemmPlan {
    int T, panels, grid; // tiles per panel, panels, CTAs
    int64_t units;       // panels * K / 256
};
int qgemm_debug_read(long long *host, int n); // profiling builds (-DQG_DBG) only; 0 otherwise
void qgemm_plan(int n_tiles, int nkb, int n_sm, bool allow_pairs, QGemmPlan *pl);
int64_t qgemm_plan_check(int n_tiles, int nkb, int n_sm); // host-only replay of the kernel's work split; throws on an inconsistency
inline size_t qgemm_partial_bytes(int n_sm) { return (size_t)2 * n_sm * 2 * 64 * 128 * 4; } // two partial segments (of two 128 x 64 tiles) per CTA at most
constexpr size_t QGEMM_COUNTERS = 1 << 16;                                   // tile quarters: 16384 tiles = 2 M rows
// out[t][r] = W[r] . x[t] (+ residual) for up to three matrices that share x; 1..64 tokens; one launch, split-K reduced inside
void launch_qgemm_multi(const QGemmJob *jobs, int n_jobs, const void *x_tiles, int n_tok, const float *residual, int64_t res_stride,
                        float *partial, size_t partial_bytes, unsigned *counters, size_t n_counters, int n_sm, cudaStream_t s);

// --- sampling -------------------------------------------------------------------------------------------
// out[r] = first argmax of row r.  If feed_tok/feed_pos are given (device-side decode loop) the sampled id is also
// written to feed_tok[r] and feed_pos[r] is incremented, so the next step needs no host round trip.
void launch_argmax(const float *logits, int64_t stride, int n_rows, int n_vocab, int32_t *out, int32_t *feed_tok,
                   int32_t *feed_pos, cudaStream_t s);

} // namespace b2
