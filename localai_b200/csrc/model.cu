// model.cu -- LoadModel (GGUF -> HBM) and the llama forward graph.
//
// Replaces llama_init_from_gpt_params + llama_decode as called from the reference engine
// (backend/cpp/llama/grpc-server.cpp:461, :1910).  Graph = llm_build_llama.
#include "model.h"
#include "sampling.h"

#include <cmath>
#include <cstdlib>
#include <cstring>

namespace b2 {

void *Model::dalloc(size_t bytes) {
    void *p = nullptr;
    B2_CUDA(cudaMalloc(&p, bytes ? bytes : 16));
    allocs_.push_back(p);
    return p;
}

DTensor Model::upload(const GGUFFile &f, const std::string &name, int64_t want_rows, int64_t want_cols, bool required,
                      int64_t row0, int64_t nrows, int64_t col0, int64_t ncols, int pair_heads) {
    const GGUFTensorInfo *ti = f.tensor(name);
    if (!ti) {
        B2_CHECK(!required, B2_EIO, "gguf: missing tensor " + name);
        return DTensor{};
    }
    B2_CHECK(ti->n_bytes >= 0, B2_EUNSUPPORTED, "gguf: tensor " + name + " has unsupported ggml type " + std::to_string(ti->type));
    const int64_t cols_all = ti->ne[0];
    const int64_t rows_all = ti->ne.size() > 1 ? ti->ne[1] : 1;
    B2_CHECK(cols_all == want_cols && rows_all == want_rows, B2_EIO,
             "gguf: tensor " + name + " has shape [" + std::to_string(rows_all) + "," + std::to_string(cols_all) + "], expected [" +
                 std::to_string(want_rows) + "," + std::to_string(want_cols) + "]");
    // the slice this shard keeps (everything when the model is not tensor-parallel)
    const int64_t rows = nrows < 0 ? rows_all : nrows, cols = ncols < 0 ? cols_all : ncols;
    const int be = type_block_elems(ti->type);
    B2_CHECK(row0 >= 0 && row0 + rows <= rows_all && col0 >= 0 && col0 + cols <= cols_all && col0 % be == 0 && cols % be == 0, B2_EINVAL,
             "tensor " + name + ": slice outside the tensor or off the block boundaries");
    if (type_tiles_only(ti->type)) {
        B2_CHECK(cols % std::max(64, be) == 0, B2_EUNSUPPORTED, "tensor " + name + ": row length must be a multiple of 64 and of the block size");
        tiles_only_ = true; // these formats have no integer decode kernel: everything goes through the fp16 tiles
    } else if (ti->type != B2_T_F32 && ti->type != B2_T_F16) {
        B2_CHECK(cols % 256 == 0, B2_EUNSUPPORTED, "tensor " + name + ": quantised row length must be a multiple of 256");
    }
    DTensor t;
    t.type = ti->type;
    t.rows = rows;
    t.cols = cols;
    t.row_stride = device_row_stride(ti->type, cols);
    uint8_t *dst = (uint8_t *)dalloc((size_t)t.bytes());
    const int64_t row_bytes_all = gguf_row_bytes(ti->type, cols_all), row_bytes = gguf_row_bytes(ti->type, cols);
    const size_t n_bytes = (size_t)(rows * row_bytes);
    if (n_bytes > staging_bytes_) {
        if (staging_) cudaFree(staging_);
        staging_bytes_ = n_bytes;
        B2_CUDA(cudaMalloc((void **)&staging_, staging_bytes_));
    }
    const uint8_t *src = ti->data + row0 * row_bytes_all;
    std::vector<uint8_t> shuffled;
    if (pair_heads > 0) {
        // NEOX rotary pairs (i, i + h/2) of every head become neighbours (2i, 2i + 1): the rotation the llama kernels apply to
        // neighbouring rows is then the architecture's own, on the same numbers, and q . k does not care in which order its
        // dimensions are summed as long as q and k agree (llama.cpp's converter does the same to HF llama checkpoints).
        // A row is a byte range whatever the quantisation, so this is a copy of rows.
        B2_CHECK(cols == cols_all && rows % pair_heads == 0 && pair_heads % 2 == 0, B2_EINVAL, "tensor " + name + ": cannot interleave the rows of its heads");
        shuffled.resize(n_bytes);
        const int64_t half = pair_heads / 2;
        for (int64_t j = 0; j < rows; j++) {
            const int64_t h0 = j / pair_heads * pair_heads, i = j % pair_heads;
            const int64_t from = h0 + (i & 1 ? i / 2 + half : i / 2);
            memcpy(shuffled.data() + j * row_bytes, src + from * row_bytes, (size_t)row_bytes);
        }
        src = shuffled.data();
    }
    if (cols == cols_all) { // whole rows: one contiguous range of the file
        B2_CUDA(cudaMemcpyAsync(staging_, src, n_bytes, cudaMemcpyHostToDevice, stream));
    } else { // a K-slice: the blocks [col0 / be, (col0 + cols) / be) of every row
        B2_CUDA(cudaMemcpy2DAsync(staging_, (size_t)row_bytes, src + gguf_row_bytes(ti->type, col0), (size_t)row_bytes_all, (size_t)row_bytes,
                                  (size_t)rows, cudaMemcpyHostToDevice, stream));
    }
    launch_repack(t.type, staging_, dst, rows, cols, stream);
    B2_CUDA(cudaStreamSynchronize(stream)); // staging buffer is reused by the next tensor
    t.data = dst;
    weight_bytes += t.bytes();
    return t;
}

Model::Model(const GGUFFile &f, const b2_load_opts &o, int tp_rank_arg, int tp_size_arg) {
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) throw Error(B2_ENODEV, "no CUDA device visible (this library has no CPU fallback)");
    tp_size = tp_size_arg > 0 ? tp_size_arg : (o.struct_size >= offsetof(b2_load_opts, tp_size) + sizeof(o.tp_size) && o.tp_size > 1 ? o.tp_size : 1);
    tp_rank = tp_rank_arg;
    device = o.device + tp_rank;
    B2_CHECK(device >= 0 && device < ndev, B2_EINVAL, tp_size > 1 ? "tensor-parallel model: not enough devices" : "device ordinal out of range");
    B2_CUDA(cudaSetDevice(device));
    cudaDeviceProp prop;
    B2_CUDA(cudaGetDeviceProperties(&prop, device));
    B2_CHECK(prop.major == 10, B2_ENODEV,
             std::string("device is sm_") + std::to_string(prop.major * 10 + prop.minor) + "; this library is built for sm_100a only");
    B2_CUDA(cudaStreamCreateWithFlags(&stream, cudaStreamNonBlocking));
    unfused_ = getenv("B200LLM_UNFUSED") && atoi(getenv("B200LLM_UNFUSED")) != 0;
    no_graph_ = getenv("B200LLM_NO_GRAPH") && atoi(getenv("B200LLM_NO_GRAPH")) != 0;
    no_mega_ = getenv("B200LLM_NO_MEGA") && atoi(getenv("B200LLM_NO_MEGA")) != 0;
    no_gemm_ = getenv("B200LLM_NO_GEMM") && atoi(getenv("B200LLM_NO_GEMM")) != 0;
    no_qgemm_ = getenv("B200LLM_NO_QGEMM") && atoi(getenv("B200LLM_NO_QGEMM")) != 0;
    no_fused_att_ = getenv("B200LLM_NO_FUSED_ATT") && atoi(getenv("B200LLM_NO_FUSED_ATT")) != 0;
    no_fused_swiglu_ = getenv("B200LLM_NO_FUSED_SWIGLU") && atoi(getenv("B200LLM_NO_FUSED_SWIGLU")) != 0;
    no_shadow_ = getenv("B200LLM_NO_BF16_TILES") && atoi(getenv("B200LLM_NO_BF16_TILES")) != 0;
    if (getenv("B200LLM_GEMM_MIN_TOKENS")) gemm_min_tokens_ = atoi(getenv("B200LLM_GEMM_MIN_TOKENS"));
    else gemm_min_tokens_ = -1; // automatic, see gemm_threshold()

    hp.arch = f.get_s("general.architecture", "");
    B2_CHECK(hp.arch == "llama" || hp.arch == "qwen2" || hp.arch == "phi3", B2_EUNSUPPORTED,
             "unsupported architecture '" + hp.arch + "' (llama, qwen2 and phi3 GGUFs)");
    hp.rope_neox = hp.arch == "qwen2" || hp.arch == "phi3"; // llama.cpp llama_rope_type: NEOX pairs (i, i + head_dim / 2)
    const bool fused_qkv = hp.arch == "phi3"; // attn_qkv.weight = [q; k; v] rows, ffn_up.weight = [gate; up] rows
    hp.rope_neox_kernel = hp.rope_neox && getenv("B200LLM_NEOX_KERNEL") && atoi(getenv("B200LLM_NEOX_KERNEL")) != 0;
    hp.name = f.get_s("general.name", "");
    const std::string a = hp.arch + ".";
    hp.n_layer = (int)f.get_u(a + "block_count", 0);
    hp.n_embd = (int)f.get_u(a + "embedding_length", 0);
    hp.n_head = (int)f.get_u(a + "attention.head_count", 0);
    hp.n_head_kv = (int)f.get_u(a + "attention.head_count_kv", hp.n_head);
    hp.n_ff = (int)f.get_u(a + "feed_forward_length", 0);
    hp.n_ctx_train = (int)f.get_u(a + "context_length", 2048);
    hp.rms_eps = (float)f.get_f(a + "attention.layer_norm_rms_epsilon", 1e-5);
    hp.rope_base = (float)f.get_f(a + "rope.freq_base", 10000.0);
    hp.rope_scale = 1.0f;
    B2_CHECK(hp.n_layer > 0 && hp.n_embd > 0 && hp.n_head > 0 && hp.n_ff > 0, B2_EIO, "gguf: missing llama hyper-parameters");
    hp.head_dim = hp.n_embd / hp.n_head;
    const int rope_dims = (int)f.get_u(a + "rope.dimension_count", hp.head_dim);
    B2_CHECK(rope_dims == hp.head_dim, B2_EUNSUPPORTED, "partial rotary dimensions are not supported");
    B2_CHECK(hp.n_head % hp.n_head_kv == 0, B2_EIO, "head_count not a multiple of head_count_kv");
    // RoPE scaling as llama.cpp derives it from the metadata (llm_load_hparams, llama_new_context_with_model): linear and
    // YaRN scale the frequencies by 1 / rope.scaling.factor, YaRN additionally blends towards the unscaled frequencies
    // (ext_factor 1) inside its correction range; rope_freqs.weight holds per-dimension divisors (the Llama-3.1 scheme)
    {
        std::string st = f.get_s(a + "rope.scaling.type", "none");
        const float factor = (float)f.get_f(a + "rope.scaling.factor", 1.0);
        B2_CHECK(st == "none" || st == "linear" || st == "yarn", B2_EUNSUPPORTED, "gguf: unknown rope.scaling.type");
        B2_CHECK(o.rope_scaling >= B2_ROPE_FROM_GGUF && o.rope_scaling <= B2_ROPE_YARN, B2_EINVAL, "rope_scaling: 0 .. 3");
        if (o.rope_scaling == B2_ROPE_NONE) st = "none"; // ModelOptions.RopeScaling overrides the file's type (grpc-server.cpp:2279-2281)
        else if (o.rope_scaling == B2_ROPE_LINEAR) st = "linear";
        else if (o.rope_scaling == B2_ROPE_YARN) st = "yarn";
        // the trained scale 1 / factor applies whatever the type (llama.cpp: rope_freq_scale_train is read unconditionally)
        if (factor > 0) hp.rope_scale = 1.0f / factor;
        if (st == "none") hp.rope_scale = 1.0f; // LLAMA_ROPE_SCALING_TYPE_NONE: never scale contexts
        hp.rope_ext_factor = o.yarn_ext_factor != 0.f ? o.yarn_ext_factor : (st == "yarn" ? 1.0f : 0.0f);
        hp.rope_attn_factor = o.yarn_attn_factor != 0.f ? o.yarn_attn_factor : 1.0f;
        hp.rope_beta_fast = o.yarn_beta_fast != 0.f ? o.yarn_beta_fast : 32.0f;
        hp.rope_beta_slow = o.yarn_beta_slow != 0.f ? o.yarn_beta_slow : 1.0f;
        hp.rope_n_ctx_orig = (int)f.get_u(a + "rope.scaling.original_context_length", hp.n_ctx_train);
    }
    if (o.rope_freq_base > 0) hp.rope_base = o.rope_freq_base;
    if (o.rope_freq_scale > 0) hp.rope_scale = o.rope_freq_scale;

    const GGUFTensorInfo *te = f.tensor("token_embd.weight");
    B2_CHECK(te && te->ne.size() == 2, B2_EIO, "gguf: missing token_embd.weight");
    hp.n_vocab = (int)te->ne[1];

    n_ctx = o.n_ctx > 0 ? o.n_ctx : 1024;
    n_parallel = o.n_parallel > 0 ? o.n_parallel : 1;
    n_batch = o.n_batch > 0 ? o.n_batch : 512;
    page_size = o.kv_page_size > 0 ? o.kv_page_size : 32;
    n_ctx_slot = n_ctx / n_parallel; // grpc-server.cpp:512
    B2_CHECK(n_ctx_slot > 0, B2_EINVAL, "context size smaller than the number of slots");
    max_pages = (n_ctx_slot + page_size - 1) / page_size;

    const int d = hp.n_embd;
    B2_CHECK(tp_size == 1 || (hp.n_head % tp_size == 0 && hp.n_head_kv % tp_size == 0 && (hp.n_ff / tp_size) % 256 == 0 && hp.n_ff % tp_size == 0 &&
                              (d / tp_size) % 256 == 0 && hp.n_vocab % (2 * tp_size) == 0),
             B2_EUNSUPPORTED, "tensor-parallel degree does not divide the heads / KV heads / 256-weight blocks of this model");
    nh_l = hp.n_head / tp_size;
    nkv_l = hp.n_head_kv / tp_size;
    ff_l = hp.n_ff / tp_size;
    dq_l = nh_l * hp.head_dim;
    kvd_l = nkv_l * hp.head_dim;
    vocab_l = hp.n_vocab / tp_size;
    const int ff = hp.n_ff, kvd = hp.n_head_kv * hp.head_dim, r = tp_rank;
    // ---- weights, in execution order (a tensor-parallel shard keeps its rows / its K-slice only)
    tok_embd = upload(f, "token_embd.weight", hp.n_vocab, d, true);
    const int64_t embd_bytes = tok_embd.bytes();
    layers.resize(hp.n_layer);
    for (int l = 0; l < hp.n_layer; l++) {
        const std::string p = "blk." + std::to_string(l) + ".";
        LayerW &L = layers[l];
        L.attn_norm = upload(f, p + "attn_norm.weight", 1, d, true);
        // NEOX architectures: the rows of wq / wk (and their biases) are interleaved per head, after which the model is a
        // llama model to every kernel (see upload)
        const int pair = (hp.rope_neox && !hp.rope_neox_kernel) ? hp.head_dim : 0;
        if (fused_qkv) { // rows are contiguous in the file: the three projections are row ranges of the one tensor
            L.wq = upload(f, p + "attn_qkv.weight", d + 2 * kvd, d, true, 0, d, 0, -1, pair);
            L.wk = upload(f, p + "attn_qkv.weight", d + 2 * kvd, d, true, d, kvd, 0, -1, pair);
            L.wv = upload(f, p + "attn_qkv.weight", d + 2 * kvd, d, true, d + kvd, kvd);
        } else {
            L.wq = upload(f, p + "attn_q.weight", d, d, true, (int64_t)r * dq_l, dq_l, 0, -1, pair);
            L.wk = upload(f, p + "attn_k.weight", kvd, d, true, (int64_t)r * kvd_l, kvd_l, 0, -1, pair);
            L.wv = upload(f, p + "attn_v.weight", kvd, d, true, (int64_t)r * kvd_l, kvd_l);
        }
        L.wo = upload(f, p + "attn_output.weight", d, d, true, 0, -1, (int64_t)r * dq_l, dq_l);
        if (f.tensor(p + "attn_q.bias")) { // qwen2: biases of the three projections (F32 rows)
            B2_CHECK(tp_size == 1, B2_EUNSUPPORTED, "tensor-parallel loading of models with QKV biases is not built");
            L.bq = upload(f, p + "attn_q.bias", 1, d, true);
            L.bk = upload(f, p + "attn_k.bias", 1, kvd, true);
            L.bv = upload(f, p + "attn_v.bias", 1, kvd, true);
            B2_CHECK(L.bq.type == B2_T_F32 && L.bk.type == B2_T_F32 && L.bv.type == B2_T_F32, B2_EUNSUPPORTED, "QKV biases must be F32");
            qkv_bias_ = true;
            if (pair > 0) { // the biases of q and k follow their rows
                for (const auto &nb : {std::make_pair("attn_q.bias", &L.bq), std::make_pair("attn_k.bias", &L.bk)}) {
                    const float *src = (const float *)f.tensor(p + nb.first)->data;
                    const int64_t n = nb.second->cols, half = pair / 2;
                    std::vector<float> v((size_t)n);
                    for (int64_t j = 0; j < n; j++) {
                        const int64_t h0 = j / pair * pair, i = j % pair;
                        v[(size_t)j] = src[h0 + (i & 1 ? i / 2 + half : i / 2)];
                    }
                    B2_CUDA(cudaMemcpy((void *)nb.second->data, v.data(), (size_t)n * 4, cudaMemcpyHostToDevice));
                }
            }
        }
        L.ffn_norm = upload(f, p + "ffn_norm.weight", 1, d, true);
        if (fused_qkv) { // llama.cpp build_phi3 / LLM_FFN_SWIGLU on one tensor: silu(first half) * second half
            L.gate = upload(f, p + "ffn_up.weight", 2 * (int64_t)ff, d, true, 0, ff);
            L.up = upload(f, p + "ffn_up.weight", 2 * (int64_t)ff, d, true, ff, ff);
        } else {
            L.gate = upload(f, p + "ffn_gate.weight", ff, d, true, (int64_t)r * ff_l, ff_l);
            L.up = upload(f, p + "ffn_up.weight", ff, d, true, (int64_t)r * ff_l, ff_l);
        }
        L.down = upload(f, p + "ffn_down.weight", d, ff, true, 0, -1, (int64_t)r * ff_l, ff_l);
        B2_CHECK(L.attn_norm.type == B2_T_F32 && L.ffn_norm.type == B2_T_F32, B2_EUNSUPPORTED, "norm weights must be F32");
    }
    if (fused_qkv) {
        // the long-context variants carry per-dimension frequency factors chosen by context length, and the model attends
        // through a sliding window beyond phi3.attention.sliding_window positions: neither is built
        B2_CHECK(!f.tensor("rope_factors_long.weight") && !f.tensor("rope_factors_short.weight"), B2_EUNSUPPORTED,
                 "phi3: long-context rope factors (rope_factors_long / _short) are not supported");
        const int64_t window = (int64_t)f.get_u(a + "attention.sliding_window", 0);
        B2_CHECK(window <= 0 || n_ctx_slot <= window, B2_EUNSUPPORTED,
                 "phi3: contexts beyond the model's sliding window (" + std::to_string(window) + " positions per slot) are not supported");
    }
    B2_CHECK(tp_size == 1 || hp.arch == "llama", B2_EUNSUPPORTED, "tensor-parallel loading is built for the llama architecture only");
    if (qkv_bias_) { // the biases are added by rope_kv_store_kernel and by the fused mat-vec's QKV epilogue: no persistent decode
        no_mega_ = true; // kernel, no fused RoPE + attention kernel for such models; everything else is shared
        no_fused_att_ = true;
    }
    if (hp.rope_neox_kernel) { // A/B switch: NEOX pairs rotated by rope_kv_store_kernel on un-interleaved rows, per-operator path only
        no_mega_ = true;
        unfused_ = true;
        no_fused_att_ = true;
    }
    output_norm = upload(f, "output_norm.weight", 1, d, true);
    B2_CHECK(output_norm.type == B2_T_F32, B2_EUNSUPPORTED, "norm weights must be F32");
    B2_CHECK(!(tiles_only_ && no_gemm_), B2_EUNSUPPORTED, "Q4_0 / Q4_1 / Q5_0 / Q5_1 / Q2_K / Q3_K / IQ4 weights need the tensor-core path (B200LLM_NO_GEMM is set)");
    B2_CHECK(!(tiles_only_ && tp_size > 1), B2_EUNSUPPORTED, "tensor-parallel models need weights with an integer decode kernel (Q4_K, Q5_K, Q6_K, Q8_0)");
    output = upload(f, "output.weight", hp.n_vocab, d, false, (int64_t)r * vocab_l, vocab_l);
    if (!output.data) { // tied embeddings
        B2_CHECK(tp_size == 1, B2_EUNSUPPORTED, "tensor-parallel models need an output.weight tensor");
        output = tok_embd;
        decode_stream_bytes = weight_bytes;
    } else {
        decode_stream_bytes = weight_bytes - embd_bytes;
    }
    if (staging_) {
        cudaFree(staging_);
        staging_ = nullptr;
        staging_bytes_ = 0;
    }

    // ---- KV cache (F16, paged; slot s statically owns pages [s*max_pages, (s+1)*max_pages))
    const int64_t n_pages = (int64_t)n_parallel * max_pages;
    pool_layer_stride_ = n_pages * page_size * kvd_l;
    const size_t pool_bytes = (size_t)pool_layer_stride_ * hp.n_layer * sizeof(__half);
    k_pool_ = (__half *)dalloc(pool_bytes);
    v_pool_ = (__half *)dalloc(pool_bytes);
    B2_CUDA(cudaMemsetAsync(k_pool_, 0, pool_bytes, stream));
    B2_CUDA(cudaMemsetAsync(v_pool_, 0, pool_bytes, stream));
    kv_bytes_per_pos = (int64_t)2 * hp.n_layer * kvd_l * sizeof(__half);
    {
        std::vector<int32_t> pt((size_t)n_parallel * max_pages);
        for (size_t i = 0; i < pt.size(); i++) pt[i] = (int32_t)i;
        page_table_ = (int32_t *)dalloc(pt.size() * 4);
        B2_CUDA(cudaMemcpyAsync(page_table_, pt.data(), pt.size() * 4, cudaMemcpyHostToDevice, stream));
        B2_CUDA(cudaStreamSynchronize(stream));
    }
    // ---- RoPE table, computed on the host exactly as ggml_rope_cache_init does (theta *= theta_scale, cosf/sinf)
    {
        const int hd = hp.head_dim;
        std::vector<float> tab((size_t)n_ctx_slot * hd);
        const float theta_scale = powf(hp.rope_base, -2.0f / hd);
        // per-dimension frequency divisors (rope_freqs.weight, f32 [hd / 2]), read straight from the mapping
        const float *ff = nullptr;
        if (const GGUFTensorInfo *tf = f.tensor("rope_freqs.weight")) {
            B2_CHECK(tf->type == B2_T_F32 && tf->n_bytes == (int64_t)(hd / 2) * 4 && tf->data, B2_EIO, "gguf: rope_freqs.weight must be f32 [head_dim / 2]");
            ff = (const float *)tf->data;
        }
        // YaRN correction range (ggml_rope_yarn_corr_dims, beta_fast 32, beta_slow 1)
        float corr[2] = {0.f, 0.f};
        if (hp.rope_ext_factor != 0.0f) {
            auto corr_dim = [&](float n_rot) { return hd * logf(hp.rope_n_ctx_orig / (n_rot * 2 * (float)M_PI)) / (2 * logf(hp.rope_base)); };
            corr[0] = std::max(0.f, floorf(corr_dim(hp.rope_beta_fast)));
            corr[1] = std::min((float)hd - 1, ceilf(corr_dim(hp.rope_beta_slow)));
        }
        for (int p = 0; p < n_ctx_slot; p++) {
            float theta = (float)p;
            for (int i0 = 0; i0 < hd; i0 += 2) {
                float t, mscale = hp.rope_attn_factor;
                if (!ff && hp.rope_ext_factor == 0.0f && hp.rope_attn_factor == 1.0f) {
                    t = hp.rope_scale * theta; // the unscaled / linear path keeps its exact arithmetic
                } else {
                    const float extrap = theta / (ff ? ff[i0 / 2] : 1.0f), interp = hp.rope_scale * extrap;
                    t = interp;
                    if (hp.rope_ext_factor != 0.0f) { // rope_yarn
                        const float y = (i0 / 2 - corr[0]) / std::max(0.001f, corr[1] - corr[0]);
                        const float ramp_mix = (1.0f - std::min(1.0f, std::max(0.0f, y))) * hp.rope_ext_factor;
                        t = interp * (1 - ramp_mix) + extrap * ramp_mix;
                        mscale *= 1.0f + 0.1f * logf(1.0f / hp.rope_scale);
                    }
                }
                tab[(size_t)p * hd + i0] = cosf(t) * mscale;
                tab[(size_t)p * hd + i0 + 1] = sinf(t) * mscale;
                theta *= theta_scale;
            }
        }
        rope_table_ = (float *)dalloc(tab.size() * 4);
        B2_CUDA(cudaMemcpyAsync(rope_table_, tab.data(), tab.size() * 4, cudaMemcpyHostToDevice, stream));
        B2_CUDA(cudaStreamSynchronize(stream));
    }
    // ---- scratch
    const size_t nb = (size_t)n_batch;
    x_ = (float *)dalloc(nb * d * 4);
    xn_ = (float *)dalloc(nb * d * 4);
    q_ = (float *)dalloc(nb * dq_l * 4);
    k_ = (float *)dalloc(nb * kvd_l * 4);
    v_ = (float *)dalloc(nb * kvd_l * 4);
    att_ = (float *)dalloc(nb * dq_l * 4);
    gate_ = (float *)dalloc(nb * ff_l * 4);
    up_ = (float *)dalloc(nb * ff_l * 4);
    if (tp_size > 1) xp_ = (float *)dalloc(nb * d * 4);
    const int64_t kmax = std::max(d, ff_l);
    for (int fmt = ACT_F16; fmt <= ACT_Q8_K; fmt++) actq_[fmt] = (uint8_t *)dalloc(nb * (size_t)act_row_bytes((ActFmt)fmt, kmax));
    tok_ = (int32_t *)dalloc(3 * nb * 4); // tokens | positions | slots, one allocation: a step's inputs go up in ONE copy
    pos_ = tok_ + nb;
    slot_ = pos_ + nb;
    in_host_.resize(3 * nb);
    out_rows_ = (int32_t *)dalloc(nb * 4);
    next_tok_ = (int32_t *)dalloc(nb * 4);
    xn_out_ = (float *)dalloc(nb * d * 4);
    sample_specs_ = (SampleSpec *)dalloc(nb * sizeof(SampleSpec));
    sample_mu_ = (float *)dalloc(nb * sizeof(float));
    if (tp_size > 1) {
        B2_CUDA(cudaEventCreateWithFlags(&tp_done_, cudaEventDisableTiming));
        B2_CUDA(cudaEventCreateWithFlags(&tp_in_, cudaEventDisableTiming));
    }
    // a tensor-parallel shard never re-allocates its logits rows (the fused decode kernels' phase lists hold the address):
    // one row per token of the largest batch, plus a spare one for rows nobody asked for
    ensure_logits_rows(tp_size > 1 ? n_batch + 1 : std::max(n_parallel, 1));
    if (tp_size == 1) build_mega();
    decode_graphs_.assign((size_t)std::min(n_parallel, 64) + 1, nullptr);
    decode_graph_launches_.assign(decode_graphs_.size(), 0);
    if (tp_size > 1 && tp_rank == 0) tp_init(f, o); // creates the other shards, the communicators and the fused decode kernels' phase lists
}

Model::~Model() {
    cudaSetDevice(device);
    if (stream) cudaStreamSynchronize(stream);
    for (auto g : decode_graphs_)
        if (g) cudaGraphExecDestroy(g);
    tp_destroy(); // joins the shard threads and frees the other shards first
    if (tp_done_) cudaEventDestroy(tp_done_);
    if (tp_in_) cudaEventDestroy(tp_in_);
    for (void *p : allocs_) cudaFree(p);
    for (auto e : mega_ev_) cudaEventDestroy(e);
    mega_streams_.free();
    if (logits_) cudaFree(logits_);
    for (auto &kv : w_tiles_) cudaFree(kv.second);
    for (auto &kv : q_tiles_) cudaFree(kv.second);
    for (auto &kv : pair_tiles_) cudaFree(kv.second);
    if (gemm_x2_) cudaFree(gemm_x2_);
    if (gemm_w_) cudaFree(gemm_w_);
    if (gemm_x_) cudaFree(gemm_x_);
    if (staging_) cudaFree(staging_);
    if (stream) cudaStreamDestroy(stream);
}

void Model::ensure_logits_rows(int n) {
    if (n <= logits_rows_cap_) return;
    B2_CUDA(cudaStreamSynchronize(stream));
    if (logits_) cudaFree(logits_);
    logits_ = nullptr;
    B2_CUDA(cudaMalloc((void **)&logits_, (size_t)n * hp.n_vocab * 4));
    logits_rows_cap_ = n;
    B2_CHECK(tp_size == 1 || !mega_phases_, B2_ESTATE, "tensor-parallel shard: the logits rows are fixed");
    if (mega_phases_) build_mega(); // the LM-head phase holds the logits pointer
    for (auto &g : decode_graphs_) { // captured graphs hold the old logits pointer
        if (g) cudaGraphExecDestroy(g);
        g = nullptr;
    }
}

SampleSpec *Model::sample_specs_dev(int n) {
    B2_CHECK(n <= n_batch, B2_EINVAL, "too many rows to sample");
    return sample_specs_;
}

uint32_t *Model::token_mask_dev() {
    if (!token_mask_) token_mask_ = (uint32_t *)dalloc((size_t)((hp.n_vocab + 31) / 32) * 4);
    return token_mask_;
}

void Model::enable_debug_hidden() {
    if (!dbg_hidden_) dbg_hidden_ = (float *)dalloc((size_t)(hp.n_layer + 1) * hp.n_embd * 4);
}

// quantise activations the way ggml converts src1 for weights of W's type; cached per format until invalidated
void Model::quantize_for(const DTensor &W, const float *x, int64_t x_stride, int n, uint8_t **out, ActFmt *fmt) {
    const ActFmt f = act_format_for(W.type);
    *fmt = f;
    if (f == ACT_F32) {
        B2_CHECK(x_stride == W.cols, B2_EINVAL, "f32 weights need densely packed activations");
        *out = (uint8_t *)x;
        return;
    }
    if (!actq_valid_[f]) {
        launch_quantize_act(f, x, x_stride, n, W.cols, actq_[f], stream);
        actq_valid_[f] = true;
    }
    *out = actq_[f];
}

void Model::matmul(const DTensor &W, const float *x, int64_t x_stride, int n, float *out, int64_t out_stride,
                   const float *residual, int64_t res_stride) {
    uint8_t *xq;
    ActFmt f;
    quantize_for(W, x, x_stride, n, &xq, &f);
    launch_gemv(W, xq, f, n, out, out_stride, residual, res_stride, stream);
}

static inline void invalidate(bool (&v)[4]) { v[0] = v[1] = v[2] = v[3] = false; }
void Model::invalidate_acts() { invalidate(actq_valid_); }

// attention over a batch: consecutive tokens of one sequence (a prompt chunk) go through the flash kernel, single decode
// tokens through the per-token kernel
// every run of the batch is long enough for the flash kernel: it writes attn_output's operand tiles itself
bool Model::attention_to_tiles(const KVLayout &kv, int n, void *x_tiles) {
    static const bool off = getenv("B200LLM_NO_ATT_TILES") && atoi(getenv("B200LLM_NO_ATT_TILES")) != 0;
    if (off || runs_.empty() || !fattn_supported(hp.head_dim) || hp.n_embd % 64 != 0 || hp.n_head * hp.head_dim != hp.n_embd) return false;
    for (const Run &r : runs_)
        if (r.n < 16) return false;
    (void)n;
    for (const Run &r : runs_)
        launch_fattn_prefill(q_, r.t0, r.n, r.pos0, r.slot, hp.n_head, hp.n_head_kv, hp.head_dim, kv, att_, stream, x_tiles);
    return true;
}

void Model::attention(const KVLayout &kv, int n) {
    const int hd = hp.head_dim, d = hp.n_embd;
    if (runs_.empty()) {
        launch_attention(q_, pos_, slot_, n, hp.n_head, hp.n_head_kv, hd, kv, att_, n_ctx_slot, stream);
        return;
    }
    for (const Run &r : runs_) {
        if (r.n >= 16 && fattn_supported(hd))
            launch_fattn_prefill(q_, r.t0, r.n, r.pos0, r.slot, hp.n_head, hp.n_head_kv, hd, kv, att_, stream);
        else
            launch_attention(q_ + (size_t)r.t0 * d, pos_ + r.t0, slot_ + r.t0, r.n, hp.n_head, hp.n_head_kv, hd, kv,
                             att_ + (size_t)r.t0 * d, n_ctx_slot, stream);
    }
}

static FusedJob job_of(const DTensor &W) { return FusedJob{W.data, W.type, W.rows, W.row_stride}; }

// The original, one-op-per-kernel path.  Kept for layers whose matrices need different activation formats (the
// fused kernel quantises the activation row once) and as the cross-check of the fused path in the tests
// (B200LLM_UNFUSED=1).
void Model::run_layer_unfused(const LayerW &L, const KVLayout &kv, int n) {
    const int d = hp.n_embd, ff = hp.n_ff, hd = hp.head_dim, kvd = hp.n_head_kv * hd;
    launch_rmsnorm(x_, d, nullptr, (const float *)L.attn_norm.data, xn_, d, n, d, hp.rms_eps, stream);
    invalidate(actq_valid_);
    matmul(L.wq, xn_, d, n, q_, d, nullptr, 0);
    matmul(L.wk, xn_, d, n, k_, kvd, nullptr, 0);
    matmul(L.wv, xn_, d, n, v_, kvd, nullptr, 0);
    launch_rope_kv_store(q_, k_, v_, pos_, slot_, n, hp.n_head, hp.n_head_kv, hd, rope_table_, kv, stream, (const float *)L.bq.data,
                         (const float *)L.bk.data, (const float *)L.bv.data, hp.rope_neox_kernel);
    attention(kv, n);
    invalidate(actq_valid_);
    matmul(L.wo, att_, d, n, x_, d, x_, d); // x = x + wo.att
    launch_rmsnorm(x_, d, nullptr, (const float *)L.ffn_norm.data, xn_, d, n, d, hp.rms_eps, stream);
    invalidate(actq_valid_);
    matmul(L.up, xn_, d, n, up_, ff, nullptr, 0);
    matmul(L.gate, xn_, d, n, gate_, ff, nullptr, 0);
    launch_swiglu(gate_, up_, gate_, (int64_t)n * ff, stream);
    invalidate(actq_valid_);
    matmul(L.down, gate_, ff, n, x_, d, x_, d); // x = x + down.act
}

// 5 launches per layer: [norm+quant+QKV+RoPE+KV-store] [attention] [quant+wo+residual] [norm+quant+gate/up+SwiGLU]
// [quant+down+residual]
void Model::run_layer_fused(const LayerW &L, const KVLayout &kv, int n) {
    const int d = hp.n_embd, ff = hp.n_ff, hd = hp.head_dim;
    FusedGemv a;
    a.x = x_; a.x_stride = d; a.norm_w = (const float *)L.attn_norm.data; a.eps = hp.rms_eps; a.K = d;
    a.fmt = act_format_for(L.wq.type);
    a.n_jobs = 3; a.job[0] = job_of(L.wq); a.job[1] = job_of(L.wk); a.job[2] = job_of(L.wv);
    a.epi = EPI_QKV; a.out = q_; a.out_stride = d; a.pos = pos_; a.slot = slot_; a.rope_table = rope_table_; a.kv = kv; a.hd = hd;
    a.qkv_bias[0] = (const float *)L.bq.data; a.qkv_bias[1] = (const float *)L.bk.data; a.qkv_bias[2] = (const float *)L.bv.data;
    launch_fused_gemv(a, n, stream);
    attention(kv, n);
    FusedGemv o;
    o.x = att_; o.x_stride = d; o.K = d; o.fmt = act_format_for(L.wo.type);
    o.n_jobs = 1; o.job[0] = job_of(L.wo); o.epi = EPI_PLAIN; o.out = x_; o.out_stride = d; o.residual = x_; o.res_stride = d;
    launch_fused_gemv(o, n, stream);
    FusedGemv g;
    g.x = x_; g.x_stride = d; g.norm_w = (const float *)L.ffn_norm.data; g.eps = hp.rms_eps; g.K = d;
    g.fmt = act_format_for(L.gate.type);
    g.n_jobs = 2; g.job[0] = job_of(L.gate); g.job[1] = job_of(L.up); g.epi = EPI_SWIGLU; g.out = gate_; g.out_stride = ff;
    launch_fused_gemv(g, n, stream);
    FusedGemv w;
    w.x = gate_; w.x_stride = ff; w.K = ff; w.fmt = act_format_for(L.down.type);
    w.n_jobs = 1; w.job[0] = job_of(L.down); w.epi = EPI_PLAIN; w.out = x_; w.out_stride = d; w.residual = x_; w.res_stride = d;
    launch_fused_gemv(w, n, stream);
}

void Model::ensure_gemm_scratch(size_t w_bytes, size_t x_bytes, size_t x2_bytes) {
    bool realloc = false;
    if (x2_bytes > gemm_x2_bytes_) {
        B2_CUDA(cudaStreamSynchronize(stream));
        if (gemm_x2_) cudaFree(gemm_x2_);
        gemm_x2_ = nullptr;
        B2_CUDA(cudaMalloc(&gemm_x2_, x2_bytes));
        gemm_x2_bytes_ = x2_bytes;
        realloc = true;
    }
    if (w_bytes > gemm_w_bytes_) {
        realloc = true;
        B2_CUDA(cudaStreamSynchronize(stream));
        for (auto &kv : w_tiles_) cudaFree(kv.second); // memory is short (no_shadow_): give the resident tiles back too
        w_tiles_.clear();
        for (auto &kv : pair_tiles_) cudaFree(kv.second);
        pair_tiles_.clear();
        if (gemm_w_) cudaFree(gemm_w_);
        gemm_w_ = nullptr;
        B2_CUDA(cudaMalloc(&gemm_w_, w_bytes));
        gemm_w_bytes_ = w_bytes;
    }
    if (x_bytes > gemm_x_bytes_) {
        B2_CUDA(cudaStreamSynchronize(stream));
        if (gemm_x_) cudaFree(gemm_x_);
        gemm_x_ = nullptr;
        B2_CUDA(cudaMalloc(&gemm_x_, x_bytes));
        gemm_x_bytes_ = x_bytes;
        realloc = true;
    }
    if (realloc)
        for (auto &g : decode_graphs_) { // captured graphs hold the old scratch pointers
            if (g) cudaGraphExecDestroy(g);
            g = nullptr;
        }
}

// fp16 operand tiles of W for the tensor-core path.  180 GB of HBM is room to keep them resident (15 GB for an 8B
// model) next to the quantised weights the decode kernels stream: prefill chunks and batched decode steps then skip
// the dequantisation pass.  Built on first use (never inside a graph capture: run_decode_graphed warms up first);
// when memory is short the tiles are rebuilt into a scratch buffer on every call instead.
// Batches above this many tokens take the tensor-core path.  With the fp16 tiles resident it wins from 2 tokens on
// (measured on Llama-3-8B: 4.3 ms per step for 2..16 streams against 5.4 / 6.9 / 8.5 ms of the integer kernels at 2 / 3 / 4);
// when the tiles have to be rebuilt on every call the integer kernels keep the small batches.
int Model::gemm_threshold() const {
    if (tiles_only_) return 0; // Q4_0 / Q4_1 / Q5_0 / Q5_1 / Q2_K / Q3_K / IQ4 weights: the tensor-core path is the only one, from 1 token on
    return gemm_min_tokens_ >= 0 ? gemm_min_tokens_ : (no_shadow_ ? 4 : 1);
}

void *Model::resident_tiles(const DTensor &W) {
    if (no_shadow_) return nullptr;
    auto it = w_tiles_.find(W.data);
    if (it != w_tiles_.end()) return it->second;
    const size_t bytes = gemm_w_tiles_bytes(W.rows, W.cols);
    size_t free_b = 0, total_b = 0;
    cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
    cudaStreamIsCapturing(stream, &cs);
    if (cs != cudaStreamCaptureStatusNone) return nullptr;
    if (cudaMemGetInfo(&free_b, &total_b) != cudaSuccess || free_b < bytes + shadow_reserve_) { no_shadow_ = true; return nullptr; }
    void *p = nullptr;
    if (cudaMalloc(&p, bytes) != cudaSuccess) { cudaGetLastError(); no_shadow_ = true; return nullptr; }
    launch_gemm_dequant_w(W, p, stream);
    w_tiles_[W.data] = p;
    return p;
}

// interleaved ffn_gate | ffn_up operand tiles (the GEMM over them applies SwiGLU in its epilogue), resident like the others
void *Model::resident_pair_tiles(const LayerW &L) {
    if (no_shadow_ || no_fused_swiglu_ || L.gate.rows % 128 != 0 || L.gate.rows != L.up.rows || L.gate.cols != L.up.cols) return nullptr;
    auto it = pair_tiles_.find(L.gate.data);
    if (it != pair_tiles_.end()) return it->second;
    if (w_tiles_.count(L.gate.data)) return nullptr; // the separate tiles of this layer exist already: keep using them
    cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
    cudaStreamIsCapturing(stream, &cs);
    if (cs != cudaStreamCaptureStatusNone) return nullptr;
    const size_t bytes = gemm_w_pair_tiles_bytes(L.gate.rows, L.gate.cols);
    size_t free_b = 0, total_b = 0;
    if (cudaMemGetInfo(&free_b, &total_b) != cudaSuccess || free_b < bytes + shadow_reserve_) return nullptr;
    void *p = nullptr;
    if (cudaMalloc(&p, bytes) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    launch_gemm_dequant_w_pair(L.gate, L.up, p, stream);
    pair_tiles_[L.gate.data] = p;
    return p;
}

void Model::gemm(const DTensor &W, int n, float *out, int64_t out_stride, const float *residual, int64_t res_stride, const void *x_tiles) {
    if (!x_tiles) x_tiles = gemm_x_;
    if (!gemm_part_) { // split-K partial tiles for batched decode (<= 128 tokens): 148 slices x 128 tokens x 256 rows at most
        gemm_part_bytes_ = (size_t)32 << 20;
        gemm_part_ = (float *)dalloc(gemm_part_bytes_);
    }
    if (void *t = resident_tiles(W)) {
        launch_gemm_tc(W, x_tiles, n, out, out_stride, residual, res_stride, t, true, gemm_part_, gemm_part_bytes_, stream);
        return;
    }
    ensure_gemm_scratch(gemm_w_tiles_bytes(W.rows, W.cols), 0);
    launch_gemm_tc(W, x_tiles, n, out, out_stride, residual, res_stride, gemm_w_, false, gemm_part_, gemm_part_bytes_, stream);
}

// Several matrices that share the activation tiles in ONE launch (wq|wk|wv, gate|up) when their fp16 tiles are
// resident; otherwise one launch each through the scratch tile buffer.
void Model::gemm_multi(const DTensor *const *Ws, float *const *outs, const int64_t *strides, int n_jobs, int n) {
    if (!gemm_part_) {
        gemm_part_bytes_ = (size_t)32 << 20;
        gemm_part_ = (float *)dalloc(gemm_part_bytes_);
    }
    GemmJob jobs[3];
    bool all = n_jobs <= 3;
    for (int j = 0; j < n_jobs && all; j++) {
        void *t = resident_tiles(*Ws[j]);
        if (!t) all = false;
        else jobs[j] = GemmJob{Ws[j], t, true, outs[j], strides[j]};
    }
    if (all) {
        launch_gemm_tc_multi(jobs, n_jobs, gemm_x_, n, nullptr, 0, gemm_part_, gemm_part_bytes_, stream);
        return;
    }
    for (int j = 0; j < n_jobs; j++) gemm(*Ws[j], n, outs[j], strides[j], nullptr, 0);
}

// Prefill / batched decode: every mat-mul is a tcgen05 tensor-core GEMM on fp16 operands (weights dequantised into
// operand tiles on the fly); attention, norms and RoPE as in the other paths.
void Model::run_layer_gemm(const LayerW &L, const KVLayout &kv, int n) {
    const int d = hp.n_embd, ff = hp.n_ff, hd = hp.head_dim, kvd = hp.n_head_kv * hd;
    ensure_gemm_scratch(0, gemm_x_tiles_bytes(n, std::max(d, ff)));
    launch_gemm_prepare_x_rmsnorm(x_, d, (const float *)L.attn_norm.data, n, d, hp.rms_eps, gemm_x_, nullptr, stream);
    {
        const DTensor *Ws[3] = {&L.wq, &L.wk, &L.wv};
        float *outs[3] = {q_, k_, v_};
        const int64_t strides[3] = {d, kvd, kvd};
        gemm_multi(Ws, outs, strides, 3, n);
    }
    launch_rope_kv_store(q_, k_, v_, pos_, slot_, n, hp.n_head, hp.n_head_kv, hd, rope_table_, kv, stream, (const float *)L.bq.data,
                         (const float *)L.bk.data, (const float *)L.bv.data, hp.rope_neox_kernel);
    if (!attention_to_tiles(kv, n, gemm_x_)) {
        attention(kv, n);
        launch_gemm_prepare_x(att_, d, n, d, gemm_x_, stream);
    }
    gemm(L.wo, n, x_, d, x_, d);
    launch_gemm_prepare_x_rmsnorm(x_, d, (const float *)L.ffn_norm.data, n, d, hp.rms_eps, gemm_x_, nullptr, stream);
    if (void *pt = resident_pair_tiles(L)) { // gate | up interleaved: SwiGLU in the GEMM's epilogue, which writes ffn_down's operand tiles
        ensure_gemm_scratch(0, 0, gemm_x_tiles_bytes(n, ff));
        launch_gemm_tc_swiglu(ff, d, pt, gemm_x_, n, gemm_x2_, stream);
        gemm(L.down, n, x_, d, x_, d, gemm_x2_);
        return;
    }
    {
        const DTensor *Ws[2] = {&L.gate, &L.up};
        float *outs[2] = {gate_, up_};
        const int64_t strides[2] = {ff, ff};
        gemm_multi(Ws, outs, strides, 2, n);
    }
    launch_gemm_prepare_x_swiglu(gate_, up_, n, ff, gemm_x_, stream);
    gemm(L.down, n, x_, d, x_, d);
}

// ---- batched decode, 2..64 tokens: the weights stay quantised all the way into the tensor-core kernel (kernels_qgemm.cu)
void *Model::resident_qtiles(const DTensor &W) {
    auto it = q_tiles_.find(W.data);
    if (it != q_tiles_.end()) return it->second;
    if (!qgemm_type_supported(W.type) || W.cols % 256 != 0) return nullptr;
    cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
    cudaStreamIsCapturing(stream, &cs);
    if (cs != cudaStreamCaptureStatusNone) return nullptr;
    const size_t bytes = qgemm_w_bytes(W.type, W.rows, W.cols);
    size_t free_b = 0, total_b = 0;
    if (cudaMemGetInfo(&free_b, &total_b) != cudaSuccess || free_b < bytes + shadow_reserve_) return nullptr;
    void *p = nullptr;
    if (cudaMalloc(&p, bytes) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    launch_qgemm_repack(W, p, stream);
    q_tiles_[W.data] = p;
    return p;
}

bool Model::qgemm_layer_ready(const LayerW &L) {
    const DTensor *ws[7] = {&L.wq, &L.wk, &L.wv, &L.wo, &L.gate, &L.up, &L.down};
    for (const DTensor *w : ws)
        if (!resident_qtiles(*w)) return false;
    return true;
}

bool Model::qgemm_multi(const DTensor *const *Ws, float *const *outs, const int64_t *strides, int n_jobs, int n, const float *residual,
                        int64_t res_stride) {
    if (n_sm_ <= 0) B2_CUDA(cudaDeviceGetAttribute(&n_sm_, cudaDevAttrMultiProcessorCount, device));
    const size_t part_bytes = qgemm_partial_bytes(n_sm_);
    if (!qg_part_) {
        qg_part_ = (float *)dalloc(part_bytes);
        qg_counters_ = (unsigned *)dalloc(QGEMM_COUNTERS * sizeof(unsigned));
        B2_CUDA(cudaMemsetAsync(qg_counters_, 0, QGEMM_COUNTERS * sizeof(unsigned), stream));
    }
    QGemmJob jobs[3];
    for (int j = 0; j < n_jobs; j++) {
        void *t = resident_qtiles(*Ws[j]);
        if (!t) return false;
        jobs[j] = QGemmJob{Ws[j], t, outs[j], strides[j]};
    }
    launch_qgemm_multi(jobs, n_jobs, gemm_x_, n, residual, res_stride, qg_part_, part_bytes, qg_counters_, QGEMM_COUNTERS, n_sm_, stream);
    return true;
}

void Model::run_layer_qgemm(const LayerW &L, const KVLayout &kv, int n) {
    const int d = hp.n_embd, ff = hp.n_ff, hd = hp.head_dim, kvd = hp.n_head_kv * hd;
    ensure_gemm_scratch(0, gemm_x_tiles_bytes(n, std::max(d, ff)));
    launch_qgemm_prepare_x_rmsnorm(x_, d, (const float *)L.attn_norm.data, n, d, hp.rms_eps, gemm_x_, stream);
    {
        const DTensor *Ws[3] = {&L.wq, &L.wk, &L.wv};
        float *outs[3] = {q_, k_, v_};
        const int64_t strides[3] = {d, kvd, kvd};
        B2_CHECK(qgemm_multi(Ws, outs, strides, 3, n, nullptr, 0), B2_ESTATE, "qgemm: tiles missing");
    }
    // tokens in different slots (a decode step over many streams): RoPE, KV store, attention and the conversion of its
    // output into attn_output's activation tiles are one launch; otherwise (several tokens of one slot) three
    if (!(slots_distinct_ && !no_fused_att_ &&
          launch_attention_decode_fused(q_, k_, v_, pos_, slot_, n, hp.n_head, hp.n_head_kv, hd, rope_table_, kv, n_ctx_slot, gemm_x_,
                                        (int)align_up(n, 16), stream))) {
        launch_rope_kv_store(q_, k_, v_, pos_, slot_, n, hp.n_head, hp.n_head_kv, hd, rope_table_, kv, stream, (const float *)L.bq.data,
                         (const float *)L.bk.data, (const float *)L.bv.data, hp.rope_neox_kernel);
        attention(kv, n);
        launch_qgemm_prepare_x(att_, d, n, d, gemm_x_, stream);
    }
    {
        const DTensor *Ws[1] = {&L.wo};
        float *outs[1] = {x_};
        const int64_t strides[1] = {d};
        B2_CHECK(qgemm_multi(Ws, outs, strides, 1, n, x_, d), B2_ESTATE, "qgemm: tiles missing");
    }
    launch_qgemm_prepare_x_rmsnorm(x_, d, (const float *)L.ffn_norm.data, n, d, hp.rms_eps, gemm_x_, stream);
    {
        const DTensor *Ws[2] = {&L.gate, &L.up};
        float *outs[2] = {gate_, up_};
        const int64_t strides[2] = {ff, ff};
        B2_CHECK(qgemm_multi(Ws, outs, strides, 2, n, nullptr, 0), B2_ESTATE, "qgemm: tiles missing");
    }
    launch_qgemm_prepare_x_swiglu(gate_, up_, n, ff, gemm_x_, stream);
    {
        const DTensor *Ws[1] = {&L.down};
        float *outs[1] = {x_};
        const int64_t strides[1] = {d};
        B2_CHECK(qgemm_multi(Ws, outs, strides, 1, n, x_, d), B2_ESTATE, "qgemm: tiles missing");
    }
}

static bool layer_fusable(const LayerW &L) {
    const ActFmt f = act_format_for(L.wq.type);
    return f != ACT_F32 && act_format_for(L.wk.type) == f && act_format_for(L.wv.type) == f && L.gate.type == L.up.type &&
           act_format_for(L.gate.type) != ACT_F32 && act_format_for(L.wo.type) != ACT_F32 && act_format_for(L.down.type) != ACT_F32 &&
           L.wq.rows % 2 == 0 && L.wk.rows % 2 == 0 && L.wv.rows % 2 == 0;
}

void Model::run_graph(int n, const int32_t *out_rows_dev, int n_out) {
    const int d = hp.n_embd, hd = hp.head_dim, kvd = hp.n_head_kv * hd;
    launch_dequant_rows(tok_embd, tok_, 0, n, x_, d, stream); // get_rows(tok_embd)
    for (int l = 0; l < hp.n_layer; l++) {
        const LayerW &L = layers[l];
        KVLayout kv{k_pool_ + (size_t)l * pool_layer_stride_, v_pool_ + (size_t)l * pool_layer_stride_, page_table_, page_size,
                    max_pages, kvd};
        if (!no_gemm_ && !no_qgemm_ && n >= 2 && n <= QGEMM_MAX_TOKENS && qgemm_layer_ready(L)) run_layer_qgemm(L, kv, n);
        else if (!no_gemm_ && n > gemm_threshold()) run_layer_gemm(L, kv, n);
        else if (!unfused_ && n <= 4 && layer_fusable(L)) run_layer_fused(L, kv, n); // larger batches: quantise once, not per CTA
        else run_layer_unfused(L, kv, n);
        if (dbg_hidden_)
            B2_CUDA(cudaMemcpyAsync(dbg_hidden_ + (size_t)l * d, x_ + (size_t)(n - 1) * d, d * 4, cudaMemcpyDeviceToDevice, stream));
    }
    if (n_out != 0) {
        const int rows = n_out < 0 ? n : n_out;
        ensure_logits_rows(rows);
        if (!no_gemm_ && !no_qgemm_ && rows >= 2 && rows <= QGEMM_MAX_TOKENS && resident_qtiles(output)) {
            launch_rmsnorm(x_, d, n_out < 0 ? nullptr : out_rows_dev, (const float *)output_norm.data, xn_out_, d, rows, d, hp.rms_eps,
                           stream);
            ensure_gemm_scratch(0, gemm_x_tiles_bytes(rows, d));
            launch_qgemm_prepare_x(xn_out_, d, rows, d, gemm_x_, stream);
            const DTensor *Ws[1] = {&output};
            float *outs[1] = {logits_};
            const int64_t strides[1] = {hp.n_vocab};
            B2_CHECK(qgemm_multi(Ws, outs, strides, 1, rows, nullptr, 0), B2_ESTATE, "qgemm: tiles missing");
        } else if (!no_gemm_ && rows > gemm_threshold()) {
            launch_rmsnorm(x_, d, n_out < 0 ? nullptr : out_rows_dev, (const float *)output_norm.data, xn_out_, d, rows, d, hp.rms_eps,
                           stream);
            ensure_gemm_scratch(0, gemm_x_tiles_bytes(rows, d));
            launch_gemm_prepare_x(xn_out_, d, rows, d, gemm_x_, stream);
            gemm(output, rows, logits_, hp.n_vocab, nullptr, 0);
        } else if (!unfused_ && rows <= 4 && act_format_for(output.type) != ACT_F32) {
            FusedGemv a;
            a.x = x_; a.x_stride = d; a.x_rows = n_out < 0 ? nullptr : out_rows_dev;
            a.norm_w = (const float *)output_norm.data; a.eps = hp.rms_eps; a.xn_out = xn_out_; a.K = d;
            a.fmt = act_format_for(output.type);
            a.n_jobs = 1; a.job[0] = job_of(output); a.epi = EPI_PLAIN; a.out = logits_; a.out_stride = hp.n_vocab;
            launch_fused_gemv(a, rows, stream);
        } else {
            launch_rmsnorm(x_, d, n_out < 0 ? nullptr : out_rows_dev, (const float *)output_norm.data, xn_out_, d, rows, d, hp.rms_eps,
                           stream);
            invalidate(actq_valid_);
            matmul(output, xn_out_, d, rows, logits_, hp.n_vocab, nullptr, 0);
        }
        if (dbg_hidden_)
            B2_CUDA(cudaMemcpyAsync(dbg_hidden_ + (size_t)hp.n_layer * d, xn_out_ + (size_t)(rows - 1) * d, d * 4,
                                    cudaMemcpyDeviceToDevice, stream));
    }
}

// Build the phase list of the persistent decode kernel (once): 5 phases per layer + the LM head.
void Model::build_mega() {
    mega_ready_ = false;
    if (no_mega_ || unfused_) return;
    const int d = hp.n_embd, ff = hp.n_ff, hd = hp.head_dim, kvd = hp.n_head_kv * hp.head_dim;
    cudaDeviceGetAttribute(&n_sm_, cudaDevAttrMultiProcessorCount, device);
    if (!x_ll_) {
        // attention of one head is shared by up to 4 CTAs (B200LLM_ATT_PARTS overrides; 1 = one CTA per head)
        mega_att_parts_ = 1;
        while (mega_att_parts_ < 4 && hp.n_head * mega_att_parts_ * 2 <= n_sm_) mega_att_parts_ *= 2;
        if (getenv("B200LLM_ATT_PARTS")) mega_att_parts_ = std::min(4, std::max(1, atoi(getenv("B200LLM_ATT_PARTS"))));
        while (mega_att_parts_ > 1 && (hp.n_head * mega_att_parts_ > n_sm_ || (mega_att_parts_ & (mega_att_parts_ - 1)))) mega_att_parts_--;
        // The tags only ever grow and are compared for equality modulo 2^32; B200LLM_MEGA_COUNTER0 starts them at an
        // arbitrary value so that a test can drive them across the wrap-around.  Every word starts with the tag the
        // first launch can not expect (its base).
        const unsigned c0 = getenv("B200LLM_MEGA_COUNTER0") ? (unsigned)strtoul(getenv("B200LLM_MEGA_COUNTER0"), nullptr, 0) : 0u;
        mega_tag_base_ = c0;
        const size_t kmax = (size_t)std::max(d, ff);
        const size_t n_xq = kmax / 4 + kmax / 256 + kmax / 32 + 8; // words of the quantised activation row image (+ slack for the paired loads)
        const size_t sizes[9] = {(size_t)d, (size_t)d, (size_t)d, (size_t)ff, (size_t)kvd, (size_t)kvd, n_xq,
                                 (size_t)hp.n_head * n_ctx_slot, (size_t)hp.n_head * mega_att_parts_ * hd};
        uint2 **bufs[9] = {&x_ll_, &q_ll_, &att_ll_, &gate_ll_, &kcur_ll_, &vcur_ll_, &xq_ll_, &mega_att_scores_, &mega_att_part_};
        size_t most = 0;
        for (size_t n : sizes) most = std::max(most, n);
        std::vector<uint2> init(most, make_uint2(0u, c0));
        for (int i = 0; i < 9; i++) {
            *bufs[i] = (uint2 *)dalloc(sizes[i] * sizeof(uint2));
            B2_CUDA(cudaMemcpyAsync(*bufs[i], init.data(), sizes[i] * sizeof(uint2), cudaMemcpyHostToDevice, stream));
        }
        mega_hint_ = (unsigned *)dalloc(256);
        mega_hint_base_ = c0;
        {
            std::vector<unsigned> h0(64, c0);
            B2_CUDA(cudaMemcpyAsync(mega_hint_, h0.data(), 256, cudaMemcpyHostToDevice, stream));
        }
        B2_CUDA(cudaStreamSynchronize(stream));
    }
    // ffn_down's input (n_ff >= 8192) is quantised once per GPU, not once per CTA (B200LLM_NO_DIST_QUANT=1: per CTA)
    const bool dist_quant = !(getenv("B200LLM_NO_DIST_QUANT") && atoi(getenv("B200LLM_NO_DIST_QUANT")));
    std::vector<MegaPhase> ph;
    int prev = -1; // the phase that produced x_ll_ (-1: the embedding row, plain fp32 in x_)
    for (int l = 0; l < hp.n_layer; l++) {
        const LayerW &L = layers[l];
        __half *kp = k_pool_ + (size_t)l * pool_layer_stride_, *vp = v_pool_ + (size_t)l * pool_layer_stride_;
        if (act_format_for(L.wq.type) != act_format_for(L.wk.type) || act_format_for(L.wq.type) != act_format_for(L.wv.type) ||
            L.gate.type != L.up.type)
            return;
        MegaPhase q;
        q.kind = PH_QKV; q.K = d; q.fmt = act_format_for(L.wq.type); q.n_jobs = 3;
        q.job[0] = job_of(L.wq); q.job[1] = job_of(L.wk); q.job[2] = job_of(L.wv);
        q.x = x_; q.x_ll = x_ll_; q.src_p = prev;
        q.norm_w = (const float *)L.attn_norm.data; q.out_ll = q_ll_; q.k_pool = kp; q.v_pool = vp;
        q.units = (int)((L.wq.rows + L.wk.rows + L.wv.rows) / 2);
        const int iq = (int)ph.size();
        ph.push_back(q);
        MegaPhase at;
        at.kind = PH_ATT; at.x_ll = q_ll_; at.src_p = iq; at.out_ll = att_ll_; at.k_pool = kp; at.v_pool = vp;
        const int iat = (int)ph.size();
        ph.push_back(at);
        MegaPhase o;
        o.kind = PH_PLAIN; o.K = d; o.fmt = act_format_for(L.wo.type); o.n_jobs = 1; o.job[0] = job_of(L.wo);
        o.x_ll = att_ll_; o.src_p = iat; o.out_ll = x_ll_; o.units = (int)(L.wo.rows / 2);
        if (prev < 0) o.residual = x_; else o.res_ll = x_ll_;
        const int io = (int)ph.size();
        ph.push_back(o);
        MegaPhase g;
        g.kind = PH_SWIGLU; g.K = d; g.fmt = act_format_for(L.gate.type); g.n_jobs = 2; g.job[0] = job_of(L.gate); g.job[1] = job_of(L.up);
        g.x_ll = x_ll_; g.src_p = io; g.norm_w = (const float *)L.ffn_norm.data; g.out_ll = gate_ll_; g.units = (int)L.gate.rows;
        const int ig = (int)ph.size();
        ph.push_back(g);
        MegaPhase w;
        w.kind = PH_PLAIN; w.K = ff; w.fmt = act_format_for(L.down.type); w.n_jobs = 1; w.job[0] = job_of(L.down);
        w.x_ll = gate_ll_; w.src_p = ig; w.res_ll = x_ll_; w.out_ll = x_ll_; w.units = (int)(L.down.rows / 2);
        if (dist_quant && w.fmt == ACT_Q8_K && w.K / 256 > 32 && w.K <= 32768) w.xq_ll = xq_ll_;
        prev = (int)ph.size();
        ph.push_back(w);
    }
    MegaPhase h;
    h.kind = PH_PLAIN; h.K = d; h.fmt = act_format_for(output.type); h.n_jobs = 1; h.job[0] = job_of(output);
    h.x = x_; h.x_ll = x_ll_; h.src_p = prev;
    h.norm_w = (const float *)output_norm.data; h.out = logits_; h.xn_out = xn_out_; h.units = (int)(output.rows / 2);
    ph.push_back(h);
    if (!mega_streams_.data) mega_geo_ = mega_geometry(ph.data(), (int)ph.size(), n_ctx_slot, hd); // (the streams are cut for this geometry)
    for (auto &p : ph)
        if (!mega_phase_prepare(p, n_sm_, mega_geo_)) return;
    if ((size_t)n_ctx_slot > mega_attention_ctx_limit(hd, mega_geo_) || hd % 64 != 0 || hd > 128 || page_size < 16 || (page_size & (page_size - 1)) ||
        (uint64_t)pool_layer_stride_ >= (1ull << 31) || kvd % 2 != 0)
        return;
    if (!mega_streams_.data) {
        // The decode kernel reads its own copy of the weights (MegaStreams): same bytes, laid out per (CTA, warp) in
        // consumption order.  180 GB of HBM is room for it next to the GGUF-layout rows the other paths read; when it
        // is not (B200LLM_NO_MEGA_STREAMS=1 forces that), batch-1 decode takes the fused per-layer kernels instead.
        size_t free_b = 0, total_b = 0;
        const bool off = getenv("B200LLM_NO_MEGA_STREAMS") && atoi(getenv("B200LLM_NO_MEGA_STREAMS"));
        if (off || cudaMemGetInfo(&free_b, &total_b) != cudaSuccess || free_b < (size_t)decode_stream_bytes + shadow_reserve_) return;
        mega_streams_.build(ph.data(), (int)ph.size(), n_sm_, stream, mega_geo_);
    }
    if (!mega_phases_) mega_phases_ = (MegaPhase *)dalloc(ph.size() * sizeof(MegaPhase));
    B2_CUDA(cudaMemcpyAsync(mega_phases_, ph.data(), ph.size() * sizeof(MegaPhase), cudaMemcpyHostToDevice, stream));
    B2_CUDA(cudaStreamSynchronize(stream));
    mega_n_phases_ = (int)ph.size();
    int coop = 0;
    cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, device);
    mega_ready_ = coop != 0;
    if (getenv("B200LLM_MEGA_DBG") && !mega_dbg_) mega_dbg_ = (long long *)dalloc((size_t)n_sm_ * mega_n_phases_ * 16 * sizeof(long long));
}

// Batch-1 decode step through the persistent kernel: [embedding gather] [mega] = 2 launches (+ sampling).
void Model::run_decode_mega() {
    const int d = hp.n_embd;
    launch_dequant_rows(tok_embd, tok_, 0, 1, x_, d, stream);
    MegaArgs a;
    a.phases = mega_phases_; a.n_phases = mega_n_phases_; a.starts = mega_streams_.starts;
    a.pos = pos_; a.slot = slot_; a.rope_table = rope_table_; a.page_table = page_table_;
    a.page_size = page_size; a.max_pages = max_pages; a.kvd = hp.n_head_kv * hp.head_dim;
    a.n_head = hp.n_head; a.n_head_kv = hp.n_head_kv; a.hd = hp.head_dim;
    a.eps = hp.rms_eps; a.att_scale = 1.0f / sqrtf((float)hp.head_dim);
    a.page_shift = 0;
    while ((1 << a.page_shift) < page_size) a.page_shift++;
    a.att_parts = mega_att_parts_;
    a.att_parts_log = 0;
    while ((1 << a.att_parts_log) < a.att_parts) a.att_parts_log++;
    a.att_scores = mega_att_scores_; a.att_scores_stride = n_ctx_slot;
    a.att_part_out = mega_att_part_;
    a.kcur_ll = kcur_ll_; a.vcur_ll = vcur_ll_;
    a.tag_base = mega_tag_base_;
    mega_tag_base_ += (unsigned)mega_n_phases_;
    a.geo = mega_geo_;
    mega_fill_knobs(a);
    a.hint = mega_hint_;
    a.hint_base = mega_hint_base_;
    if (a.ll_probe == 2) mega_hint_base_ += (unsigned)n_sm_ * (unsigned)(mega_n_phases_ - 1);
    a.dbg = mega_dbg_;
    const bool timed = mega_timing_ && mega_ev_n_ + 2 <= mega_ev_.size();
    if (timed) B2_CUDA(cudaEventRecord(mega_ev_[mega_ev_n_++], stream));
    launch_decode_mega(a, n_sm_, stream);
    if (timed) B2_CUDA(cudaEventRecord(mega_ev_[mega_ev_n_++], stream));
}

void Model::mega_timing_begin(int max_launches) {
    while ((int)mega_ev_.size() < 2 * max_launches) {
        cudaEvent_t e;
        B2_CUDA(cudaEventCreate(&e));
        mega_ev_.push_back(e);
    }
    mega_ev_n_ = 0;
    mega_timing_ = true;
}

double Model::mega_timing_end() {
    mega_timing_ = false;
    B2_CUDA(cudaStreamSynchronize(stream));
    double tot = 0;
    for (size_t i = 0; i + 1 < mega_ev_n_; i += 2) {
        float ms = 0;
        B2_CUDA(cudaEventElapsedTime(&ms, mega_ev_[i], mega_ev_[i + 1]));
        tot += ms;
    }
    const double mean = mega_ev_n_ ? tot / (double)(mega_ev_n_ / 2) : 0.0;
    mega_ev_n_ = 0;
    return mean;
}

// Decode step for `n` sequences whose inputs are in the device buffers, replayed from a captured CUDA graph
// (one graph per batch size; ~165 kernel launches collapse into one cudaGraphLaunch).
void Model::run_decode_graphed(int n) {
    if (n == 1 && mega_ready_ && !dbg_hidden_) {
        run_decode_mega();
        return;
    }
    if (no_graph_ || dbg_hidden_ || n > (int)decode_graphs_.size() - 1) {
        run_graph(n, nullptr, -1);
        return;
    }
    if (!decode_graphs_[n]) {
        ensure_logits_rows(std::max(n, n_parallel));
        run_graph(n, nullptr, -1); // warm-up outside capture: sets function attributes, validates shapes
        B2_CUDA(cudaStreamSynchronize(stream));
        cudaGraph_t g = nullptr;
        const int64_t lc = g_launch_count;
        B2_CUDA(cudaStreamBeginCapture(stream, cudaStreamCaptureModeThreadLocal));
        try {
            run_graph(n, nullptr, -1);
        } catch (...) {
            cudaStreamEndCapture(stream, &g);
            if (g) cudaGraphDestroy(g);
            throw;
        }
        B2_CUDA(cudaStreamEndCapture(stream, &g));
        decode_graph_launches_[n] = g_launch_count - lc;
        g_launch_count = lc;
        cudaGraphExec_t ge = nullptr;
        B2_CUDA(cudaGraphInstantiate(&ge, g, 0));
        cudaGraphDestroy(g);
        decode_graphs_[n] = ge;
    }
    B2_CUDA(cudaGraphLaunch(decode_graphs_[n], stream));
    g_launch_count += decode_graph_launches_[n];
}

void Model::forward(const Batch &b) {
    B2_CHECK(b.n > 0 && b.n <= n_batch, B2_EINVAL, "batch size must be in [1, n_batch]");
    B2_CUDA(cudaSetDevice(device));
    for (int i = 0; i < b.n; i++) {
        B2_CHECK(b.slot[i] >= 0 && b.slot[i] < n_parallel, B2_EINVAL, "slot out of range");
        B2_CHECK(b.pos[i] >= 0 && b.pos[i] < n_ctx_slot, B2_EINVAL, "position exceeds the slot's context");
        if (b.tokens) B2_CHECK(b.tokens[i] >= 0 && b.tokens[i] < hp.n_vocab, B2_EINVAL, "token id out of range");
    }
    if (tp_size > 1) {
        forward_tp(b);
        return;
    }
    { // one host->device copy for the step's inputs (three rows of n ints into the three device arrays, n_batch apart)
        const size_t n = (size_t)b.n;
        const int rows0 = b.tokens ? 0 : 1; // without tokens (embeddings fed on the device) only positions and slots go up
        if (b.tokens) memcpy(in_host_.data(), b.tokens, n * 4);
        memcpy(in_host_.data() + n, b.pos, n * 4);
        memcpy(in_host_.data() + 2 * n, b.slot, n * 4);
        B2_CUDA(cudaMemcpy2DAsync(tok_ + (size_t)rows0 * n_batch, (size_t)n_batch * 4, in_host_.data() + rows0 * n, n * 4, n * 4, 3 - rows0,
                                  cudaMemcpyHostToDevice, stream));
    }
    runs_.clear();
    for (int i = 0; i < b.n;) { // maximal runs of consecutive positions of one slot
        int j = i + 1;
        while (j < b.n && b.slot[j] == b.slot[i] && b.pos[j] == b.pos[i] + (j - i)) j++;
        runs_.push_back(Run{i, j - i, b.pos[i], b.slot[i]});
        i = j;
    }
    int n_out = b.n_out;
    bool pure_decode = n_out == b.n && b.n < (int)decode_graphs_.size(); // every row wants logits, in order
    for (int i = 0; pure_decode && i < n_out; i++) pure_decode = b.out_rows[i] == i;
    // every token in its own slot?  (then no token of the batch attends to another one's K/V row, and the captured decode
    // graphs -- which hold the fused RoPE + attention kernel -- apply)
    slots_distinct_ = true;
    for (int i = 1; slots_distinct_ && i < b.n; i++)
        for (int j = 0; j < i; j++)
            if (b.slot[j] == b.slot[i]) { slots_distinct_ = false; break; }
    if (pure_decode && !slots_distinct_ && b.n > 1) pure_decode = false; // several tokens of one slot: the un-captured path, all rows
    if (pure_decode) {
        runs_.clear(); // one per-token attention launch for the whole batch (and what the captured graph holds)
        run_decode_graphed(b.n);
        return;
    }
    if (n_out > 0) B2_CUDA(cudaMemcpyAsync(out_rows_, b.out_rows, (size_t)n_out * 4, cudaMemcpyHostToDevice, stream));
    run_graph(b.n, out_rows_, n_out);
}

void Model::set_device_inputs(const int32_t *tokens, const int32_t *pos, const int32_t *slot, int n) {
    B2_CUDA(cudaSetDevice(device)); // (a tensor-parallel model: shard 0's buffers; forward_device_inputs hands them to the other shards)
    B2_CUDA(cudaMemcpyAsync(tok_, tokens, (size_t)n * 4, cudaMemcpyHostToDevice, stream));
    B2_CUDA(cudaMemcpyAsync(pos_, pos, (size_t)n * 4, cudaMemcpyHostToDevice, stream));
    B2_CUDA(cudaMemcpyAsync(slot_, slot, (size_t)n * 4, cudaMemcpyHostToDevice, stream));
}

void Model::forward_device_inputs(int n) { // (device-resident decode loops: one token per slot by construction)
    B2_CUDA(cudaSetDevice(device));
    slots_distinct_ = true;
    if (tp_size > 1) {
        forward_device_inputs_tp(n);
        return;
    }
    runs_.clear();
    run_decode_graphed(n);
}

void Model::sample_greedy(int n_rows, bool feed) {
    launch_argmax(logits_, hp.n_vocab, n_rows, hp.n_vocab, next_tok_, feed ? tok_ : nullptr, feed ? pos_ : nullptr, stream);
}

} // namespace b2

namespace b2 {
// time the GEMV kernel alone, cycling over the layers so the weights of a launch are never L2-resident
int Model::bench_gemv(int which, int n_tok, int n_rounds, double *ms_per_launch, int64_t *bytes_per_launch, int *ggml_type) {
    B2_CUDA(cudaSetDevice(device));
    std::vector<const DTensor *> ws;
    for (auto &L : layers) {
        const DTensor *t = which == 0 ? &L.wq : which == 1 ? &L.wk : which == 2 ? &L.wv : which == 3 ? &L.wo
                         : which == 4 ? &L.gate : which == 5 ? &L.up : which == 6 ? &L.down : nullptr;
        if (t) ws.push_back(t);
    }
    if (which == 7) ws.push_back(&output);
    if (which >= 8) ws.push_back(which == 8 ? &layers[0].wq : &layers[0].gate); // 8 = wq|wk|wv, 9 = gate|up as one launch (batched decode)
    B2_CHECK(!ws.empty(), B2_EINVAL, "bench_gemv: no such tensor");
    const DTensor &W0 = *ws[0];
    const int64_t K = W0.cols;
    float *xin = which == 6 ? gate_ : xn_;
    float *out = which == 7 ? logits_ : (which >= 4 && which <= 5 ? up_ : q_);
    if (which == 7) ensure_logits_rows(n_tok);
    // deterministic, O(1)-magnitude activations
    std::vector<float> hx((size_t)n_tok * K);
    for (size_t i = 0; i < hx.size(); i++) hx[i] = sinf(0.37f * (float)i) * 1.5f;
    B2_CUDA(cudaMemcpyAsync(xin, hx.data(), hx.size() * 4, cudaMemcpyHostToDevice, stream));
    if (n_tok > 4 || which >= 8) { // batched decode: the GEMM that dequantises inside the kernel, one launch per layer
        std::vector<std::vector<const DTensor *>> jobs;
        if (which >= 8) {
            for (auto &L : layers) jobs.push_back(which == 8 ? std::vector<const DTensor *>{&L.wq, &L.wk, &L.wv} : std::vector<const DTensor *>{&L.gate, &L.up});
        } else {
            for (auto *w : ws) jobs.push_back({w});
        }
        ensure_gemm_scratch(0, gemm_x_tiles_bytes(n_tok, K));
        launch_qgemm_prepare_x(xin, K, n_tok, K, gemm_x_, stream);
        float *outs[3] = {which == 7 ? logits_ : q_, k_, v_};
        if (which == 9 || which == 4 || which == 5) { outs[0] = gate_; outs[1] = up_; }
        int64_t bytes = 0;
        auto run = [&](bool count) {
            for (auto &jb : jobs) {
                int64_t strides[3];
                for (size_t j = 0; j < jb.size(); j++) {
                    strides[j] = jb[j]->rows;
                    if (count) bytes += jb[j]->rows * gguf_row_bytes(jb[j]->type, jb[j]->cols);
                }
                B2_CHECK(qgemm_multi(jb.data(), outs, strides, (int)jb.size(), n_tok, nullptr, 0), B2_EUNSUPPORTED, "bench_gemv: no qgemm tiles for this tensor");
            }
        };
        run(true); // warm-up round (builds the tile-major copies)
        cudaEvent_t e0, e1;
        B2_CUDA(cudaEventCreate(&e0));
        B2_CUDA(cudaEventCreate(&e1));
        B2_CUDA(cudaEventRecord(e0, stream));
        for (int r = 0; r < n_rounds; r++) run(false);
        B2_CUDA(cudaEventRecord(e1, stream));
        B2_CUDA(cudaStreamSynchronize(stream));
        float ms = 0;
        B2_CUDA(cudaEventElapsedTime(&ms, e0, e1));
        cudaEventDestroy(e0);
        cudaEventDestroy(e1);
        if (ms_per_launch) *ms_per_launch = ms / ((double)n_rounds * jobs.size());
        if (bytes_per_launch) *bytes_per_launch = bytes / (int64_t)jobs.size();
        if (ggml_type) *ggml_type = W0.type;
        return 0;
    }
    uint8_t *xq;
    ActFmt f;
    invalidate_acts();
    quantize_for(W0, xin, K, n_tok, &xq, &f);
    for (auto *w : ws) B2_CHECK(w->type == W0.type, B2_EUNSUPPORTED, "bench_gemv: layers use different types for this tensor; pick another");
    cudaEvent_t e0, e1;
    B2_CUDA(cudaEventCreate(&e0));
    B2_CUDA(cudaEventCreate(&e1));
    for (auto *w : ws) launch_gemv(*w, xq, f, n_tok, out, w->rows, nullptr, 0, stream); // warm-up round
    B2_CUDA(cudaEventRecord(e0, stream));
    for (int r = 0; r < n_rounds; r++)
        for (auto *w : ws) launch_gemv(*w, xq, f, n_tok, out, w->rows, nullptr, 0, stream);
    B2_CUDA(cudaEventRecord(e1, stream));
    B2_CUDA(cudaStreamSynchronize(stream));
    float ms = 0;
    B2_CUDA(cudaEventElapsedTime(&ms, e0, e1));
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    if (ms_per_launch) *ms_per_launch = ms / ((double)n_rounds * ws.size());
    if (bytes_per_launch) *bytes_per_launch = W0.rows * gguf_row_bytes(W0.type, W0.cols);
    if (ggml_type) *ggml_type = W0.type;
    return 0;
}
} // namespace b2
