// bert.cu -- BertModel: load a GGUF "bert" encoder into HBM and run the embeddings forward (see bert.h).
#include "bert.h"

#include <algorithm>
#include <cstdlib>
#include <cstring>

namespace b2 {

// ---- WordPiece ---------------------------------------------------------------------------------------------------------
WpmTokenizer::WpmTokenizer(const GGUFFile &f) {
    const GGUFValue *tv = f.get("tokenizer.ggml.tokens");
    B2_CHECK(tv && !tv->arr_s.empty(), B2_EIO, "gguf: tokenizer.ggml.tokens is missing");
    const std::string model = f.get_s("tokenizer.ggml.model", "");
    B2_CHECK(model == "bert", B2_EUNSUPPORTED, "bert: tokenizer.ggml.model must be \"bert\" (WordPiece), got \"" + model + "\"");
    tokens_ = tv->arr_s;
    for (size_t i = 0; i < tokens_.size(); i++) ids_.emplace(tokens_[i], (int32_t)i); // the first id of a duplicate wins
    cls = (int32_t)f.get_u("tokenizer.ggml.cls_token_id", f.get_u("tokenizer.ggml.bos_token_id", 101));
    sep = (int32_t)f.get_u("tokenizer.ggml.seperator_token_id", f.get_u("tokenizer.ggml.eos_token_id", 102));
    unk = (int32_t)f.get_u("tokenizer.ggml.unknown_token_id", 100);
}

namespace {
inline bool wpm_punct(uint32_t c) { return (c >= 33 && c <= 47) || (c >= 58 && c <= 64) || (c >= 91 && c <= 96) || (c >= 123 && c <= 126); }
inline bool wpm_cjk(uint32_t c) {
    return (c >= 0x4E00 && c <= 0x9FFF) || (c >= 0x3400 && c <= 0x4DBF) || (c >= 0x3040 && c <= 0x30FF) || (c >= 0xF900 && c <= 0xFAFF);
}
inline int utf8_len(uint8_t b) { return b < 0x80 ? 1 : (b >> 5) == 6 ? 2 : (b >> 4) == 14 ? 3 : (b >> 3) == 30 ? 4 : 1; }
} // namespace

std::vector<int32_t> WpmTokenizer::tokenize(const std::string &text, int max_tokens) const {
    // words: runs between white space; punctuation and CJK characters stand alone; ASCII upper case is folded
    std::vector<std::string> words;
    std::string cur;
    for (size_t i = 0; i < text.size();) {
        const int n = std::min<size_t>(utf8_len((uint8_t)text[i]), text.size() - i);
        uint32_t cp = (uint8_t)text[i];
        if (n > 1) {
            cp &= 0xFFu >> (n + 1);
            for (int j = 1; j < n; j++) cp = (cp << 6) | ((uint8_t)text[i + j] & 0x3F);
        }
        if (cp == ' ' || cp == '\t' || cp == '\n' || cp == '\r') {
            if (!cur.empty()) words.push_back(cur);
            cur.clear();
        } else if (wpm_punct(cp) || wpm_cjk(cp)) {
            if (!cur.empty()) words.push_back(cur);
            cur.clear();
            words.push_back(text.substr(i, n));
        } else if (cp >= 'A' && cp <= 'Z') {
            cur.push_back((char)(cp + 32));
        } else {
            cur.append(text, i, n);
        }
        i += n;
    }
    if (!cur.empty()) words.push_back(cur);

    std::vector<int32_t> out;
    out.push_back(cls);
    std::vector<size_t> at; // code point boundaries of the word
    std::vector<int32_t> pieces;
    for (const std::string &w : words) {
        at.clear();
        for (size_t i = 0; i < w.size(); i += std::min<size_t>(utf8_len((uint8_t)w[i]), w.size() - i)) at.push_back(i);
        at.push_back(w.size());
        pieces.clear();
        bool ok = true;
        for (size_t a = 0; a + 1 < at.size();) { // greedy longest match; the first piece of a word carries U+2581
            size_t b = at.size() - 1;
            int32_t id = -1;
            for (; b > a; b--) {
                std::string cand = (a == 0 ? std::string("\xE2\x96\x81") : std::string()) + w.substr(at[a], at[b] - at[a]);
                auto it = ids_.find(cand);
                if (it != ids_.end()) {
                    id = it->second;
                    break;
                }
            }
            if (id < 0) {
                ok = false;
                break;
            }
            pieces.push_back(id);
            a = b;
        }
        if (ok) out.insert(out.end(), pieces.begin(), pieces.end());
        else out.push_back(unk);
    }
    out.push_back(sep);
    if (max_tokens > 1 && (int)out.size() > max_tokens) { // bert.cpp cuts at the model's maximum; the closing [SEP] stays
        out.resize(max_tokens);
        out.back() = sep;
    }
    return out;
}

// ---- model --------------------------------------------------------------------------------------------------------------
void *BertModel::dalloc(size_t bytes) {
    void *p = nullptr;
    if (cudaMalloc(&p, bytes ? bytes : 16) != cudaSuccess) throw Error(B2_ENOMEM, "bert: device allocation of " + std::to_string(bytes) + " bytes failed");
    allocs_.push_back(p);
    return p;
}

DTensor BertModel::upload(const GGUFFile &f, const std::string &name, int64_t rows, int64_t cols) {
    const GGUFTensorInfo *ti = f.tensor(name);
    B2_CHECK(ti, B2_EIO, "gguf: missing tensor " + name);
    B2_CHECK(ti->n_bytes >= 0, B2_EUNSUPPORTED, "gguf: tensor " + name + " has unsupported ggml type " + std::to_string(ti->type));
    const int64_t c = ti->ne[0], r = ti->ne.size() > 1 ? ti->ne[1] : 1;
    B2_CHECK(c == cols && r == rows, B2_EIO, "gguf: tensor " + name + " has an unexpected shape");
    B2_CHECK(cols % std::max(64, type_block_elems(ti->type)) == 0, B2_EUNSUPPORTED, "tensor " + name + ": row length must be a multiple of 64 and of the block size");
    DTensor t;
    t.type = ti->type;
    t.rows = rows;
    t.cols = cols;
    t.row_stride = device_row_stride(ti->type, cols);
    uint8_t *dst = (uint8_t *)dalloc((size_t)t.bytes());
    void *staging = nullptr;
    B2_CUDA(cudaMalloc(&staging, (size_t)ti->n_bytes));
    B2_CUDA(cudaMemcpyAsync(staging, ti->data, (size_t)ti->n_bytes, cudaMemcpyHostToDevice, stream_));
    launch_repack(t.type, (const uint8_t *)staging, dst, rows, cols, stream_);
    B2_CUDA(cudaStreamSynchronize(stream_));
    cudaFree(staging);
    t.data = dst;
    weight_bytes += t.bytes();
    return t;
}

float *BertModel::upload_f32(const GGUFFile &f, const std::string &name, int64_t rows, int64_t cols) {
    const GGUFTensorInfo *ti = f.tensor(name);
    B2_CHECK(ti, B2_EIO, "gguf: missing tensor " + name);
    float *out = (float *)dalloc((size_t)(rows * cols) * 4);
    if (ti->type == B2_T_F32) { // biases, LayerNorm parameters, position / type tables: as they are
        const int64_t n = ti->ne[0] * (ti->ne.size() > 1 ? ti->ne[1] : 1);
        B2_CHECK(n == rows * cols, B2_EIO, "gguf: tensor " + name + " has an unexpected shape");
        B2_CUDA(cudaMemcpyAsync(out, ti->data, (size_t)n * 4, cudaMemcpyHostToDevice, stream_));
        B2_CUDA(cudaStreamSynchronize(stream_));
    } else {
        DTensor t = upload(f, name, rows, cols);
        launch_dequant_rows(t, nullptr, 0, (int)rows, out, cols, stream_);
        B2_CUDA(cudaStreamSynchronize(stream_));
    }
    weight_bytes += rows * cols * 4;
    return out;
}

void *BertModel::make_tiles(const DTensor &W) {
    void *t = dalloc(gemm_w_tiles_bytes(W.rows, W.cols));
    launch_gemm_dequant_w(W, t, stream_);
    return t;
}

BertModel::BertModel(const std::string &path, int device, bool vocab_only) : tok_(GGUFFile(path)), vocab_only_(vocab_only), device_(device) {
    GGUFFile f(path);
    const std::string arch = f.get_s("general.architecture", "");
    B2_CHECK(arch == "bert", B2_EUNSUPPORTED, "bert: general.architecture is '" + arch + "', expected 'bert'");
    n_layer = (int)f.get_u("bert.block_count", 0);
    n_embd = (int)f.get_u("bert.embedding_length", 0);
    n_head = (int)f.get_u("bert.attention.head_count", 0);
    n_ff = (int)f.get_u("bert.feed_forward_length", 0);
    n_ctx = (int)f.get_u("bert.context_length", 512);
    eps = (float)f.get_f("bert.attention.layer_norm_epsilon", 1e-12);
    n_vocab = tok_.n_vocab();
    B2_CHECK(n_layer > 0 && n_embd > 0 && n_head > 0 && n_ff > 0 && n_embd % n_head == 0, B2_EIO, "bert: bad hyper-parameters");
    head_dim = n_embd / n_head;
    B2_CHECK(head_dim == 32 || head_dim == 64, B2_EUNSUPPORTED, "bert: head size must be 32 or 64");
    B2_CHECK(f.get_u("bert.pooling_type", 1) == 1, B2_EUNSUPPORTED, "bert: only mean pooling is built");
    if (vocab_only) return;

    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) throw Error(B2_ENODEV, "no CUDA device visible (this library has no CPU fallback)");
    B2_CHECK(device >= 0 && device < ndev, B2_EINVAL, "device ordinal out of range");
    B2_CUDA(cudaSetDevice(device));
    cudaDeviceProp prop;
    B2_CUDA(cudaGetDeviceProperties(&prop, device));
    B2_CHECK(prop.major == 10, B2_ENODEV, "this library is built for sm_100a only");
    B2_CUDA(cudaStreamCreateWithFlags(&stream_, cudaStreamNonBlocking));

    const int d = n_embd, ff = n_ff;
    tok_embd_ = upload(f, "token_embd.weight", n_vocab, d);
    pos_embd_ = upload_f32(f, "position_embd.weight", n_ctx, d);
    {
        const GGUFTensorInfo *tt = f.tensor("token_types.weight");
        B2_CHECK(tt, B2_EIO, "gguf: missing tensor token_types.weight");
        type0_ = upload_f32(f, "token_types.weight", tt->ne.size() > 1 ? tt->ne[1] : 1, d); // row 0 is the one used
    }
    ln0_w_ = upload_f32(f, "token_embd_norm.weight", 1, d);
    ln0_b_ = upload_f32(f, "token_embd_norm.bias", 1, d);
    layers_.resize(n_layer);
    for (int i = 0; i < n_layer; i++) {
        BertLayer &L = layers_[i];
        const std::string p = "blk." + std::to_string(i) + ".";
        L.wq = upload(f, p + "attn_q.weight", d, d);
        L.wk = upload(f, p + "attn_k.weight", d, d);
        L.wv = upload(f, p + "attn_v.weight", d, d);
        L.wo = upload(f, p + "attn_output.weight", d, d);
        L.up = upload(f, p + "ffn_up.weight", ff, d);
        L.down = upload(f, p + "ffn_down.weight", d, ff);
        L.tq = make_tiles(L.wq); L.tk = make_tiles(L.wk); L.tv = make_tiles(L.wv); L.to = make_tiles(L.wo);
        L.tup = make_tiles(L.up); L.tdown = make_tiles(L.down);
        L.bq = upload_f32(f, p + "attn_q.bias", 1, d); L.bk = upload_f32(f, p + "attn_k.bias", 1, d);
        L.bv = upload_f32(f, p + "attn_v.bias", 1, d); L.bo = upload_f32(f, p + "attn_output.bias", 1, d);
        L.bup = upload_f32(f, p + "ffn_up.bias", 1, ff); L.bdown = upload_f32(f, p + "ffn_down.bias", 1, d);
        L.ln1_w = upload_f32(f, p + "attn_output_norm.weight", 1, d); L.ln1_b = upload_f32(f, p + "attn_output_norm.bias", 1, d);
        L.ln2_w = upload_f32(f, p + "layer_output_norm.weight", 1, d); L.ln2_b = upload_f32(f, p + "layer_output_norm.bias", 1, d);
    }
    const size_t T = MAX_TOKENS;
    ids_dev_ = (int32_t *)dalloc(T * 4);
    pos_dev_ = (int32_t *)dalloc(T * 4);
    offs_dev_ = (int32_t *)dalloc((T + 1) * 4);
    for (float **b : {&x_, &q_, &k_, &v_, &att_, &h_}) *b = (float *)dalloc(T * d * 4);
    u_ = (float *)dalloc(T * ff * 4);
    out_ = (float *)dalloc(T * d * 4);
    x_tiles_ = dalloc(gemm_x_tiles_bytes((int)T, std::max(d, ff)));
    x2_tiles_ = dalloc(gemm_x_tiles_bytes((int)T, ff)); // written by the feed-forward GEMM's epilogue (gelu(u + b) as operand tiles)
    splitk_bytes_ = (size_t)32 << 20;
    splitk_ = (float *)dalloc(splitk_bytes_);
    B2_CUDA(cudaStreamSynchronize(stream_));
}

BertModel::~BertModel() {
    if (vocab_only_) return;
    cudaSetDevice(device_);
    if (stream_) cudaStreamSynchronize(stream_);
    for (cudaGraphExec_t g : graphs_)
        if (g) cudaGraphExecDestroy(g);
    for (void *p : allocs_) cudaFree(p);
    if (stream_) cudaStreamDestroy(stream_);
}

void BertModel::forward(const int32_t *ids, const int32_t *offs, int n_seq) {
    B2_CHECK(!vocab_only_, B2_ESTATE, "bert: loaded with vocab_only, no model on the device");
    B2_CHECK(n_seq > 0 && offs[0] == 0, B2_EINVAL, "bert: bad sequence offsets");
    const int T = offs[n_seq];
    B2_CHECK(T > 0 && T <= MAX_TOKENS, B2_EINVAL, "bert: between 1 and " + std::to_string(MAX_TOKENS) + " tokens per call");
    int max_len = 0;
    for (int s = 0; s < n_seq; s++) {
        const int len = offs[s + 1] - offs[s];
        B2_CHECK(len > 0 && len <= n_ctx, B2_EINVAL, "bert: a sequence has " + std::to_string(len) + " tokens; the model takes 1.." + std::to_string(n_ctx));
        max_len = std::max(max_len, len);
    }
    for (int i = 0; i < T; i++) B2_CHECK(ids[i] >= 0 && ids[i] < n_vocab, B2_EINVAL, "bert: token id out of range");
    B2_CUDA(cudaSetDevice(device_));
    // One text per call (the reference API's shape) is launch-bound below ~128 tokens: its ~110 kernels are replayed from
    // a CUDA graph captured per length bucket of 32 tokens.  The rows between the text's end and the bucket's are
    // padding: token 0 at the following positions, computed and ignored (attention and pooling read the true length).
    const bool debug_env = (getenv("B200LLM_BERT_UNFUSED_LN") && atoi(getenv("B200LLM_BERT_UNFUSED_LN"))) ||
                           (getenv("B200LLM_BERT_ATT_SIMPLE") && atoi(getenv("B200LLM_BERT_ATT_SIMPLE"))) ||
                           (getenv("B200LLM_BERT_NO_GRAPH") && atoi(getenv("B200LLM_BERT_NO_GRAPH")));
    const bool bucketed = n_seq == 1 && !debug_env;
    const int Tp = bucketed ? std::min((T + 31) / 32 * 32, n_ctx) : T;
    std::vector<int32_t> pos(Tp), idp(Tp, 0);
    memcpy(idp.data(), ids, (size_t)T * 4);
    for (int s = 0; s < n_seq; s++)
        for (int i = offs[s]; i < offs[s + 1]; i++) pos[i] = i - offs[s];
    for (int i = T; i < Tp; i++) pos[i] = i;
    cudaStream_t s = stream_;
    B2_CUDA(cudaMemcpyAsync(ids_dev_, idp.data(), (size_t)Tp * 4, cudaMemcpyHostToDevice, s));
    B2_CUDA(cudaMemcpyAsync(pos_dev_, pos.data(), (size_t)Tp * 4, cudaMemcpyHostToDevice, s));
    B2_CUDA(cudaMemcpyAsync(offs_dev_, offs, (size_t)(n_seq + 1) * 4, cudaMemcpyHostToDevice, s));
    if (!bucketed) {
        run_kernels(T, n_seq, max_len);
        return;
    }
    B2_CUDA(cudaStreamSynchronize(s)); // the host vectors above go out of scope; also keeps the copies out of the capture
    const int bi = Tp / 32;
    if ((int)graphs_.size() <= bi) { graphs_.resize(bi + 1, nullptr); graph_seen_.resize(bi + 1, 0); }
    if (graphs_[bi]) {
        B2_CUDA(cudaGraphLaunch(graphs_[bi], s));
    } else if (!graph_seen_[bi]) { // first time: eagerly (one-time kernel attributes are set outside any capture)
        graph_seen_[bi] = 1;
        run_kernels(Tp, 1, Tp);
    } else if (graph_seen_[bi] == 2) { // capture failed once for this bucket: stay eager
        run_kernels(Tp, 1, Tp);
    } else {
        cudaGraph_t gr = nullptr;
        bool ok = cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal) == cudaSuccess;
        if (ok) {
            try {
                run_kernels(Tp, 1, Tp);
            } catch (...) {
                ok = false;
            }
            ok = (cudaStreamEndCapture(s, &gr) == cudaSuccess) && ok && gr;
        }
        if (ok) ok = cudaGraphInstantiate(&graphs_[bi], gr, 0) == cudaSuccess;
        if (gr) cudaGraphDestroy(gr);
        if (!ok) { // nothing ran: forget the error state and do this call the plain way
            graphs_[bi] = nullptr;
            graph_seen_[bi] = 2;
            cudaGetLastError();
            run_kernels(Tp, 1, Tp);
        } else {
            B2_CUDA(cudaGraphLaunch(graphs_[bi], s));
        }
    }
}

// the kernels of one forward; every input (ids, positions, offsets) is already in its device buffer
void BertModel::run_kernels(int T, int n_seq, int max_len) {
    cudaStream_t s = stream_;
    const int d = n_embd, ff = n_ff;
    launch_dequant_rows(tok_embd_, ids_dev_, 0, T, x_, d, s);
    launch_bert_embed_ln(x_, pos_embd_, type0_, pos_dev_, ln0_w_, ln0_b_, T, d, eps, s);
    // LayerNorm rows of up to 2048 features are produced together with their operand tiles (one pass over the row)
    const bool fused_ln = d <= 2048 && !(getenv("B200LLM_BERT_UNFUSED_LN") && atoi(getenv("B200LLM_BERT_UNFUSED_LN")) != 0);
    // A/B switch, off: the projections back into the residual stream add their bias and the residual in the GEMM's epilogue, in
    // place in x_, and the LayerNorm reads one fp32 row instead of two.  Measured slower (BERT-base, 64 x 512 tokens per call:
    // 5.27 k against 5.84 k seq/s): these K = 768 GEMMs are bound by their epilogue, which the extra reads lengthen by more
    // than the LayerNorm saves.
    const bool fused_res = fused_ln && T > 128 && getenv("B200LLM_BERT_FUSED_RES") && atoi(getenv("B200LLM_BERT_FUSED_RES")) != 0;
    auto to_residual = [&](const DTensor &W, void *tiles, const void *in_tiles, const float *bias) {
        if (fused_res) {
            const GemmJob j = {&W, tiles, true, x_, d, bias};
            launch_gemm_tc_multi(&j, 1, in_tiles, T, x_, d, nullptr, 0, s);
        } else {
            launch_gemm_tc(W, in_tiles, T, h_, d, nullptr, 0, tiles, true, splitk_, splitk_bytes_, s);
        }
    };
    auto ln = [&](const float *bias, const float *w, const float *b) {
        if (fused_res) {
            launch_gemm_prepare_x_bias_residual_ln(x_, nullptr, nullptr, w, b, T, d, eps, x_tiles_, s);
        } else if (fused_ln) {
            launch_gemm_prepare_x_bias_residual_ln(x_, h_, bias, w, b, T, d, eps, x_tiles_, s);
        } else {
            launch_bert_bias_residual_ln(x_, h_, bias, w, b, T, d, eps, s);
            launch_gemm_prepare_x(x_, d, T, d, x_tiles_, s);
        }
    };
    // the fused epilogue runs without split-K: only where the feed-forward GEMM has enough output tiles for the SMs anyway
    const bool fused_gelu = ff % 256 == 0 && (int64_t)((T + 127) / 128) * (ff / 256) > 74 &&
                            !(getenv("B200LLM_BERT_UNFUSED_GELU") && atoi(getenv("B200LLM_BERT_UNFUSED_GELU")) != 0);
    // More than one token tile (the projection is not split over K then) and heads of 64: attention on fp16 rows
    const bool f16_att = head_dim == 64 && T > 128 && d % 8 == 0 &&
                         !(getenv("B200LLM_BERT_ATT_F32") && atoi(getenv("B200LLM_BERT_ATT_F32")) != 0) &&
                         !(getenv("B200LLM_BERT_ATT_SIMPLE") && atoi(getenv("B200LLM_BERT_ATT_SIMPLE")) != 0);
    launch_gemm_prepare_x(x_, d, T, d, x_tiles_, s);
    for (const BertLayer &L : layers_) { // on entry x_tiles_ holds x_
        if (f16_att) { // q | k | v leave the GEMM as fp16 rows with their biases; the attention writes attn_output's operand tiles
            const GemmJob qkv[3] = {{&L.wq, L.tq, true, q_, d, L.bq}, {&L.wk, L.tk, true, k_, d, L.bk}, {&L.wv, L.tv, true, v_, d, L.bv}};
            launch_gemm_tc_multi(qkv, 3, x_tiles_, T, nullptr, 0, nullptr, 0, s, true);
            launch_bert_attention_f16(q_, k_, v_, offs_dev_, n_seq, max_len, n_head, x_tiles_, s);
        } else {
            const GemmJob qkv[3] = {{&L.wq, L.tq, true, q_, d}, {&L.wk, L.tk, true, k_, d}, {&L.wv, L.tv, true, v_, d}};
            launch_gemm_tc_multi(qkv, 3, x_tiles_, T, nullptr, 0, splitk_, splitk_bytes_, s);
            launch_bert_attention(q_, k_, v_, L.bq, L.bk, L.bv, offs_dev_, n_seq, max_len, n_head, head_dim, att_, s);
            launch_gemm_prepare_x(att_, d, T, d, x_tiles_, s);
        }
        to_residual(L.wo, L.to, x_tiles_, L.bo);
        ln(L.bo, L.ln1_w, L.ln1_b);
        if (fused_gelu) { // bias + GELU in the GEMM's epilogue, which writes the down projection's operand tiles
            launch_gemm_tc_bias_gelu(ff, d, L.tup, x_tiles_, T, L.bup, x2_tiles_, s);
            to_residual(L.down, L.tdown, x2_tiles_, L.bdown);
        } else {
            launch_gemm_tc(L.up, x_tiles_, T, u_, ff, nullptr, 0, L.tup, true, splitk_, splitk_bytes_, s);
            launch_gemm_prepare_x_bias_gelu(u_, L.bup, T, ff, x_tiles_, s); // gelu(u + b) straight into the operand tiles
            to_residual(L.down, L.tdown, x_tiles_, L.bdown);
        }
        ln(L.bdown, L.ln2_w, L.ln2_b);
    }
}

void BertModel::embed(const int32_t *ids, const int32_t *offs, int n_seq, float *out_host) {
    std::lock_guard<std::mutex> lk(mu_);
    forward(ids, offs, n_seq);
    launch_bert_mean_pool_l2(x_, offs_dev_, n_seq, n_embd, out_, stream_);
    B2_CUDA(cudaMemcpyAsync(out_host, out_, (size_t)n_seq * n_embd * 4, cudaMemcpyDeviceToHost, stream_));
    B2_CUDA(cudaStreamSynchronize(stream_));
}

void BertModel::hidden(const int32_t *ids, int n, float *out_host) {
    std::lock_guard<std::mutex> lk(mu_);
    const int32_t offs[2] = {0, n};
    forward(ids, offs, 1);
    B2_CUDA(cudaMemcpyAsync(out_host, x_, (size_t)n * n_embd * 4, cudaMemcpyDeviceToHost, stream_));
    B2_CUDA(cudaStreamSynchronize(stream_));
}

} // namespace b2
