// bert.h -- BERT encoder embeddings: the `bert-embeddings` backend of the reference (backend/go/llm/bert/bert.go:18-34:
// bert.New, Embeddings, TokenEmbeddings -> go-bert.cpp -> bert.cpp bert_eval) and llama.cpp's build_bert for the same
// model stored as GGUF (general.architecture = "bert"), which the llama-cpp backend serves through `Embedding`
// (backend/cpp/llama/grpc-server.cpp:2410-2432).  One text -> one vector: mean over tokens, L2-normalised.
//
// Device layout: every linear weight is dequantised once into resident fp16 UMMA operand tiles (the tcgen05 GEMM of
// kernels_gemm.cu runs all six mat-muls of a layer); biases, LayerNorm parameters and the position / token-type tables
// are fp32; activations fp32 rows [token][feature].  Several sequences are packed into one forward (the GEMMs see
// the total token count, attention and pooling work per sequence) -- there is no KV cache and no causal mask.
#pragma once
#include <mutex>
#include <string>
#include <unordered_map>
#include <vector>

#include "common.h"
#include "gguf.h"
#include "kernels.h"

namespace b2 {

// WordPiece as llama.cpp's llm_tokenizer_wpm does it (vocabulary convention: word-initial pieces start with U+2581)
class WpmTokenizer {
  public:
    explicit WpmTokenizer(const GGUFFile &f);
    std::vector<int32_t> tokenize(const std::string &text, int max_tokens) const; // [CLS] pieces... [SEP]
    int n_vocab() const { return (int)tokens_.size(); }
    const std::string &text(int32_t id) const { return tokens_.at((size_t)id); }
    int32_t cls = 2, sep = 3, unk = 1;

  private:
    std::vector<std::string> tokens_;
    std::unordered_map<std::string, int32_t> ids_;
};

struct BertLayer {
    DTensor wq, wk, wv, wo, up, down;
    void *tq = nullptr, *tk = nullptr, *tv = nullptr, *to = nullptr, *tup = nullptr, *tdown = nullptr; // fp16 tiles
    float *bq = nullptr, *bk = nullptr, *bv = nullptr, *bo = nullptr, *bup = nullptr, *bdown = nullptr;
    float *ln1_w = nullptr, *ln1_b = nullptr, *ln2_w = nullptr, *ln2_b = nullptr;
};

class BertModel {
  public:
    BertModel(const std::string &path, int device, bool vocab_only);
    ~BertModel();
    BertModel(const BertModel &) = delete;

    std::vector<int32_t> tokenize(const std::string &text) const { return tok_.tokenize(text, n_ctx); }
    int32_t cls_id() const { return tok_.cls; }
    int32_t sep_id() const { return tok_.sep; }
    std::string token_text(int32_t id) const { return id >= 0 && id < n_vocab ? tok_.text(id) : std::string(); }
    // ids: the sequences back to back; offs[n_seq + 1]: where each starts; out_host: [n_seq][n_embd]
    void embed(const int32_t *ids, const int32_t *offs, int n_seq, float *out_host);
    // the last hidden states [n_tokens][n_embd] of one sequence (parity tests)
    void hidden(const int32_t *ids, int n, float *out_host);

    int n_layer = 0, n_embd = 0, n_head = 0, head_dim = 0, n_ff = 0, n_ctx = 0, n_vocab = 0;
    float eps = 1e-12f;
    int64_t weight_bytes = 0;
    static constexpr int MAX_TOKENS = 32768; // per forward (64 full-length sequences; 1.5 GB of scratch for BERT-base)

  private:
    void *dalloc(size_t bytes);
    DTensor upload(const GGUFFile &f, const std::string &name, int64_t rows, int64_t cols);
    float *upload_f32(const GGUFFile &f, const std::string &name, int64_t rows, int64_t cols); // any type -> fp32 rows
    void *make_tiles(const DTensor &W);
    void forward(const int32_t *ids, const int32_t *offs, int n_seq);
    void run_kernels(int T, int n_seq, int max_len);
    std::vector<cudaGraphExec_t> graphs_; // one text per call: the forward per 32-token length bucket
    std::vector<uint8_t> graph_seen_;

    WpmTokenizer tok_;
    bool vocab_only_ = false;
    int device_ = 0;
    cudaStream_t stream_ = nullptr;
    std::mutex mu_; // the reference backend is single-threaded (base.SingleThread, bert.go:13); so is the scratch here
    std::vector<void *> allocs_;
    DTensor tok_embd_;
    float *pos_embd_ = nullptr, *type0_ = nullptr, *ln0_w_ = nullptr, *ln0_b_ = nullptr;
    std::vector<BertLayer> layers_;
    // scratch for MAX_TOKENS tokens
    int32_t *ids_dev_ = nullptr, *pos_dev_ = nullptr, *seq_dev_ = nullptr, *offs_dev_ = nullptr;
    float *x_ = nullptr, *q_ = nullptr, *k_ = nullptr, *v_ = nullptr, *att_ = nullptr, *h_ = nullptr, *u_ = nullptr, *out_ = nullptr;
    void *x_tiles_ = nullptr, *x2_tiles_ = nullptr;
    float *splitk_ = nullptr;
    size_t splitk_bytes_ = 0;
};

void launch_bert_embed_ln(float *x, const float *pos_embd, const float *type0, const int32_t *pos, const float *w, const float *b,
                          int n_tok, int d, float eps, cudaStream_t s);
// x = LayerNorm(x + h + bias) * w + b (rows wider than 2048 features; narrower rows take the fused kernel of kernels_gemm.cu)
void launch_bert_bias_residual_ln(float *x, const float *h, const float *bias, const float *w, const float *b, int n_tok, int d,
                                  float eps, cudaStream_t s);
// bidirectional attention inside each sequence; q/k/v: [n_tok][n_head * hd] without their biases (added on load)
// head size 64, q | k | v as fp16 rows with their biases (the projection GEMM's out_f16 epilogue): writes attn_output's operand tiles
void launch_bert_attention_f16(const void *q, const void *k, const void *v, const int32_t *offs, int n_seq, int max_len, int n_head,
                               void *x_tiles, cudaStream_t s);
void launch_bert_attention(const float *q, const float *k, const float *v, const float *bq, const float *bk, const float *bv,
                           const int32_t *offs, int n_seq, int max_len, int n_head, int hd, float *out, cudaStream_t s);
void launch_bert_mean_pool_l2(const float *x, const int32_t *offs, int n_seq, int d, float *out, cudaStream_t s);

} // namespace b2
