// engine.h -- the engine behind the C ABI: model + tokenizer + slot scheduler.
// Replaces llama_server_context (backend/cpp/llama/grpc-server.cpp:394-2036) and the task/result queues
// (backend/cpp/llama/utils.hpp:192-409).
#pragma once
#include <map>
#include <memory>
#include <mutex>
#include <string>
#include <vector>

#include "common.h"
#include "model.h"

namespace b2 {

class Tokenizer;
class Scheduler;
class BertModel;

class Engine {
  public:
    explicit Engine(const b2_load_opts &o);
    ~Engine();

    void model_info(b2_model_info_t *out);
    std::vector<int32_t> tokenize(const std::string &text, bool add_bos, bool parse_special);
    std::string token_to_piece(int32_t tok);

    int64_t submit(const b2_request &r);
    int poll(int64_t id, b2_event *ev, int timeout_ms);
    int cancel(int64_t id);
    int embed_result(int64_t id, float *out, int cap);
    int release(int64_t id);
    void metrics(b2_metrics_t *out);

    // synchronous low-level path (holds the device mutex)
    int eval(int slot, const int32_t *tokens, int n, int n_past, float *logits, bool logits_all, float *embd);
    int decode_greedy(int slot, const int32_t *prompt, int n_prompt, int n_gen, int32_t *out, double *prefill_ms,
                      double *decode_ms);
    int debug_hidden(float *out, int cap_floats);
    int bench_decode(int batch, int ctx_len, int n_steps, int n_warmup, double *ms_per_step, int64_t *launches);
    int mega_debug(long long *out, int cap);
    double last_mega_ms() const { return last_mega_ms_; } // mean decode_mega_kernel launch of the last bench_decode(batch 1)
    double last_mega_ms_ = 0.0;
    double last_tp_allreduce_ms_ = 0.0; // tensor-parallel model, TP_NCCL mode: ms inside ncclAllReduce per step of the last bench_decode
    int eval_step(const int32_t *tokens, const int32_t *pos, const int32_t *slots, int n, float *logits);
    int bench_gemv(int which, int n_tok, int n_rounds, double *ms_per_launch, int64_t *bytes_per_launch, int *ggml_type);
    int bench_prefill(int n_tokens, int n_iters, int n_warmup, double *ms_per_iter, int64_t *launches);

    Model &model();
    Tokenizer *tokenizer() { return tok_.get(); }
    std::mutex &device_mutex() { return dev_mu_; }
    bool embeddings_enabled() const { return opts_.embeddings != 0; }

  private:
    void read_hparams_only(const GGUFFile &f);
    Scheduler &sched();
    HParams vocab_hp_;
    void prefill_chunks(int slot, const int32_t *tokens, int n, int n_past, bool want_logits);

    std::unique_ptr<Model> model_;
    // general.architecture = "bert": the engine serves Embedding requests only (what the reference's llama-cpp backend does
    // with a BERT GGUF and `embeddings: true`, grpc-server.cpp:2410-2432); requests are computed inside b2_submit
    std::unique_ptr<BertModel> bert_;
    struct BertResult { std::vector<float> embedding; int n_prompt = 0; double ms = 0; bool polled = false; };
    std::map<int64_t, BertResult> bert_results_;
    std::mutex bert_mu_;
    int64_t bert_next_id_ = 1;
    std::unique_ptr<Tokenizer> tok_;
    std::unique_ptr<Scheduler> sched_;
    std::mutex dev_mu_; // serialises users of the device-side scratch buffers
    b2_load_opts opts_;
    int32_t bos_ = -1, eos_ = -1, eot_ = -1;
    bool add_bos_ = true;
    std::string tok_model_;
};

} // namespace b2
