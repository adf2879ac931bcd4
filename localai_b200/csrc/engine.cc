// engine.cc -- Engine: LoadModel + the synchronous low-level entry points.  The request scheduler
// (submit / poll) lives in scheduler.cc, the tokenisers in tokenizer.cc.
#include "engine.h"

#include <algorithm>
#include <cstring>

#include <chrono>

#include "bert.h"
#include "scheduler.h"
#include "tokenizer.h"

namespace b2 {

Engine::Engine(const b2_load_opts &o) : opts_(o) {
    GGUFFile f(o.model_path);
    if (f.get_s("general.architecture", "") == "bert") { // encoder: its own loader, tokeniser and forward (bert.cu)
        bert_.reset(new BertModel(o.model_path, o.device, o.vocab_only != 0));
        read_hparams_only(f);
        vocab_hp_.n_vocab = bert_->n_vocab;
        vocab_hp_.n_ctx_train = bert_->n_ctx;
        tok_model_ = "bert";
        bos_ = bert_->cls_id();
        eos_ = bert_->sep_id();
        return;
    }
    if (!o.vocab_only) model_.reset(new Model(f, o));
    else read_hparams_only(f);
    tok_.reset(new Tokenizer(f));
    bos_ = tok_->bos_id();
    eos_ = tok_->eos_id();
    eot_ = tok_->eot_id();
    add_bos_ = tok_->add_bos();
    tok_model_ = tok_->model_name();
    if (model_) sched_.reset(new Scheduler(this));
}

void Engine::read_hparams_only(const GGUFFile &f) {
    const std::string a = f.get_s("general.architecture", "llama") + ".";
    vocab_hp_.arch = f.get_s("general.architecture", "");
    vocab_hp_.name = f.get_s("general.name", "");
    vocab_hp_.n_layer = (int)f.get_u(a + "block_count", 0);
    vocab_hp_.n_embd = (int)f.get_u(a + "embedding_length", 0);
    vocab_hp_.n_head = (int)f.get_u(a + "attention.head_count", 0);
    vocab_hp_.n_head_kv = (int)f.get_u(a + "attention.head_count_kv", vocab_hp_.n_head);
    vocab_hp_.n_ff = (int)f.get_u(a + "feed_forward_length", 0);
    vocab_hp_.n_ctx_train = (int)f.get_u(a + "context_length", 0);
    vocab_hp_.head_dim = vocab_hp_.n_head ? vocab_hp_.n_embd / vocab_hp_.n_head : 0;
    vocab_hp_.rms_eps = (float)f.get_f(a + "attention.layer_norm_rms_epsilon", 1e-5);
    vocab_hp_.rope_base = (float)f.get_f(a + "rope.freq_base", 10000.0);
    if (const GGUFTensorInfo *te = f.tensor("token_embd.weight")) vocab_hp_.n_vocab = te->ne.size() > 1 ? (int)te->ne[1] : 0;
}

Model &Engine::model() {
    B2_CHECK(!bert_, B2_EUNSUPPORTED, "this entry point needs a decoder model; a BERT encoder serves Embedding requests only");
    B2_CHECK(model_ != nullptr, B2_ESTATE, "engine was created with vocab_only: no model is loaded");
    return *model_;
}

Engine::~Engine() {
    bert_.reset();
    sched_.reset(); // joins the engine thread before the model goes away
    tok_.reset();
    model_.reset();
}

void Engine::model_info(b2_model_info_t *o) {
    memset(o, 0, sizeof(*o));
    const HParams &h = model_ ? model_->hp : vocab_hp_;
    o->n_layer = h.n_layer; o->n_embd = h.n_embd; o->n_head = h.n_head; o->n_head_kv = h.n_head_kv;
    o->head_dim = h.head_dim; o->n_ff = h.n_ff; o->n_vocab = h.n_vocab; o->n_ctx_train = h.n_ctx_train;
    if (model_) {
        o->n_ctx = model_->n_ctx; o->n_ctx_slot = model_->n_ctx_slot; o->n_parallel = model_->n_parallel;
        o->n_batch = model_->n_batch;
        o->weight_bytes_device = model_->weight_bytes;
        o->decode_stream_bytes = model_->decode_stream_bytes;
        o->kv_bytes_per_pos = model_->kv_bytes_per_pos;
    }
    if (bert_) o->weight_bytes_device = bert_->weight_bytes;
    o->bos_id = bos_; o->eos_id = eos_; o->eot_id = eot_; o->add_bos = add_bos_ ? 1 : 0;
    o->rms_eps = h.rms_eps; o->rope_freq_base = h.rope_base; o->rope_freq_scale = h.rope_scale;
    snprintf(o->arch, sizeof(o->arch), "%s", h.arch.c_str());
    snprintf(o->tokenizer_model, sizeof(o->tokenizer_model), "%s", tok_model_.c_str());
    snprintf(o->name, sizeof(o->name), "%s", h.name.c_str());
}

std::vector<int32_t> Engine::tokenize(const std::string &text, bool add_bos, bool parse_special) {
    if (bert_) return bert_->tokenize(text); // [CLS] ... [SEP] always, as bert.cpp and llama.cpp's WordPiece do
    return tok_->tokenize(text, add_bos, parse_special);
}
// special = true, like common's llama_token_to_piece(ctx, token) default used at grpc-server.cpp:1012
std::string Engine::token_to_piece(int32_t t) { return bert_ ? bert_->token_text(t) : tok_->token_to_piece(t, true); }

int64_t Engine::submit(const b2_request &r) {
    if (!bert_) return sched().submit(r);
    B2_CHECK(r.embedding_mode, B2_EUNSUPPORTED, "a BERT encoder serves Embedding requests only");
    std::vector<int32_t> ids;
    if (r.prompt_ids && r.n_prompt_ids > 0) ids.assign(r.prompt_ids, r.prompt_ids + r.n_prompt_ids);
    else ids = bert_->tokenize(std::string(r.prompt ? r.prompt : "", r.prompt ? r.prompt_len : 0));
    BertResult res;
    res.n_prompt = (int)ids.size();
    const auto t0 = std::chrono::steady_clock::now();
    if (opts_.embeddings) { // embeddings disabled -> zeros, as send_embedding does (grpc-server.cpp:1301-1310)
        res.embedding.resize(bert_->n_embd);
        const int32_t offs[2] = {0, (int32_t)ids.size()};
        bert_->embed(ids.data(), offs, 1, res.embedding.data());
    } else {
        res.embedding.assign(bert_->n_embd, 0.f);
    }
    res.ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    std::lock_guard<std::mutex> lk(bert_mu_);
    const int64_t id = bert_next_id_++;
    bert_results_.emplace(id, std::move(res));
    return id;
}

int Engine::poll(int64_t id, b2_event *ev, int timeout_ms) {
    if (!bert_) return sched().poll(id, ev, timeout_ms);
    std::lock_guard<std::mutex> lk(bert_mu_);
    auto it = bert_results_.find(id);
    B2_CHECK(it != bert_results_.end(), B2_ESTATE, "unknown request id");
    if (it->second.polled) return B2_ETIMEOUT; // the one (final) event was handed out already
    it->second.polled = true;
    const uint32_t sz = ev->struct_size ? std::min<uint32_t>(ev->struct_size, sizeof(b2_event)) : (uint32_t)sizeof(b2_event);
    b2_event out;
    memset(&out, 0, sizeof(out));
    out.struct_size = sz;
    out.flags = B2_EV_FINAL | B2_EV_EMBEDDING;
    out.text = "";
    out.n_prompt_tokens = it->second.n_prompt;
    out.n_embd = (int)it->second.embedding.size();
    out.t_prompt_ms = it->second.ms;
    memcpy(ev, &out, sz);
    return 0;
}

int Engine::cancel(int64_t id) {
    if (!bert_) return sched().cancel(id);
    std::lock_guard<std::mutex> lk(bert_mu_);
    B2_CHECK(bert_results_.count(id), B2_ESTATE, "unknown request id");
    return 0;
}

int Engine::embed_result(int64_t id, float *out, int cap) {
    if (!bert_) return sched().embed_result(id, out, cap);
    std::lock_guard<std::mutex> lk(bert_mu_);
    auto it = bert_results_.find(id);
    B2_CHECK(it != bert_results_.end(), B2_ESTATE, "unknown request id");
    const int n = (int)it->second.embedding.size();
    if (n > cap) return -n;
    memcpy(out, it->second.embedding.data(), (size_t)n * 4);
    return n;
}

int Engine::release(int64_t id) {
    if (!bert_) return sched().release(id);
    std::lock_guard<std::mutex> lk(bert_mu_);
    B2_CHECK(bert_results_.erase(id) == 1, B2_ESTATE, "unknown request id");
    return 0;
}

void Engine::metrics(b2_metrics_t *out) {
    if (bert_) {
        memset(out, 0, sizeof(*out));
        return;
    }
    sched().metrics(out);
}

Scheduler &Engine::sched() {
    B2_CHECK(sched_ != nullptr, B2_ESTATE, "engine was created with vocab_only: no model is loaded");
    return *sched_;
}

// ---------------------------------------------------------------------------------------------------------
void Engine::prefill_chunks(int slot, const int32_t *tokens, int n, int n_past, bool want_logits) {
    Model &m = model();
    std::vector<int32_t> pos(m.n_batch), sl(m.n_batch, slot);
    for (int i0 = 0; i0 < n; i0 += m.n_batch) { // chunked prefill (grpc-server.cpp:1864-1866)
        const int c = std::min(m.n_batch, n - i0);
        for (int i = 0; i < c; i++) pos[i] = n_past + i0 + i;
        Batch b;
        b.n = c;
        b.tokens = tokens + i0;
        b.pos = pos.data();
        b.slot = sl.data();
        const bool last = i0 + c == n;
        int32_t row = c - 1;
        if (last && want_logits) { // logits only for the last prompt token (grpc-server.cpp:1847-1850)
            b.out_rows = &row;
            b.n_out = 1;
        }
        m.forward(b);
        m.sync(); // host arrays above are reused by the next chunk
    }
}

int Engine::eval(int slot, const int32_t *tokens, int n, int n_past, float *logits, bool logits_all, float *embd) {
    Model &m = model();
    B2_CHECK(tokens && n > 0, B2_EINVAL, "eval: no tokens");
    B2_CHECK(slot >= 0 && slot < m.n_parallel, B2_EINVAL, "eval: slot out of range");
    B2_CHECK(n_past >= 0 && n_past + n <= m.n_ctx_slot, B2_ENOSPC, "eval: context exhausted");
    std::lock_guard<std::mutex> lk(dev_mu_);
    const int V = m.hp.n_vocab, d = m.hp.n_embd;
    if (!logits_all) {
        prefill_chunks(slot, tokens, n, n_past, true);
        if (logits) B2_CUDA(cudaMemcpyAsync(logits, m.logits_dev(), (size_t)V * 4, cudaMemcpyDeviceToHost, m.stream));
        if (embd) B2_CUDA(cudaMemcpyAsync(embd, m.embd_dev(), (size_t)d * 4, cudaMemcpyDeviceToHost, m.stream));
        m.sync();
        return 0;
    }
    std::vector<int32_t> pos(m.n_batch), sl(m.n_batch, slot);
    for (int i0 = 0; i0 < n; i0 += m.n_batch) {
        const int c = std::min(m.n_batch, n - i0);
        for (int i = 0; i < c; i++) pos[i] = n_past + i0 + i;
        Batch b;
        b.n = c;
        b.tokens = tokens + i0;
        b.pos = pos.data();
        b.slot = sl.data();
        b.n_out = -1;
        m.forward(b);
        if (logits)
            B2_CUDA(cudaMemcpyAsync(logits + (size_t)i0 * V, m.logits_dev(), (size_t)c * V * 4, cudaMemcpyDeviceToHost, m.stream));
        if (embd && i0 + c == n)
            B2_CUDA(cudaMemcpyAsync(embd, m.embd_dev() + (size_t)(c - 1) * d, (size_t)d * 4, cudaMemcpyDeviceToHost, m.stream));
        m.sync();
    }
    return 0;
}

// one decode step over several slots (what the scheduler issues for concurrent streams): token i at position pos[i] of
// slot slots[i], every row's logits back.  Test / measurement hook (b2_eval_step).
int Engine::eval_step(const int32_t *tokens, const int32_t *pos, const int32_t *slots, int n, float *logits) {
    Model &m = model();
    B2_CHECK(tokens && pos && slots && n > 0 && n <= m.n_batch, B2_EINVAL, "eval_step: 1..n_batch tokens");
    std::lock_guard<std::mutex> lk(dev_mu_);
    Batch b;
    b.n = n;
    b.tokens = tokens;
    b.pos = pos;
    b.slot = slots;
    b.n_out = -1;
    m.forward(b);
    if (logits) B2_CUDA(cudaMemcpyAsync(logits, m.logits_dev(), (size_t)n * m.hp.n_vocab * 4, cudaMemcpyDeviceToHost, m.stream));
    m.sync();
    return 0;
}

int Engine::decode_greedy(int slot, const int32_t *prompt, int n_prompt, int n_gen, int32_t *out, double *prefill_ms,
                          double *decode_ms) {
    Model &m = model();
    B2_CHECK(prompt && n_prompt > 0 && n_gen >= 0 && out, B2_EINVAL, "decode_greedy: bad arguments");
    B2_CHECK(slot >= 0 && slot < m.n_parallel, B2_EINVAL, "decode_greedy: slot out of range");
    B2_CHECK(n_prompt <= m.n_ctx_slot, B2_ENOSPC, "decode_greedy: prompt longer than the slot context");
    std::lock_guard<std::mutex> lk(dev_mu_);
    cudaEvent_t e0, e1, e2;
    B2_CUDA(cudaEventCreate(&e0));
    B2_CUDA(cudaEventCreate(&e1));
    B2_CUDA(cudaEventCreate(&e2));
    int32_t *host_out = nullptr;
    B2_CUDA(cudaMallocHost((void **)&host_out, std::max(1, n_gen) * sizeof(int32_t)));
    int n_done = 0;
    try {
        B2_CUDA(cudaEventRecord(e0, m.stream));
        prefill_chunks(slot, prompt, n_prompt, 0, true);
        B2_CUDA(cudaEventRecord(e1, m.stream));
        // device-side loop: argmax feeds the token buffer and bumps the position; only ids come back
        int32_t z = 0, p = n_prompt - 1, s = slot;
        m.set_device_inputs(&z, &p, &s, 1);
        m.sync();
        const int max_gen = std::min(n_gen, m.n_ctx_slot - n_prompt + 1);
        for (int i = 0; i < max_gen; i++) {
            m.sample_greedy(1, true);
            B2_CUDA(cudaMemcpyAsync(host_out + i, m.next_tok_dev(), 4, cudaMemcpyDeviceToHost, m.stream));
            n_done++;
            if (i + 1 < max_gen) m.forward_device_inputs(1);
        }
        B2_CUDA(cudaEventRecord(e2, m.stream));
        m.sync();
        memcpy(out, host_out, (size_t)n_done * 4);
        float a = 0, b = 0;
        B2_CUDA(cudaEventElapsedTime(&a, e0, e1));
        B2_CUDA(cudaEventElapsedTime(&b, e1, e2));
        if (prefill_ms) *prefill_ms = a;
        if (decode_ms) *decode_ms = b;
    } catch (...) {
        cudaFreeHost(host_out);
        cudaEventDestroy(e0); cudaEventDestroy(e1); cudaEventDestroy(e2);
        throw;
    }
    cudaFreeHost(host_out);
    cudaEventDestroy(e0); cudaEventDestroy(e1); cudaEventDestroy(e2);
    return n_done;
}

int Engine::debug_hidden(float *out, int cap_floats) {
    Model &m = model();
    std::lock_guard<std::mutex> lk(dev_mu_);
    const int n = (m.hp.n_layer + 1) * m.hp.n_embd;
    if (!m.debug_hidden_dev()) { // first call only arms the capture
        m.enable_debug_hidden();
        return 0;
    }
    B2_CHECK(out && cap_floats >= n, B2_ENOSPC, "debug_hidden: buffer too small");
    m.sync();
    B2_CUDA(cudaMemcpy(out, m.debug_hidden_dev(), (size_t)n * 4, cudaMemcpyDeviceToHost));
    return n;
}

int Engine::bench_decode(int batch, int ctx_len, int n_steps, int n_warmup, double *ms_per_step, int64_t *launches) {
    Model &m = model();
    B2_CHECK(batch > 0 && batch <= m.n_parallel && batch <= m.n_batch, B2_EINVAL, "bench_decode: batch must be <= n_parallel");
    B2_CHECK(ctx_len >= 0 && ctx_len < m.n_ctx_slot, B2_EINVAL, "bench_decode: ctx_len must be < n_ctx_slot");
    B2_CHECK(n_steps > 0, B2_EINVAL, "bench_decode: n_steps");
    std::lock_guard<std::mutex> lk(dev_mu_);
    std::vector<int32_t> tok(batch), pos(batch, ctx_len), sl(batch);
    for (int i = 0; i < batch; i++) { tok[i] = (int32_t)((1 + 7919 * i) % m.hp.n_vocab); sl[i] = i; }
    m.set_device_inputs(tok.data(), pos.data(), sl.data(), batch);
    m.sync();
    cudaEvent_t e0, e1;
    B2_CUDA(cudaEventCreate(&e0));
    B2_CUDA(cudaEventCreate(&e1));
    auto step = [&]() {
        m.forward_device_inputs(batch);
        launch_argmax(m.logits_dev(), m.hp.n_vocab, batch, m.hp.n_vocab, const_cast<int32_t *>(m.next_tok_dev()), nullptr, nullptr,
                      m.stream);
    };
    for (int i = 0; i < n_warmup; i++) step();
    m.sync();
    const int64_t l0 = g_launch_count;
    if (batch == 1 && m.tp_size == 1) m.mega_timing_begin(n_steps);
    if (m.tp_size > 1) m.tp_timing(true);
    B2_CUDA(cudaEventRecord(e0, m.stream));
    for (int i = 0; i < n_steps; i++) step();
    B2_CUDA(cudaEventRecord(e1, m.stream));
    m.sync();
    last_mega_ms_ = batch == 1 && m.tp_size == 1 ? m.mega_timing_end() : 0.0;
    if (m.tp_size > 1) {
        m.tp_sync();
        last_tp_allreduce_ms_ = m.tp_allreduce_ms() / n_steps;
        m.tp_timing(false);
    }
    float ms = 0;
    B2_CUDA(cudaEventElapsedTime(&ms, e0, e1));
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    if (ms_per_step) *ms_per_step = ms / n_steps;
    if (launches) *launches = (g_launch_count - l0) / n_steps;
    return 0;
}

int Engine::mega_debug(long long *out, int cap) {
    Model &m = model();
    std::lock_guard<std::mutex> lk(dev_mu_);
    const int n = m.n_sm() * m.mega_n_phases() * 16;
    if (!m.mega_dbg()) return 0;
    if (!out) return n;
    B2_CHECK(cap >= n, B2_ENOSPC, "mega_debug: buffer too small");
    m.sync();
    B2_CUDA(cudaMemcpy(out, m.mega_dbg(), (size_t)n * sizeof(long long), cudaMemcpyDeviceToHost));
    return n;
}

int Engine::bench_gemv(int which, int n_tok, int n_rounds, double *ms_per_launch, int64_t *bytes_per_launch, int *ggml_type) {
    Model &m = model();
    B2_CHECK(which >= 0 && which <= 9 && n_tok >= 1 && n_tok <= QGEMM_MAX_TOKENS && n_rounds > 0, B2_EINVAL, "bench_gemv: bad arguments");
    std::lock_guard<std::mutex> lk(dev_mu_);
    return m.bench_gemv(which, n_tok, n_rounds, ms_per_launch, bytes_per_launch, ggml_type);
}

int Engine::bench_prefill(int n_tokens, int n_iters, int n_warmup, double *ms_per_iter, int64_t *launches) {
    Model &m = model();
    B2_CHECK(n_tokens > 0 && n_tokens <= m.n_ctx_slot, B2_EINVAL, "bench_prefill: n_tokens must be <= n_ctx_slot");
    B2_CHECK(n_iters > 0, B2_EINVAL, "bench_prefill: n_iters");
    std::lock_guard<std::mutex> lk(dev_mu_);
    std::vector<int32_t> tok(n_tokens);
    for (int i = 0; i < n_tokens; i++) tok[i] = (int32_t)((1 + 7919LL * i) % m.hp.n_vocab);
    cudaEvent_t e0, e1;
    B2_CUDA(cudaEventCreate(&e0));
    B2_CUDA(cudaEventCreate(&e1));
    for (int i = 0; i < n_warmup; i++) prefill_chunks(0, tok.data(), n_tokens, 0, true);
    m.sync();
    const int64_t l0 = g_launch_count;
    B2_CUDA(cudaEventRecord(e0, m.stream));
    for (int i = 0; i < n_iters; i++) prefill_chunks(0, tok.data(), n_tokens, 0, true);
    B2_CUDA(cudaEventRecord(e1, m.stream));
    m.sync();
    float ms = 0;
    B2_CUDA(cudaEventElapsedTime(&ms, e0, e1));
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    if (ms_per_iter) *ms_per_iter = ms / n_iters;
    if (launches) *launches = (g_launch_count - l0) / n_iters;
    return 0;
}

} // namespace b2
