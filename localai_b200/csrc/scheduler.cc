// scheduler.cc -- see scheduler.h.
#include "scheduler.h"

#include <algorithm>
#include <cstring>

#include "engine.h"
#include "sampling.h"
#include "tokenizer.h"

namespace b2 {

using clk = std::chrono::steady_clock;
static double ms_since(clk::time_point t0) { return std::chrono::duration<double, std::milli>(clk::now() - t0).count(); }

Scheduler::Scheduler(Engine *e) : eng_(e) {
    const int n = e->model().n_parallel;
    slots_.resize(n);
    for (int i = 0; i < n; i++) {
        slots_[i].id = i;
        slots_[i].t_last_used = clk::now();
    }
    thread_ = std::thread([this] { loop(); });
}

Scheduler::~Scheduler() {
    {
        std::lock_guard<std::mutex> lk(mu_);
        stop_ = true;
    }
    cv_.notify_all();
    if (thread_.joinable()) thread_.join();
    // wake any poller still blocked
    for (auto &kv : requests_) {
        std::lock_guard<std::mutex> lk(kv.second->mu);
        if (!kv.second->finished) {
            EventData ev;
            ev.flags = B2_EV_FINAL | B2_EV_ERROR;
            ev.text = "engine shut down";
            kv.second->events.push_back(std::move(ev));
            kv.second->finished = true;
        }
        kv.second->cv.notify_all();
    }
}

int64_t Scheduler::submit(const b2_request &r) {
    auto q = std::make_shared<Request>();
    if (r.prompt && r.prompt_len) q->prompt.assign(r.prompt, r.prompt_len);
    if (r.prompt_ids && r.n_prompt_ids > 0) {
        q->prompt_ids.assign(r.prompt_ids, r.prompt_ids + r.n_prompt_ids);
        q->has_ids = true;
    }
    q->n_predict = r.n_predict == 0 ? -1 : r.n_predict; // grpc-server.cpp:2098
    q->temperature = r.temperature;
    q->top_k = r.top_k;
    q->top_p = r.top_p > 0 ? r.top_p : 1.f;
    q->min_p = r.min_p;
    q->repeat_penalty = r.repeat_penalty > 0 ? r.repeat_penalty : 1.f;
    q->repeat_last_n = r.repeat_last_n;
    q->frequency_penalty = r.frequency_penalty;
    q->presence_penalty = r.presence_penalty;
    q->seed = r.seed;
    q->n_keep = r.n_keep;
    B2_CHECK(r.mirostat >= 0 && r.mirostat <= 2, B2_EINVAL, "mirostat must be 0 (off), 1 or 2");
    q->typical_p = (r.typical_p > 0.f && r.typical_p < 1.f) ? r.typical_p : 1.f;
    q->tfs_z = (r.tfs_z > 0.f && r.tfs_z < 1.f) ? r.tfs_z : 1.f;
    q->mirostat = r.mirostat;
    q->mirostat_tau = r.mirostat_tau > 0.f ? r.mirostat_tau : 5.f;
    q->mirostat_eta = r.mirostat_eta > 0.f ? r.mirostat_eta : 0.1f;
    q->ignore_eos = r.ignore_eos != 0;
    q->cache_prompt = r.cache_prompt != 0;
    q->embedding_mode = r.embedding_mode != 0;
    q->stream = r.stream != 0;
    q->penalize_nl = r.penalize_nl != 0; // grpc-server.cpp:2111
    if (q->embedding_mode) q->n_predict = 0;
    if (r.struct_size >= offsetof(b2_request, grammar) + sizeof(r.grammar) && r.grammar && r.grammar[0] && !q->embedding_mode)
        q->grammar = grammar_parse(r.grammar); // throws B2_EINVAL: the caller gets the parse error, nothing is queued
    for (int i = 0; i < r.n_stop; i++) {
        if (!r.stop || !r.stop[i]) continue;
        const size_t n = r.stop_len ? r.stop_len[i] : strlen(r.stop[i]);
        if (n) q->stop.emplace_back(r.stop[i], n);
    }
    {
        std::lock_guard<std::mutex> lk(mu_);
        B2_CHECK(!stop_, B2_ESTATE, "engine is shutting down");
        q->id = next_id_++;
        requests_[q->id] = q;
        pending_.push_back(q);
    }
    cv_.notify_all();
    return q->id;
}

int Scheduler::poll(int64_t id, b2_event *ev, int timeout_ms) {
    std::shared_ptr<Request> q;
    {
        std::lock_guard<std::mutex> lk(mu_);
        auto it = requests_.find(id);
        B2_CHECK(it != requests_.end(), B2_ESTATE, "unknown request id");
        q = it->second;
    }
    std::unique_lock<std::mutex> lk(q->mu);
    auto ready = [&] { return !q->events.empty(); };
    if (timeout_ms < 0) q->cv.wait(lk, ready);
    else if (!q->cv.wait_for(lk, std::chrono::milliseconds(timeout_ms), ready)) return B2_ETIMEOUT;
    EventData d = std::move(q->events.front());
    q->events.pop_front();
    q->last_text = std::move(d.text);
    const uint32_t sz = ev->struct_size ? std::min<uint32_t>(ev->struct_size, sizeof(b2_event)) : (uint32_t)sizeof(b2_event);
    b2_event out;
    memset(&out, 0, sizeof(out));
    out.struct_size = sz;
    out.flags = d.flags;
    out.text = q->last_text.data();
    out.text_len = q->last_text.size();
    out.n_prompt_tokens = d.n_prompt;
    out.n_decoded = d.n_decoded;
    out.slot_id = d.slot_id;
    out.n_embd = d.n_embd;
    out.t_prompt_ms = d.t_prompt_ms;
    out.t_gen_ms = d.t_gen_ms;
    memcpy(ev, &out, sz);
    return 0;
}

int Scheduler::cancel(int64_t id) {
    std::lock_guard<std::mutex> lk(mu_);
    auto it = requests_.find(id);
    B2_CHECK(it != requests_.end(), B2_ESTATE, "unknown request id");
    it->second->cancelled = true;
    cv_.notify_all();
    return 0;
}

int Scheduler::embed_result(int64_t id, float *out, int cap) {
    std::shared_ptr<Request> q;
    {
        std::lock_guard<std::mutex> lk(mu_);
        auto it = requests_.find(id);
        B2_CHECK(it != requests_.end(), B2_ESTATE, "unknown request id");
        q = it->second;
    }
    std::lock_guard<std::mutex> lk(q->mu);
    const int n = (int)q->embedding.size();
    if (n > cap) return -n;
    if (n) memcpy(out, q->embedding.data(), (size_t)n * 4);
    return n;
}

int Scheduler::release(int64_t id) {
    std::lock_guard<std::mutex> lk(mu_);
    auto it = requests_.find(id);
    B2_CHECK(it != requests_.end(), B2_ESTATE, "unknown request id");
    it->second->cancelled = true; // if still running, the engine thread drops it at the next step
    requests_.erase(it);
    return 0;
}

void Scheduler::metrics(b2_metrics_t *out) {
    std::lock_guard<std::mutex> lk(metrics_mu_);
    *out = metrics_;
}

void Scheduler::push_event(Request &r, EventData &&ev) {
    {
        std::lock_guard<std::mutex> lk(r.mu);
        if (ev.flags & B2_EV_FINAL) r.finished = true;
        r.events.push_back(std::move(ev));
    }
    r.cv.notify_all();
}

// ---------------------------------------------------------------------------------------------------------
void Scheduler::loop() {
    for (;;) {
        {
            std::unique_lock<std::mutex> lk(mu_);
            cv_.wait(lk, [&] {
                if (stop_ || !pending_.empty()) return true;
                for (auto &s : slots_)
                    if (s.state != IDLE) return true;
                return false;
            });
            if (stop_) return;
        }
        try {
            step();
        } catch (const std::exception &e) { // a device error must not wedge pollers: fail every active request
            for (auto &s : slots_)
                if (s.state != IDLE) finish_slot(s, true, e.what());
        }
    }
}

// LRU free slot (get_slot, grpc-server.cpp:609-628)
Scheduler::Slot *Scheduler::get_free_slot() {
    Slot *best = nullptr;
    for (auto &s : slots_)
        if (s.state == IDLE && (!best || s.t_last_used < best->t_last_used)) best = &s;
    return best;
}

const GrammarVocab &Scheduler::grammar_vocab() {
    if (!grammar_vocab_) { // pieces as llama_grammar_apply sees them: special tokens rendered, end-of-generation apart
        const Tokenizer &tk = *eng_->tokenizer();
        auto v = std::make_unique<GrammarVocab>();
        const int n = tk.n_vocab();
        v->cps.resize(n);
        v->tails.resize(n);
        v->is_eog.resize(n);
        v->is_empty.resize(n);
        v->pieces.resize(n);
        for (int t = 0; t < n; t++) {
            v->pieces[t] = tk.token_to_piece(t, true);
            v->is_eog[t] = tk.is_eog(t);
            v->is_empty[t] = v->pieces[t].empty() || v->pieces[t][0] == '\0';
            Utf8Tail tail;
            utf8_decode_piece(v->pieces[t], tail, v->cps[t]);
            v->tails[t] = tail;
        }
        grammar_vocab_ = std::move(v);
    }
    return *grammar_vocab_;
}

bool Scheduler::has_budget(const Slot &s) const { // grpc-server.cpp:260-278
    if (s.req->n_predict == -1) return true;
    return s.req->n_predict - s.n_decoded > 0;
}

void Scheduler::launch_slot(Slot &s, std::shared_ptr<Request> r) {
    Model &m = eng_->model();
    const int n_ctx = m.n_ctx_slot;
    s.req = r;
    s.state = PROMPT;
    s.n_decoded = 0;
    s.out.reset(&r->stop);
    s.has_next_token = true;
    s.stopped_eos = s.stopped_word = s.stopped_limit = s.truncated = false;
    s.t_start_prompt = clk::now();
    s.t_prompt_ms = s.t_gen_ms = 0;
    s.mirostat_mu = 2.f * r->mirostat_tau; // llama_sampler_init_mirostat_v2: mu = 2 * tau
    s.grammar.reset(r->grammar ? new Grammar(r->grammar) : nullptr); // the prompt is not run through it (grpc-server.cpp:1734)
    s.mask_cache.clear();
    s.rng_state = r->seed < 0 ? (uint64_t)std::chrono::high_resolution_clock::now().time_since_epoch().count() : (uint64_t)(uint32_t)r->seed;
    // tokenize (BOS iff the vocabulary asks for it; special tokens parsed: grpc-server.cpp:564,1689)
    std::vector<int32_t> toks;
    if (r->has_ids) toks = r->prompt_ids;
    else if (!r->prompt.empty()) toks = eng_->tokenizer()->tokenize(r->prompt, true, true);
    for (int32_t t : toks) B2_CHECK(t >= 0 && t < m.hp.n_vocab, B2_EINVAL, "prompt token id out of range");
    s.n_prompt = (int)toks.size();
    const int n_keep = effective_n_keep(r->n_keep, s.n_prompt, n_ctx);
    if (fit_prompt_to_context(toks, n_ctx, n_keep)) { // (stream_text.h; the reference's rule, grpc-server.cpp:1700-1721)
        s.truncated = true;
        s.n_prompt = (int)toks.size();
    }
    s.n_past = r->cache_prompt ? reusable_prefix(s.cache_tokens, toks) : 0; // in-memory prefix reuse (grpc-server.cpp:1740-1749)
    s.n_prompt_processed = s.n_prompt - s.n_past;
    s.cache_tokens = toks;
    s.prompt_tokens.swap(toks);
    if (s.n_past == s.n_prompt && s.n_past > 0) s.n_past--; // at least one token must be evaluated (after the metric above, as in the reference)
}

void Scheduler::finish_slot(Slot &s, bool error, const std::string &msg) {
    if (s.n_decoded > 0) s.t_gen_ms = ms_since(s.t_start_gen);
    EventData ev;
    ev.flags = B2_EV_FINAL | (error ? B2_EV_ERROR : 0) | (s.stopped_eos ? B2_EV_STOP_EOS : 0) | (s.stopped_word ? B2_EV_STOP_WORD : 0) |
               (s.stopped_limit ? B2_EV_STOP_LIMIT : 0);
    if (error) ev.text = msg;
    else if (!s.req->stream) ev.text = s.out.text(); // send_final_response: content only when not streaming
    ev.n_prompt = s.n_prompt;
    ev.n_decoded = s.n_decoded;
    ev.slot_id = s.id;
    ev.t_prompt_ms = s.t_prompt_ms;
    ev.t_gen_ms = s.t_gen_ms;
    {
        std::lock_guard<std::mutex> lk(metrics_mu_);
        metrics_.slot_id = s.id;
        metrics_.tokens_generated = s.n_decoded;
        metrics_.prompt_tokens_processed = s.n_prompt_processed;
        metrics_.tokens_per_second = s.t_gen_ms > 0 ? (float)(1e3 / s.t_gen_ms * s.n_decoded) : 0.f;
        metrics_.n_tokens_decoded_total += s.n_decoded;
        metrics_.n_prompt_tokens_total += s.n_prompt_processed;
    }
    push_event(*s.req, std::move(ev));
    s.req.reset();
    s.state = IDLE;
    s.t_last_used = clk::now();
}

// The token just sampled for a slot: its bytes go through the slot's StreamText (UTF-8 hold-back, stop words), then the
// budget and end-of-sequence rules decide whether the slot goes on (the behaviour of grpc-server.cpp:1010-1123).
bool Scheduler::process_token(Slot &s, int32_t tok) {
    Request &r = *s.req;
    s.sampled = tok;
    StreamText::Step step = s.out.append(eng_->tokenizer()->token_to_piece(tok, true));
    bool go_on = !step.stop_word;
    s.stopped_word = s.stopped_word || step.stop_word;
    if (r.stream && !step.release.empty()) {
        EventData ev;
        ev.text = std::move(step.release);
        ev.n_prompt = s.n_prompt;
        ev.n_decoded = s.n_decoded;
        ev.slot_id = s.id;
        push_event(r, std::move(ev));
    }
    if (go_on && s.n_decoded > 0 && !has_budget(s)) {
        s.stopped_limit = true;
        go_on = false;
    }
    if (tok == eng_->tokenizer()->eos_id()) { // EOS only; EOT is the caller's stop word (grpc-server.cpp:1103)
        s.stopped_eos = true;
        go_on = false;
    }
    s.has_next_token = go_on;
    return go_on;
}

// one update_slots() (grpc-server.cpp:1546-1982)
bool Scheduler::step() {
    Model &m = eng_->model();
    // cancellations and context exhaustion
    for (auto &s : slots_) {
        if (s.state == IDLE) continue;
        if (s.req->cancelled) {
            s.stopped_limit = true;
            finish_slot(s);
            continue;
        }
        if (s.state == GENERATING && (int)s.cache_tokens.size() >= m.n_ctx_slot) { // context-shift is disabled: release
            s.stopped_limit = true;
            finish_slot(s);
            s.cache_tokens.clear();
            s.n_past = 0;
        }
    }
    // bind queued requests to free slots; requests that find none stay queued (deferred, grpc-server.cpp:1472-1478)
    for (;;) {
        std::shared_ptr<Request> r;
        Slot *s = get_free_slot();
        if (!s) break;
        {
            std::lock_guard<std::mutex> lk(mu_);
            if (pending_.empty()) break;
            r = pending_.front();
            pending_.pop_front();
        }
        if (r->cancelled) {
            EventData ev;
            ev.flags = B2_EV_FINAL | B2_EV_STOP_LIMIT;
            push_event(*r, std::move(ev));
            continue;
        }
        try {
            launch_slot(*s, r);
        } catch (const std::exception &e) {
            s->req = r;
            s->n_prompt = 0;
            finish_slot(*s, true, e.what());
            continue;
        }
        if (s->n_prompt == 0) { // empty prompt -> empty final response (grpc-server.cpp:1645-1651)
            finish_slot(*s);
        }
    }
    // assemble the batch: one token per generating slot, then all pending prompt tokens
    struct Row { int32_t tok, pos, slot; bool want; };
    std::vector<Row> rows;
    std::vector<std::pair<int, int>> out_of; // (slot index, row index) for rows that want logits
    for (auto &s : slots_) {
        if (s.state != GENERATING) continue;
        out_of.emplace_back(s.id, (int)rows.size());
        rows.push_back({s.sampled, s.n_past, s.id, true});
        s.n_past += 1;
    }
    for (auto &s : slots_) {
        if (s.state != PROMPT) continue;
        for (; s.n_past < s.n_prompt; s.n_past++) rows.push_back({s.prompt_tokens[s.n_past], s.n_past, s.id, false});
        rows.back().want = true; // logits only for the last prompt token
        out_of.emplace_back(s.id, (int)rows.size() - 1);
        s.n_decoded = 0;
    }
    if (rows.empty()) return false;

    std::lock_guard<std::mutex> dev(eng_->device_mutex());
    const int nb = m.n_batch;
    std::vector<int32_t> tok(nb), pos(nb), sl(nb), orow(nb), sampled(nb);
    for (int i0 = 0; i0 < (int)rows.size(); i0 += nb) {
        const int c = std::min(nb, (int)rows.size() - i0);
        int n_out = 0;
        std::vector<int> out_slot;
        std::vector<SampleSpec> specs;
        for (int i = 0; i < c; i++) {
            const Row &r = rows[i0 + i];
            tok[i] = r.tok; pos[i] = r.pos; sl[i] = r.slot;
            if (r.want) {
                orow[n_out++] = i;
                out_slot.push_back(r.slot);
                Slot &s = slots_[r.slot];
                SampleSpec sp;
                sp.temperature = s.req->temperature;
                sp.top_k = s.req->top_k;
                sp.top_p = s.req->top_p;
                sp.min_p = s.req->min_p;
                sp.ban_token = s.req->ignore_eos ? eng_->tokenizer()->eos_id() : -1;
                sp.rng = s.rng_state;
                s.rng_state = s.rng_state * 6364136223846793005ULL + 1442695040888963407ULL;
                // penalties over the last repeat_last_n tokens of prompt + generation (gpt_sampler_accept is fed both,
                // grpc-server.cpp:1737,1948); -1 = the whole context, clamped to the window the kernel keeps
                sp.repeat_penalty = s.req->repeat_penalty;
                sp.frequency_penalty = s.req->frequency_penalty;
                sp.presence_penalty = s.req->presence_penalty;
                sp.nl_token = s.req->penalize_nl ? -1 : eng_->tokenizer()->nl_id(); // the newline is exempt unless PenalizeNL
                {
                    const int hist = (int)s.cache_tokens.size();
                    int last_n = s.req->repeat_last_n < 0 ? hist : s.req->repeat_last_n;
                    last_n = std::min(std::min(last_n, hist), SAMPLE_MAX_PREV);
                    sp.n_prev = last_n;
                    for (int k = 0; k < last_n; k++) sp.prev[k] = s.cache_tokens[hist - last_n + k];
                }
                sp.tfs_z = s.req->tfs_z;
                sp.typical_p = s.req->typical_p;
                sp.mirostat = s.req->mirostat;
                sp.mirostat_tau = s.req->mirostat_tau;
                sp.mirostat_eta = s.req->mirostat_eta;
                sp.mirostat_mu = s.mirostat_mu;
                specs.push_back(sp);
            }
        }
        Batch b;
        b.n = c;
        b.tokens = tok.data(); b.pos = pos.data(); b.slot = sl.data();
        b.out_rows = orow.data();
        b.n_out = n_out;
        m.forward(b);
        if (n_out == 0) {
            m.sync();
            continue;
        }
        // embeddings first (they only need the final-norm rows), then sampling on device
        std::vector<float> embd;
        bool any_embd = false;
        for (int k = 0; k < n_out; k++) any_embd |= slots_[out_slot[k]].req->embedding_mode;
        if (any_embd) {
            embd.resize((size_t)n_out * m.hp.n_embd);
            B2_CUDA(cudaMemcpyAsync(embd.data(), m.embd_dev(), embd.size() * 4, cudaMemcpyDeviceToHost, m.stream));
        }
        std::vector<float> mu_new(specs.size());
        sample_rows(m, specs, sampled.data(), mu_new.data());
        m.sync();
        // grammar: keep the draw if it fits, else mask the vocabulary and draw again from the same logits -- the order
        // gpt_sampler_sample uses (grpc-server.cpp:1946), so an unconstrained step costs one host-side check
        for (int k = 0; k < n_out; k++) {
            Slot &s = slots_[out_slot[k]];
            if (s.state == IDLE || !s.grammar || s.req->embedding_mode) continue;
            const GrammarVocab &gv = grammar_vocab();
            if (s.grammar->token_fits(gv, sampled[k])) continue;
            if (s.mask_cache.size() >= 256) s.mask_cache.clear();
            auto ins = s.mask_cache.try_emplace(s.grammar->state_key());
            std::vector<uint32_t> &mask = ins.first->second;
            if (ins.second) {
                mask.resize((gv.cps.size() + 31) / 32);
                s.grammar->token_mask(gv, mask.data());
            }
            SampleSpec sp = specs[k];
            sp.n_prev = 0; // the row already carries the penalties of the first draw
            sp.mirostat_mu = mu_new[k];
            sp.rng = s.rng_state;
            s.rng_state = s.rng_state * 6364136223846793005ULL + 1442695040888963407ULL;
            resample_row_masked(m, k, sp, mask.data(), &sampled[k], &mu_new[k]);
            m.sync();
            if (!s.grammar->token_fits(gv, sampled[k])) // nothing fits (dead end, or only EOS fits and it is banned): stop
                sampled[k] = eng_->tokenizer()->eos_id();
        }
        {
            std::lock_guard<std::mutex> lk(metrics_mu_);
            metrics_.n_decode_steps++;
            int busy = 0;
            for (auto &s : slots_) busy += s.state != IDLE;
            metrics_.n_slots_busy = busy;
        }
        for (int k = 0; k < n_out; k++) {
            Slot &s = slots_[out_slot[k]];
            if (s.state == IDLE) continue;
            if (s.req->embedding_mode) { // send_embedding (grpc-server.cpp:1292-1321)
                Request &r = *s.req;
                {
                    std::lock_guard<std::mutex> lk(r.mu);
                    if (eng_->embeddings_enabled()) r.embedding.assign(embd.begin() + (size_t)k * m.hp.n_embd, embd.begin() + (size_t)(k + 1) * m.hp.n_embd);
                    else r.embedding.assign(m.hp.n_embd, 0.0f);
                }
                s.t_prompt_ms = ms_since(s.t_start_prompt);
                EventData ev;
                ev.flags = B2_EV_FINAL | B2_EV_EMBEDDING;
                ev.n_prompt = s.n_prompt;
                ev.slot_id = s.id;
                ev.n_embd = m.hp.n_embd;
                ev.t_prompt_ms = s.t_prompt_ms;
                push_event(r, std::move(ev));
                s.req.reset();
                s.state = IDLE;
                s.t_last_used = clk::now();
                continue;
            }
            const int32_t id = sampled[k];
            s.mirostat_mu = mu_new[k];
            if (s.state == PROMPT) s.state = GENERATING;
            s.n_decoded += 1;
            if (s.n_decoded == 1) {
                s.t_start_gen = clk::now();
                s.t_prompt_ms = ms_since(s.t_start_prompt);
            }
            s.cache_tokens.push_back(id);
            if (s.grammar && s.grammar->token_fits(grammar_vocab(), id)) s.grammar->accept_token(grammar_vocab(), id);
            if (!process_token(s, id)) finish_slot(s);
        }
    }
    return true;
}

} // namespace b2
