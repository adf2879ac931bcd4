// scheduler.h -- slots + continuous batching + result queues: the C++ engine loop.
// Replaces llama_server_context::{get_slot, launch_slot_with_data, update_slots, process_token,
// send_partial_response, send_final_response, send_embedding} (backend/cpp/llama/grpc-server.cpp:609-628,
// 630-904, 1546-1982, 1010-1123, 1200-1321) and llama_server_queue / llama_server_response
// (backend/cpp/llama/utils.hpp:192-409).  One engine thread owns the device; RPC threads enqueue and block.
#pragma once
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <deque>
#include <map>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "common.h"
#include "grammar.h"

namespace b2 {

class Engine;

struct EventData {
    uint32_t flags = 0;
    std::string text;
    int n_prompt = 0, n_decoded = 0, slot_id = -1, n_embd = 0;
    double t_prompt_ms = 0, t_gen_ms = 0;
};

struct Request {
    int64_t id = 0;
    // deep copy of b2_request
    std::string prompt;
    std::vector<int32_t> prompt_ids;
    bool has_ids = false;
    int n_predict = -1;
    float temperature = 0.f, top_p = 1.f, min_p = 0.f, repeat_penalty = 1.f, frequency_penalty = 0.f, presence_penalty = 0.f;
    int top_k = 0, repeat_last_n = 0, seed = -1, n_keep = 0;
    int mirostat = 0;
    float mirostat_tau = 5.f, mirostat_eta = 0.1f, typical_p = 1.f, tfs_z = 1.f;
    bool ignore_eos = false, cache_prompt = false, embedding_mode = false, stream = true, penalize_nl = false;
    std::vector<std::string> stop;
    std::shared_ptr<const GrammarRules> grammar; // parsed at submit; null = unconstrained
    // results
    std::mutex mu;
    std::condition_variable cv;
    std::deque<EventData> events;
    std::string last_text; // keeps the bytes of the event handed out by the last poll alive
    std::vector<float> embedding;
    std::atomic<bool> cancelled{false};
    bool finished = false;
};

class Scheduler {
  public:
    explicit Scheduler(Engine *e);
    ~Scheduler();

    int64_t submit(const b2_request &r);
    int poll(int64_t id, b2_event *ev, int timeout_ms);
    int cancel(int64_t id);
    int embed_result(int64_t id, float *out, int cap);
    int release(int64_t id);
    void metrics(b2_metrics_t *out);

  private:
    enum SlotState { IDLE, PROMPT, GENERATING };
    struct Slot {
        int id = 0;
        SlotState state = IDLE;
        std::shared_ptr<Request> req;
        std::vector<int32_t> cache_tokens; // tokens whose KV is (or is about to be) in the cache
        std::vector<int32_t> prompt_tokens;
        int n_past = 0, n_prompt = 0, n_prompt_processed = 0, n_decoded = 0;
        int32_t sampled = 0;
        std::string generated_text;
        size_t sent_count = 0;
        bool has_next_token = true, stopped_eos = false, stopped_word = false, stopped_limit = false, truncated = false;
        std::chrono::steady_clock::time_point t_last_used, t_start_prompt, t_start_gen;
        double t_prompt_ms = 0, t_gen_ms = 0;
        uint64_t rng_state = 0;
        float mirostat_mu = 10.f; // mirostat state of the request in the slot (2 * tau at launch)
        std::unique_ptr<Grammar> grammar; // matcher state of the request in the slot
        // token masks by matcher state: inside a JSON string, between digits ... the same state comes back token after
        // token, and a mask costs one pass over the vocabulary
        std::map<std::string, std::vector<uint32_t>> mask_cache;
    };

    void loop();
    bool step(); // one update_slots(); returns false when idle
    Slot *get_free_slot();
    void launch_slot(Slot &s, std::shared_ptr<Request> r);
    bool process_token(Slot &s, int32_t tok);
    void finish_slot(Slot &s, bool error = false, const std::string &msg = "");
    void push_event(Request &r, EventData &&ev);
    bool has_budget(const Slot &s) const;
    const GrammarVocab &grammar_vocab();

    Engine *eng_;
    std::thread thread_;
    std::mutex mu_; // protects pending_, requests_, stop_
    std::condition_variable cv_;
    std::deque<std::shared_ptr<Request>> pending_;
    std::map<int64_t, std::shared_ptr<Request>> requests_;
    std::vector<Slot> slots_;
    bool stop_ = false;
    int64_t next_id_ = 1;
    std::unique_ptr<GrammarVocab> grammar_vocab_; // built on the first grammar request (engine thread only)
    // metrics
    std::mutex metrics_mu_;
    b2_metrics_t metrics_{};
};

} // namespace b2
