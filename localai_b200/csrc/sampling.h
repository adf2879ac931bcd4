// sampling.h -- on-device token sampling: penalties -> greedy | top-k / top-p / min-p / temperature / draw | mirostat v2.
// Replaces gpt_sampler_sample as called at backend/cpp/llama/grpc-server.cpp:1946 (which runs on the CPU over
// logits copied to the host); here only the sampled id crosses PCIe.
#pragma once
#include <cstdint>
#include <vector>

#include "common.h"

namespace b2 {

class Model;

constexpr int SAMPLE_MAX_PREV = 128; // penalty window kept per row (repeat_last_n is clamped to it)

struct SampleSpec {
    float temperature = 0.f; // <= 0: greedy (first maximal logit)
    int32_t top_k = 0;       // <= 0: no top-k
    float top_p = 1.f;
    float min_p = 0.f;
    int32_t ban_token = -1;  // logit forced to -inf (ignore_eos)
    uint64_t rng = 0;        // per-call random state; the kernel derives a uniform from it
    // penalties over the last n_prev tokens (llama_sampler_penalties: repeat, frequency, presence; the newline token
    // is exempt, penalize_nl = false as the reference leaves it, grpc-server.cpp:640-647)
    float repeat_penalty = 1.f, frequency_penalty = 0.f, presence_penalty = 0.f;
    int32_t n_prev = 0;
    int32_t nl_token = -1;
    float tfs_z = 1.f, typical_p = 1.f; // tail-free / locally typical sampling; >= 1 = off
    // mirostat (0 = off, 1 = v1, 2 = v2; LocalAI's default config sends 2, core/config/backend_config.go:286-289)
    int32_t mirostat = 0;
    float mirostat_tau = 5.f, mirostat_eta = 0.1f;
    float mirostat_mu = 10.f; // state carried by the caller from draw to draw (starts at 2 * tau)
    int32_t prev[SAMPLE_MAX_PREV];
    bool has_penalties() const {
        return n_prev > 0 && (repeat_penalty != 1.f || frequency_penalty != 0.f || presence_penalty != 0.f);
    }
};

constexpr int SAMPLE_MAX_CANDIDATES = 1024; // top-k / nucleus candidates kept on chip

// Samples one token per row of the model's current logits ([specs.size()][n_vocab]); asynchronous on m.stream,
// results are copied to `out_host` (pinned or pageable) -- caller syncs the stream.
// mu_out_host (optional): the mirostat state after the draw, per row (unchanged value for rows without mirostat).
// NOTE: rows with penalties have their logits modified in place.
void sample_rows(Model &m, const std::vector<SampleSpec> &specs, int32_t *out_host, float *mu_out_host = nullptr);

// Grammar resample of ONE row of the same logits: tokens whose bit in `mask_host` ((n_vocab + 31) / 32 words) is clear
// are set to -inf in place, then the row is sampled again with `spec` (penalties are NOT applied a second time: the
// first draw already modified the row).  Asynchronous on m.stream like sample_rows.
void resample_row_masked(Model &m, int row, const SampleSpec &spec, const uint32_t *mask_host, int32_t *out_host,
                         float *mu_out_host = nullptr);

void launch_sample(float *logits, int64_t stride, int n_rows, int n_vocab, const SampleSpec *specs_dev, bool any_penalties,
                   int32_t *out, float *mu_out, cudaStream_t s, bool all_greedy = false);

} // namespace b2
