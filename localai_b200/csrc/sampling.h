// sampling.h -- on-device token sampling (greedy / temperature / top-k / top-p / min-p).
// Replaces gpt_sampler_sample as called at backend/cpp/llama/grpc-server.cpp:1946 (which runs on the CPU over
// logits copied to the host); here only the sampled id crosses PCIe.
#pragma once
#include <cstdint>
#include <vector>

#include "common.h"

namespace b2 {

class Model;

struct SampleSpec {
    float temperature = 0.f; // <= 0: greedy (first maximal logit)
    int32_t top_k = 0;       // <= 0: no top-k
    float top_p = 1.f;
    float min_p = 0.f;
    int32_t ban_token = -1;  // logit forced to -inf (ignore_eos)
    uint64_t rng = 0;        // per-call random state; the kernel derives a uniform from it
};

constexpr int SAMPLE_MAX_CANDIDATES = 1024; // top-k / nucleus candidates kept on chip

// Samples one token per row of the model's current logits ([specs.size()][n_vocab]); asynchronous on m.stream,
// results are copied to `out_host` (pinned or pageable) -- caller syncs the stream.
void sample_rows(Model &m, const std::vector<SampleSpec> &specs, int32_t *out_host);

void launch_sample(const float *logits, int64_t stride, int n_rows, int n_vocab, const SampleSpec *specs_dev, int32_t *out,
                   cudaStream_t s);

} // namespace b2
