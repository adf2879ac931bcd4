// kernels_sample.cu -- on-device sampling: penalties, then greedy | top-k, top-p, min-p, temperature, draw | mirostat v2.
// Chain order follows llama.cpp's sampler chain as gpt_sampler_init builds it (penalties -> temp <= 0 ? greedy :
// mirostat == 2 ? temperature -> mirostat_v2 : top_k -> top_p -> min_p -> temperature -> softmax -> draw); one CTA
// per logits row; the vocabulary row (<= 513 KB) is read from L2 a handful of times.
#include "kernels.h"
#include "model.h"
#include "quant.cuh"
#include "sampling.h"

namespace b2 {

constexpr int SAMPLE_THREADS = 1024;
constexpr int CAND_CAP = 2 * SAMPLE_MAX_CANDIDATES; // room for ties at the k-th value

__device__ __forceinline__ uint32_t float_key(float f) { // monotonic float -> uint
    const uint32_t u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float key_float(uint32_t k) {
    const uint32_t u = (k & 0x80000000u) ? (k & 0x7FFFFFFFu) : ~k;
    return __uint_as_float(u);
}
__device__ __forceinline__ uint64_t splitmix64(uint64_t x) {
    x += 0x9E3779B97F4A7C15ULL;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ULL;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBULL;
    return x ^ (x >> 31);
}

// llama_sampler_penalties_apply (llama.cpp src/llama-sampling.cpp @ d5cb868, un-vendored): for every distinct token
// of the window with count c: logit <= 0 ? logit * repeat : logit / repeat, then logit -= c * freq + presence.
// In place on the logits row; one thread per window entry, the first occurrence of a token does the update.
__global__ void __launch_bounds__(SAMPLE_MAX_PREV)
penalty_kernel(float *__restrict__ logits, int64_t stride, int n_vocab, const SampleSpec *__restrict__ specs) {
    const SampleSpec &sp = specs[blockIdx.x];
    if (!(sp.n_prev > 0 && (sp.repeat_penalty != 1.f || sp.frequency_penalty != 0.f || sp.presence_penalty != 0.f))) return;
    const int n = min(sp.n_prev, SAMPLE_MAX_PREV), j = threadIdx.x;
    if (j >= n) return;
    const int tok = sp.prev[j];
    if (tok < 0 || tok >= n_vocab || tok == sp.nl_token) return;
    int count = 0;
    for (int k = 0; k < n; k++) {
        if (sp.prev[k] == tok) {
            if (k < j) return; // not the first occurrence
            count++;
        }
    }
    float *x = logits + (int64_t)blockIdx.x * stride + tok;
    float v = *x;
    v = v <= 0.f ? v * sp.repeat_penalty : v / sp.repeat_penalty;
    v -= (float)count * sp.frequency_penalty + sp.presence_penalty;
    *x = v;
}

__global__ void __launch_bounds__(SAMPLE_THREADS)
sample_kernel(const float *__restrict__ logits, int64_t stride, int n_vocab, const SampleSpec *__restrict__ specs,
              int32_t *__restrict__ out, float *__restrict__ mu_out) {
    __shared__ uint32_t hist[256];
    __shared__ uint32_t s_prefix, s_remaining, s_count;
    __shared__ unsigned long long cand[CAND_CAP]; // (key << 32) | (~idx): sorting descending gives key desc, idx asc
    __shared__ float s_red[32];
    __shared__ int s_redi[32];
    __shared__ float s_scan[SAMPLE_THREADS];
    const SampleSpec &sp = specs[blockIdx.x];
    const float *x = logits + (int64_t)blockIdx.x * stride;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    // key 0 = "not a candidate": the banned token and every masked (-inf) logit (token_mask_kernel)
    auto key_at = [&](int i) -> uint32_t {
        const float v = x[i];
        return (i == sp.ban_token || v == -INFINITY) ? 0u : float_key(v);
    };
    if (mu_out && tid == 0) mu_out[blockIdx.x] = sp.mirostat_mu;

    // ---------------- greedy: first maximal element
    if (sp.temperature <= 0.f) {
        uint32_t best = 0;
        int bi = 0x7fffffff;
        for (int i = tid; i < n_vocab; i += SAMPLE_THREADS) {
            const uint32_t k = key_at(i);
            if (k > best) { best = k; bi = i; }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const uint32_t ok = __shfl_xor_sync(0xffffffffu, best, o);
            const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
            if (ok > best || (ok == best && oi < bi)) { best = ok; bi = oi; }
        }
        if (lane == 0) { s_red[warp] = __uint_as_float(best); s_redi[warp] = bi; }
        __syncthreads();
        if (warp == 0) {
            best = __float_as_uint(s_red[lane]);
            bi = s_redi[lane];
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                const uint32_t ok = __shfl_xor_sync(0xffffffffu, best, o);
                const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
                if (ok > best || (ok == best && oi < bi)) { best = ok; bi = oi; }
            }
            if (lane == 0) out[blockIdx.x] = bi == 0x7fffffff ? 0 : bi;
        }
        return;
    }

    const float u01 = (float)(splitmix64(sp.rng) >> 40) * (1.0f / 16777216.0f);
    const float inv_t = 1.0f / sp.temperature;
    int K = sp.top_k <= 0 ? n_vocab : min(sp.top_k, n_vocab);
    const bool chain = sp.mirostat != 1; // mirostat v1 replaces top-k .. min-p (gpt_sampler_init: temp -> mirostat)
    if (!chain) K = n_vocab;
    if (chain && sp.typical_p < 1.0f) K = min(K, 256); // the typicality selection below is O(n * kept) on one thread
    const bool full = chain && (K == n_vocab) && sp.top_p >= 1.0f && sp.min_p <= 0.0f && sp.tfs_z >= 1.0f && sp.typical_p >= 1.0f;

    // ---------------- mirostat v2 (llama_sampler_mirostat_v2_apply): soft-max of the temperature-scaled logits, drop
    // every token whose surprise -log2(p) exceeds mu (the most probable one always stays), renormalise, draw,
    // mu -= eta * (surprise of the drawn token - tau)
    if (sp.mirostat == 2) {
        float mx = -INFINITY;
        for (int i = tid; i < n_vocab; i += SAMPLE_THREADS)
            if (i != sp.ban_token) mx = fmaxf(mx, x[i]);
        mx = warp_max(mx);
        if (lane == 0) s_red[warp] = mx;
        __syncthreads();
        mx = s_red[0];
        for (int i = 1; i < SAMPLE_THREADS / 32; i++) mx = fmaxf(mx, s_red[i]);
        __syncthreads();
        const int per = (n_vocab + SAMPLE_THREADS - 1) / SAMPLE_THREADS;
        const int i0 = tid * per, i1 = min(n_vocab, i0 + per);
        float part = 0.f;
        for (int i = i0; i < i1; i++)
            if (i != sp.ban_token) part += expf((x[i] - mx) * inv_t);
        s_scan[tid] = part;
        __syncthreads();
        float tot = 0.f;
        for (int i = 0; i < SAMPLE_THREADS; i++) tot += s_scan[i]; // every thread: same order, same value
        __syncthreads();
        // keep token i iff -log2(e_i / tot) <= mu  <=>  (x_i - mx) / T >= ln(tot) - mu ln 2; e_max = 1 always stays
        const float thr = fminf(logf(tot) - sp.mirostat_mu * 0.69314718056f, 0.f);
        part = 0.f;
        for (int i = i0; i < i1; i++) {
            const float z = (x[i] - mx) * inv_t;
            if (i != sp.ban_token && z >= thr) part += expf(z);
        }
        s_scan[tid] = part;
        __syncthreads();
        if (tid == 0) {
            float tot2 = 0.f;
            for (int i = 0; i < SAMPLE_THREADS; i++) tot2 += s_scan[i];
            const float target = u01 * tot2;
            float acc = 0.f;
            int t = 0;
            for (; t < SAMPLE_THREADS - 1; t++) {
                if (acc + s_scan[t] > target) break;
                acc += s_scan[t];
            }
            int pick = -1;
            float e_pick = 1.f;
            const int j0 = t * per, j1 = min(n_vocab, j0 + per);
            for (int i = j0; i < j1; i++) {
                const float z = (x[i] - mx) * inv_t;
                if (i == sp.ban_token || z < thr) continue;
                const float e = expf(z);
                acc += e;
                pick = i;
                e_pick = e;
                if (acc > target) break;
            }
            if (pick < 0) { // rounding left the target past the last kept token of this chunk: take the global maximum
                for (int i = 0; i < n_vocab; i++)
                    if (i != sp.ban_token && x[i] == mx) { pick = i; break; }
                if (pick < 0) pick = 0;
                e_pick = 1.f;
            }
            out[blockIdx.x] = pick;
            const float surprise = -log2f(e_pick / tot2);
            if (mu_out) mu_out[blockIdx.x] = sp.mirostat_mu - sp.mirostat_eta * (surprise - sp.mirostat_tau);
        }
        return;
    }

    // ---------------- no truncation at all: inverse-CDF over the whole vocabulary, no sort needed
    if (full) {
        float mx = -INFINITY;
        for (int i = tid; i < n_vocab; i += SAMPLE_THREADS)
            if (i != sp.ban_token) mx = fmaxf(mx, x[i]);
        mx = warp_max(mx);
        if (lane == 0) s_red[warp] = mx;
        __syncthreads();
        mx = s_red[0];
        for (int i = 1; i < SAMPLE_THREADS / 32; i++) mx = fmaxf(mx, s_red[i]);
        __syncthreads();
        // contiguous chunk per thread so that the scan order is the index order
        const int per = (n_vocab + SAMPLE_THREADS - 1) / SAMPLE_THREADS;
        const int i0 = tid * per, i1 = min(n_vocab, i0 + per);
        float part = 0.f;
        for (int i = i0; i < i1; i++)
            if (i != sp.ban_token) part += expf((x[i] - mx) * inv_t);
        s_scan[tid] = part;
        __syncthreads();
        if (tid == 0) {
            float tot = 0.f;
            for (int i = 0; i < SAMPLE_THREADS; i++) tot += s_scan[i];
            const float target = u01 * tot;
            float acc = 0.f;
            int t = 0;
            for (; t < SAMPLE_THREADS - 1; t++) {
                if (acc + s_scan[t] > target) break;
                acc += s_scan[t];
            }
            int pick = -1;
            const int j0 = t * per, j1 = min(n_vocab, j0 + per);
            for (int i = j0; i < j1; i++) {
                if (i == sp.ban_token) continue;
                acc += expf((x[i] - mx) * inv_t);
                pick = i;
                if (acc > target) break;
            }
            if (pick < 0) pick = 0;
            out[blockIdx.x] = pick;
        }
        return;
    }

    // ---------------- radix select of the K-th largest key (K capped to what fits on chip)
    K = min(K, SAMPLE_MAX_CANDIDATES);
    if (tid == 0) { s_prefix = 0; s_remaining = (uint32_t)K; }
    __syncthreads();
    for (int pass = 0; pass < 4; pass++) {
        const int shift = 24 - 8 * pass;
        if (tid < 256) hist[tid] = 0;
        __syncthreads();
        const uint32_t prefix = s_prefix;
        const uint32_t mask = pass == 0 ? 0u : (0xFFFFFFFFu << (shift + 8));
        for (int i = tid; i < n_vocab; i += SAMPLE_THREADS) {
            const uint32_t k = key_at(i);
            if ((k & mask) == prefix) atomicAdd(&hist[(k >> shift) & 0xFF], 1u);
        }
        __syncthreads();
        if (tid == 0) {
            uint32_t rem = s_remaining;
            int b = 255;
            for (; b > 0; b--) {
                if (hist[b] >= rem) break;
                rem -= hist[b];
            }
            s_prefix = prefix | ((uint32_t)b << shift);
            s_remaining = rem;
        }
        __syncthreads();
    }
    const uint32_t kth = s_prefix; // key of the K-th largest element (0: fewer than K tokens are allowed at all)
    if (tid == 0) s_count = 0;
    __syncthreads();
    // candidates = every key above the K-th (fewer than K of them, any order: they are sorted below) ...
    for (int i = tid; i < n_vocab; i += SAMPLE_THREADS) {
        const uint32_t k = key_at(i);
        if (k > kth) {
            const uint32_t slot = atomicAdd(&s_count, 1u);
            if (slot < CAND_CAP) cand[slot] = ((unsigned long long)k << 32) | (uint32_t)(~(uint32_t)i);
        }
    }
    __syncthreads();
    // ... plus as many of the tokens tied AT the K-th key as still fit into K, lowest ids first.  A plain ">= kth"
    // gather overflows the candidate buffer when the tie is large (a grammar mask leaves fewer than K finite logits,
    // or many equal logits) and then keeps whichever tokens happened to arrive first.
    const int n_gt = min((int)s_count, CAND_CAP);
    const int need = K - n_gt;
    if (kth != 0u && need > 0) {
        int *s_cnt = (int *)s_scan;
        const int per = (n_vocab + SAMPLE_THREADS - 1) / SAMPLE_THREADS;
        const int i0 = tid * per, i1 = min(n_vocab, i0 + per);
        int c = 0;
        for (int i = i0; i < i1; i++) c += key_at(i) == kth;
        // exclusive prefix over the threads: warp scan, then the warp totals
        int incl = c;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += t;
        }
        if (lane == 31) s_redi[warp] = incl;
        __syncthreads();
        int base = 0;
        for (int w = 0; w < warp; w++) base += s_redi[w];
        int rank = base + incl - c;
        for (int i = i0; i < i1 && rank < need; i++)
            if (key_at(i) == kth) {
                cand[n_gt + rank] = ((unsigned long long)kth << 32) | (uint32_t)(~(uint32_t)i);
                rank++;
            }
        if (tid == SAMPLE_THREADS - 1) s_cnt[0] = min(need, base + incl); // ties taken
        __syncthreads();
        if (tid == 0) s_count = (uint32_t)(n_gt + s_cnt[0]);
    }
    __syncthreads();
    const int n_cand = min((int)s_count, CAND_CAP);
    for (int i = n_cand + tid; i < CAND_CAP; i += SAMPLE_THREADS) cand[i] = 0ULL;
    __syncthreads();
    // bitonic sort, descending, CAND_CAP elements, 2 per thread
    for (int k = 2; k <= CAND_CAP; k <<= 1) {
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int t = tid; t < CAND_CAP; t += SAMPLE_THREADS) {
                const int ixj = t ^ j;
                if (ixj > t) {
                    const unsigned long long a = cand[t], b = cand[ixj];
                    const bool desc = (t & k) == 0;
                    if (desc ? (a < b) : (a > b)) { cand[t] = b; cand[ixj] = a; }
                }
            }
            __syncthreads();
        }
    }
    if (tid == 0) {
        int n = min(n_cand, K);
        if (n <= 0) { out[blockIdx.x] = 0; return; }
        const float l0 = key_float((uint32_t)(cand[0] >> 32));
        auto logit = [&](int i) { return key_float((uint32_t)(cand[i] >> 32)); };
        if (!chain) {
            // ---- mirostat v1 (llama_sampler_mirostat_apply): estimate the Zipf exponent s_hat from the m = 100 most
            // probable tokens, turn mu into a top-k cut, draw, mu -= eta (surprise - tau).  Candidates beyond the
            // SAMPLE_MAX_CANDIDATES most probable are not considered (k is clamped to that).
            float tot = 0.f;
            for (int i = 0; i < n; i++) tot += expf((logit(i) - l0) * inv_t);
            const int m = min(100, n);
            float sum_ti_bi = 0.f, sum_ti_sq = 0.f;
            for (int i = 0; i < m - 1; i++) {
                const float t_i = logf((float)(i + 2) / (float)(i + 1));
                const float b_i = (logit(i) - logit(i + 1)) * inv_t; // log(p_i / p_{i+1})
                sum_ti_bi += t_i * b_i;
                sum_ti_sq += t_i * t_i;
            }
            const float s_hat = sum_ti_sq > 0.f ? sum_ti_bi / sum_ti_sq : 1.f;
            const float eps_hat = s_hat - 1.f;
            const float kf = powf((eps_hat * powf(2.f, sp.mirostat_mu)) / (1.f - powf((float)n_vocab, -eps_hat)), 1.f / s_hat);
            int k = (kf == kf && kf < (float)n) ? max((int)kf, 1) : n; // NaN / huge -> no cut
            k = min(k, n);
            float tot_k = 0.f;
            for (int i = 0; i < k; i++) tot_k += expf((logit(i) - l0) * inv_t);
            const float target = u01 * tot_k;
            float acc = 0.f, e_pick = 1.f;
            int pick = k - 1;
            for (int i = 0; i < k; i++) {
                const float e = expf((logit(i) - l0) * inv_t);
                acc += e;
                if (acc > target) { pick = i; e_pick = e; break; }
                e_pick = e;
            }
            out[blockIdx.x] = (int32_t)(~(uint32_t)(cand[pick] & 0xFFFFFFFFu));
            const float surprise = -log2f(e_pick / tot_k);
            if (mu_out) mu_out[blockIdx.x] = sp.mirostat_mu - sp.mirostat_eta * (surprise - sp.mirostat_tau);
            return;
        }
        // tail-free sampling (llama_sampler_tail_free_apply) on the T=1 softmax: cut where the cumulated, normalised
        // |second derivative| of the sorted probabilities exceeds z
        if (sp.tfs_z < 1.0f && n > 2) {
            float tot = 0.f;
            for (int i = 0; i < n; i++) { s_scan[i] = expf(logit(i) - l0); tot += s_scan[i]; }
            float sum2 = 0.f;
            for (int i = 0; i < n - 2; i++) sum2 += fabsf(((s_scan[i] - s_scan[i + 1]) - (s_scan[i + 1] - s_scan[i + 2])) / tot);
            float cum = 0.f;
            int last = n;
            for (int i = 0; i < n - 2; i++) {
                const float d2 = fabsf(((s_scan[i] - s_scan[i + 1]) - (s_scan[i + 1] - s_scan[i + 2])) / tot);
                cum += sum2 > 1e-6f ? d2 / sum2 : 1.0f / (float)(n - 2);
                if (cum > sp.tfs_z && i >= 1) { last = i; break; }
            }
            n = last;
        }
        // locally typical sampling (llama_sampler_typical_apply): keep the tokens whose surprise is closest to the
        // entropy until their mass exceeds p; the survivors stay in logit order for the samplers that follow
        if (sp.typical_p < 1.0f && n > 1) {
            float tot = 0.f;
            for (int i = 0; i < n; i++) { s_scan[i] = expf(logit(i) - l0); tot += s_scan[i]; }
            float entropy = 0.f;
            for (int i = 0; i < n; i++) {
                const float pi = s_scan[i] / tot;
                s_scan[i] = pi;
                entropy += -pi * logf(pi);
            }
            // s_scan[i] < 0 marks "taken" (probabilities are > 0)
            float cum = 0.f;
            int kept = 0;
            while (kept < n) {
                int best = -1;
                float best_s = INFINITY;
                for (int i = 0; i < n; i++) {
                    if (s_scan[i] < 0.f) continue;
                    const float sh = fabsf(-logf(s_scan[i]) - entropy);
                    if (sh < best_s) { best_s = sh; best = i; }
                }
                if (best < 0) break;
                cum += s_scan[best];
                s_scan[best] = -s_scan[best];
                kept++;
                if (cum > sp.typical_p) break;
            }
            int w = 0;
            for (int i = 0; i < n; i++)
                if (s_scan[i] < 0.f) cand[w++] = cand[i];
            n = max(w, 1);
        }
        // top-p on the T=1 softmax of the surviving candidates
        const float l0p = logit(0);
        if (sp.top_p < 1.0f) {
            float tot = 0.f;
            for (int i = 0; i < n; i++) tot += expf(key_float((uint32_t)(cand[i] >> 32)) - l0p);
            float cum = 0.f;
            int keep = n;
            for (int i = 0; i < n; i++) {
                cum += expf(key_float((uint32_t)(cand[i] >> 32)) - l0p) / tot;
                if (cum >= sp.top_p) { keep = i + 1; break; }
            }
            n = keep;
        }
        // min-p: p_i >= min_p * p_max  <=>  logit_i >= logit_max + log(min_p)
        if (sp.min_p > 0.0f) {
            const float thr = l0p + logf(sp.min_p);
            int keep = 1;
            while (keep < n && key_float((uint32_t)(cand[keep] >> 32)) >= thr) keep++;
            n = keep;
        }
        // temperature, softmax, draw
        float tot = 0.f;
        for (int i = 0; i < n; i++) tot += expf((key_float((uint32_t)(cand[i] >> 32)) - l0) * inv_t);
        const float target = u01 * tot;
        float acc = 0.f;
        int pick = n - 1;
        for (int i = 0; i < n; i++) {
            acc += expf((key_float((uint32_t)(cand[i] >> 32)) - l0) * inv_t);
            if (acc > target) { pick = i; break; }
        }
        out[blockIdx.x] = (int32_t)(~(uint32_t)(cand[pick] & 0xFFFFFFFFu));
    }
}

// Greedy rows (temperature <= 0) without the single-CTA scan: the row is split over GREEDY_SPLIT CTAs that fold their
// slice into one 64-bit key (monotone value bits | ~index: max = largest value, smallest index among equals) and merge
// with atomicMax; the last CTA of the row writes the token and re-zeroes the scratch.  Same result as the greedy
// branch of sample_kernel (first maximal element, the banned token excluded).
constexpr int GREEDY_SPLIT = 32;
constexpr int GREEDY_MAX_ROWS = 4096;
static __device__ unsigned long long g_greedy_key[GREEDY_MAX_ROWS];
static __device__ unsigned int g_greedy_done[GREEDY_MAX_ROWS];

__global__ void __launch_bounds__(256)
greedy_kernel(const float *__restrict__ logits, int64_t stride, int n_vocab, const SampleSpec *__restrict__ specs,
              int32_t *__restrict__ out, float *__restrict__ mu_out) {
    const int row = blockIdx.y;
    const SampleSpec &sp = specs[row];
    const float *x = logits + (int64_t)row * stride;
    const int per = (n_vocab + GREEDY_SPLIT - 1) / GREEDY_SPLIT;
    const int i0 = blockIdx.x * per, i1 = min(n_vocab, i0 + per);
    unsigned long long best = 0ull;
    for (int i = i0 + threadIdx.x; i < i1; i += blockDim.x) {
        if (i == sp.ban_token) continue;
        const uint32_t k = float_key(x[i]);
        if (k != 0u) best = max(best, ((unsigned long long)k << 32) | (uint32_t)(~(uint32_t)i));
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) best = max(best, __shfl_xor_sync(0xffffffffu, best, o));
    __shared__ unsigned long long sv[8];
    if ((threadIdx.x & 31) == 0) sv[threadIdx.x >> 5] = best;
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < 8; w++) best = max(best, sv[w]);
        atomicMax(&g_greedy_key[row], best);
        __threadfence();
        if (atomicAdd(&g_greedy_done[row], 1u) == (unsigned)gridDim.x - 1) {
            __threadfence();
            const unsigned long long k = atomicExch(&g_greedy_key[row], 0ull);
            g_greedy_done[row] = 0u;
            out[row] = k == 0ull ? 0 : (int)(~(uint32_t)(k & 0xFFFFFFFFull));
            if (mu_out) mu_out[row] = sp.mirostat_mu;
        }
    }
}

void launch_sample(float *logits, int64_t stride, int n_rows, int n_vocab, const SampleSpec *specs_dev, bool any_penalties,
                   int32_t *out, float *mu_out, cudaStream_t s, bool all_greedy) {
    if (n_rows <= 0) return;
    if (any_penalties) {
        penalty_kernel<<<n_rows, SAMPLE_MAX_PREV, 0, s>>>(logits, stride, n_vocab, specs_dev);
        g_launch_count++;
        B2_CUDA(cudaGetLastError());
    }
    if (all_greedy && n_rows <= GREEDY_MAX_ROWS) greedy_kernel<<<dim3(GREEDY_SPLIT, n_rows), 256, 0, s>>>(logits, stride, n_vocab, specs_dev, out, mu_out);
    else sample_kernel<<<n_rows, SAMPLE_THREADS, 0, s>>>(logits, stride, n_vocab, specs_dev, out, mu_out);
    g_launch_count++;
    B2_CUDA(cudaGetLastError());
}

// logits of the tokens whose bit is clear go to -inf (the grammar leg of the resample, see grammar.h)
__global__ void token_mask_kernel(float *__restrict__ logits, int n_vocab, const uint32_t *__restrict__ bits) {
    const int w = blockIdx.x * blockDim.x + threadIdx.x; // one 32-token word per thread
    if (w * 32 >= n_vocab) return;
    const uint32_t b = bits[w];
    if (b == 0xFFFFFFFFu) return;
    const int n = min(32, n_vocab - w * 32);
    for (int i = 0; i < n; i++)
        if (!((b >> i) & 1u)) logits[w * 32 + i] = -INFINITY;
}

void resample_row_masked(Model &m, int row, const SampleSpec &spec, const uint32_t *mask_host, int32_t *out_host, float *mu_out_host) {
    const int n_vocab = m.hp.n_vocab, n_words = (n_vocab + 31) / 32;
    uint32_t *bits = m.token_mask_dev();
    B2_CUDA(cudaMemcpyAsync(bits, mask_host, sizeof(uint32_t) * n_words, cudaMemcpyHostToDevice, m.stream));
    float *logits = m.logits_dev_mut() + (size_t)row * n_vocab;
    token_mask_kernel<<<(n_words + 255) / 256, 256, 0, m.stream>>>(logits, n_vocab, bits);
    g_launch_count++;
    B2_CUDA(cudaGetLastError());
    SampleSpec *dev = m.sample_specs_dev(row + 1) + row;
    m.sample_specs_held().clear(); // (the device copy of this row changes under sample_rows' cache)
    B2_CUDA(cudaMemcpyAsync(dev, &spec, sizeof(SampleSpec), cudaMemcpyHostToDevice, m.stream));
    int32_t *out = const_cast<int32_t *>(m.next_tok_dev()) + row;
    launch_sample(logits, n_vocab, 1, n_vocab, dev, false, out, m.sample_mu_dev() + row, m.stream, spec.temperature <= 0.f);
    B2_CUDA(cudaMemcpyAsync(out_host, out, sizeof(int32_t), cudaMemcpyDeviceToHost, m.stream));
    if (mu_out_host) B2_CUDA(cudaMemcpyAsync(mu_out_host, m.sample_mu_dev() + row, sizeof(float), cudaMemcpyDeviceToHost, m.stream));
}

void sample_rows(Model &m, const std::vector<SampleSpec> &specs, int32_t *out_host, float *mu_out_host) {
    const int n = (int)specs.size();
    if (n == 0) return;
    bool any_pen = false, all_greedy = true;
    for (const auto &sp : specs) {
        any_pen |= sp.has_penalties();
        all_greedy &= sp.temperature <= 0.f;
    }
    SampleSpec *dev = m.sample_specs_dev(n);
    // A greedy step without penalties reads two fields of a row's spec (the banned token, the mirostat state it passes
    // through): when those are what the device already holds, the 584-byte upload per row is skipped -- in a b = 1 greedy
    // stream that is every step after the first.
    bool any_miro = false;
    for (const auto &sp : specs) any_miro |= sp.mirostat != 0;
    std::vector<std::pair<int32_t, float>> &held = m.sample_specs_held();
    bool same = all_greedy && !any_pen && !any_miro && (int)held.size() == n;
    for (int i = 0; same && i < n; i++) same = held[i].first == specs[i].ban_token && held[i].second == specs[i].mirostat_mu;
    if (!same) {
        B2_CUDA(cudaMemcpyAsync(dev, specs.data(), sizeof(SampleSpec) * n, cudaMemcpyHostToDevice, m.stream));
        held.clear();
        if (all_greedy && !any_pen && !any_miro)
            for (const auto &sp : specs) held.emplace_back(sp.ban_token, sp.mirostat_mu);
    }
    launch_sample(m.logits_dev_mut(), m.hp.n_vocab, n, m.hp.n_vocab, dev, any_pen, const_cast<int32_t *>(m.next_tok_dev()),
                  m.sample_mu_dev(), m.stream, all_greedy);
    B2_CUDA(cudaMemcpyAsync(out_host, m.next_tok_dev(), sizeof(int32_t) * n, cudaMemcpyDeviceToHost, m.stream));
    if (mu_out_host) {
        if (any_miro) B2_CUDA(cudaMemcpyAsync(mu_out_host, m.sample_mu_dev(), sizeof(float) * n, cudaMemcpyDeviceToHost, m.stream));
        else
            for (int i = 0; i < n; i++) mu_out_host[i] = specs[i].mirostat_mu; // without mirostat the state passes through unchanged
    }
}

} // namespace b2
