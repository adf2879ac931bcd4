// kernels_misc.cu -- RMSNorm, SwiGLU, RoPE + paged-KV store, paged-KV attention, argmax.
// Numerics follow the ggml CPU operators the reference executes (see oracle/ggml_ref.c for the restatement):
// f32 everywhere, F16 KV cache, q and the soft-max probabilities rounded to f16 before their dot products.
#include "attention.cuh"
#include "kernels.h"
#include "quant.cuh"

namespace b2 {

bool pdl_enabled() { // B200LLM_PDL=1; read at every launch on purpose: the A/B test flips it inside one process.
    // Off by default: measured on a 64-stream decode step of Llama-3-8B, 5.00 ms with it against 4.79 ms without.
    const char *e = getenv("B200LLM_PDL");
    return e && atoi(e);
}

// ------------------------------------------------------------------------------------------ RMSNorm (* weight)
__global__ void __launch_bounds__(256)
rmsnorm_kernel(const float *__restrict__ x, int64_t x_stride, const int32_t *__restrict__ rows, const float *__restrict__ w,
               float *__restrict__ y, int64_t y_stride, int d, float eps) {
    B2_PDL_TRIGGER();
    B2_PDL_WAIT();
    const float *xr = x + (int64_t)(rows ? rows[blockIdx.x] : blockIdx.x) * x_stride;
    float *yr = y + (int64_t)blockIdx.x * y_stride;
    double sum = 0.0;
    for (int i = threadIdx.x; i < d; i += blockDim.x) {
        const float v = xr[i];
        sum += (double)__fmul_rn(v, v);
    }
    __shared__ double part[8];
    __shared__ float s_scale;
    sum = warp_sum_d(sum);
    if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = sum;
    __syncthreads();
    if (threadIdx.x == 0) {
        double tot = 0.0;
        for (int i = 0; i < (int)(blockDim.x >> 5); i++) tot += part[i];
        const float mean = (float)(tot / (double)d);
        s_scale = __fdiv_rn(1.0f, __fsqrt_rn(__fadd_rn(mean, eps)));
    }
    __syncthreads();
    const float scale = s_scale;
    for (int i = threadIdx.x; i < d; i += blockDim.x) yr[i] = __fmul_rn(__fmul_rn(xr[i], scale), w[i]);
}

void launch_rmsnorm(const float *x, int64_t x_stride, const int32_t *rows, const float *w, float *y, int64_t y_stride, int n_tok,
                    int d, float eps, cudaStream_t s) {
    if (n_tok <= 0) return;
    launch_pdl(rmsnorm_kernel, dim3(n_tok), dim3(256), 0, s, x, x_stride, rows, w, y, y_stride, d, eps);
    g_launch_count++;
    B2_CUDA(cudaGetLastError());
}

// ------------------------------------------------------------------------------------------ SwiGLU
__global__ void swiglu_kernel(const float *__restrict__ g, const float *__restrict__ u, float *__restrict__ out, int64_t n) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const float x = g[i];
        const float s = __fdiv_rn(x, __fadd_rn(1.0f, expf(-x)));
        out[i] = __fmul_rn(s, u[i]);
    }
}

void launch_swiglu(const float *gate, const float *up, float *out, int64_t n, cudaStream_t s) {
    if (n <= 0) return;
    const int blocks = (int)std::min<int64_t>((n + 255) / 256, 148 * 8);
    swiglu_kernel<<<blocks, 256, 0, s>>>(gate, up, out, n);
    g_launch_count++;
    B2_CUDA(cudaGetLastError());
}

// ------------------------------------------------------------------------------------------ RoPE (NORM pairs) + KV store
// bq / bk / bv (optional, qwen2): added to the projections before the rotation (ggml_add(Qcur, bq) in llm_build_qwen2);
// neox: the rotated pairs are (i, i + hd / 2) (LLAMA_ROPE_TYPE_NEOX) instead of (2i, 2i + 1)
__global__ void rope_kv_store_kernel(float *__restrict__ q, float *__restrict__ k, const float *__restrict__ v,
                                     const int32_t *__restrict__ pos, const int32_t *__restrict__ slot, int n_head, int n_head_kv,
                                     int hd, const float *__restrict__ rope_table, KVLayout kv, const float *__restrict__ bq,
                                     const float *__restrict__ bk, const float *__restrict__ bv, int neox) {
    B2_PDL_TRIGGER();
    B2_PDL_WAIT();
    const int t = blockIdx.x;
    const int p = pos[t];
    const int sl = slot[t];
    const float *cs = rope_table + (int64_t)p * hd;
    const int half = hd / 2;
    const int page = kv.page_table[(int64_t)sl * kv.max_pages + p / kv.page_size];
    const int64_t cell = ((int64_t)page * kv.page_size + p % kv.page_size) * kv.kvd;
    float *qt = q + (int64_t)t * n_head * hd;
    float *kt = k + (int64_t)t * n_head_kv * hd;
    const float *vt = v + (int64_t)t * n_head_kv * hd;
    const int nq = n_head * half, nk = n_head_kv * half;
    // blockIdx.y: the pairs (and the V row) are split over gridDim.y CTAs per token -- a decode batch is few tokens
    for (int i = blockIdx.y * blockDim.x + threadIdx.x; i < nq + nk; i += gridDim.y * blockDim.x) {
        const bool is_q = i < nq;
        const int ii = is_q ? i : i - nq;
        const int h = ii / half, pr = ii % half;
        const int i0 = neox ? pr : 2 * pr, i1 = neox ? pr + half : 2 * pr + 1; // the two elements of pair pr inside the head
        float *x = (is_q ? qt : kt) + h * hd;
        const float *bias = is_q ? bq : bk;
        const float c = cs[2 * pr], sn = cs[2 * pr + 1];
        float x0 = x[i0], x1 = x[i1];
        if (bias) {
            x0 = __fadd_rn(x0, bias[h * hd + i0]);
            x1 = __fadd_rn(x1, bias[h * hd + i1]);
        }
        const float r0 = __fsub_rn(__fmul_rn(x0, c), __fmul_rn(x1, sn));
        const float r1 = __fadd_rn(__fmul_rn(x0, sn), __fmul_rn(x1, c));
        if (is_q) {
            x[i0] = r0;
            x[i1] = r1;
        } else if (!neox) {
            *(__half2 *)(kv.k_pool + cell + h * hd + 2 * pr) = __halves2half2(__float2half_rn(r0), __float2half_rn(r1));
        } else {
            kv.k_pool[cell + h * hd + i0] = __float2half_rn(r0);
            kv.k_pool[cell + h * hd + i1] = __float2half_rn(r1);
        }
    }
    for (int i = blockIdx.y * blockDim.x + threadIdx.x; i < kv.kvd; i += gridDim.y * blockDim.x)
        kv.v_pool[cell + i] = __float2half_rn(bv ? __fadd_rn(vt[i], bv[i]) : vt[i]);
}

void launch_rope_kv_store(float *q, float *k, const float *v, const int32_t *pos, const int32_t *slot, int n_tok, int n_head,
                          int n_head_kv, int hd, const float *rope_table, const KVLayout &kv, cudaStream_t s, const float *bq,
                          const float *bk, const float *bv, bool neox) {
    if (n_tok <= 0) return;
    const int split = n_tok >= 512 ? 1 : n_tok >= 128 ? 2 : 8;
    launch_pdl(rope_kv_store_kernel, dim3(n_tok, split), dim3(256), 0, s, q, k, v, pos, slot, n_head, n_head_kv, hd, rope_table, kv, bq, bk, bv,
               neox ? 1 : 0);
    g_launch_count++;
    B2_CUDA(cudaGetLastError());
}

// ------------------------------------------------------------------------------------------ attention (paged KV)
// One CTA per (head, token).  Two passes so that the probabilities can be normalised and rounded to f16 before the
// P.V product exactly as ggml's mul_mat(v_f16, soft_max(kq)) does.
//   pass 1: ONE LANE PER KV POSITION -- the lane streams the whole K row (hd halves, 16-byte loads, all independent)
//           and accumulates q.k in index order; no shuffles, hd/8 loads in flight per lane
//   pass 2: a warp reads a whole V row coalesced (lane owns hd/32 dims), warps split the positions 4 at a time,
//           partial outputs are combined across warps in a fixed order (deterministic)
// HBM/L2-bound: 2 x n_kv x hd x 2 B per head.  (see attention.cuh: also used by the decode mega-kernel)
__global__ void __launch_bounds__(ATT_THREADS)
attention_kernel(const float *__restrict__ q, const int32_t *__restrict__ pos, const int32_t *__restrict__ slot, int n_head,
                 int n_head_kv, int hd, KVLayout kv, float *__restrict__ out, float scale) {
    extern __shared__ __align__(16) float smem[];
    B2_PDL_TRIGGER();
    B2_PDL_WAIT();
    const int h = blockIdx.x, t = blockIdx.y;
    attention_head(q + ((int64_t)t * n_head + h) * hd, pos[t] + 1, kv.page_table + (int64_t)slot[t] * kv.max_pages, kv,
                   h / (n_head / n_head_kv), hd, scale, out + ((int64_t)t * n_head + h) * hd, smem, threadIdx.x, ATT_THREADS);
}

// grouped-query variant: one CTA per (KV head, token) serves the G query heads of the group (attention.cuh)
template <int G>
__global__ void __launch_bounds__(ATT_THREADS)
attention_group_kernel(const float *__restrict__ q, const int32_t *__restrict__ pos, const int32_t *__restrict__ slot, int n_head,
                       int hd, KVLayout kv, float *__restrict__ out, float scale) {
    extern __shared__ __align__(16) float smem[];
    B2_PDL_TRIGGER();
    B2_PDL_WAIT();
    const int hk = blockIdx.x, t = blockIdx.y;
    const int64_t row0 = ((int64_t)t * n_head + (int64_t)hk * G) * hd; // heads hk*G .. hk*G+G-1 are consecutive
    attention_group<G>(q + row0, pos[t] + 1, kv.page_table + (int64_t)slot[t] * kv.max_pages, kv, hk, hd, scale, out + row0, smem,
                       threadIdx.x, ATT_THREADS);
}

// Decode batches whose tokens all sit in different slots: RoPE of q (in place) and k, the KV store of this token, the
// attention of the G query heads and the conversion of the output into the next GEMM's fp16 activation tiles in ONE
// launch (rope_kv_store_kernel + attention_group_kernel + qgemm_convert_x_kernel otherwise).  The CTA (KV head hk,
// token t) owns everything it touches: the G query heads of the group, row hk of k and v, one cache cell.
template <int G, int MINB>
__global__ void __launch_bounds__(ATT_THREADS, MINB)
attention_decode_fused_kernel(float *__restrict__ q, const float *__restrict__ k, const float *__restrict__ v,
                              const int32_t *__restrict__ pos, const int32_t *__restrict__ slot, int n_head, int n_head_kv, int hd,
                              const float *__restrict__ rope_table, KVLayout kv, float scale, __half *__restrict__ x16, int n_pad) {
    extern __shared__ __align__(16) float smem[];
    B2_PDL_TRIGGER();
    B2_PDL_WAIT();
    const int hk = blockIdx.x, t = blockIdx.y;
    const int p = pos[t], sl = slot[t];
    const float *cs = rope_table + (int64_t)p * hd;
    const int half = hd / 2;
    const int page = kv.page_table[(int64_t)sl * kv.max_pages + p / kv.page_size];
    const int64_t cell = ((int64_t)page * kv.page_size + p % kv.page_size) * kv.kvd;
    const int64_t row0 = ((int64_t)t * n_head + (int64_t)hk * G) * hd;
    float *qt = q + row0;
    const float *kt = k + ((int64_t)t * n_head_kv + hk) * hd, *vt = v + ((int64_t)t * n_head_kv + hk) * hd;
    for (int i = threadIdx.x; i < (G + 1) * half; i += blockDim.x) { // the arithmetic of rope_kv_store_kernel
        const bool is_q = i < G * half;
        const int ii = is_q ? i : i - G * half;
        const int h = ii / half, pr = ii % half;
        const float *x = (is_q ? qt + h * hd : kt) + 2 * pr;
        const float c = cs[2 * pr], sn = cs[2 * pr + 1];
        const float x0 = x[0], x1 = x[1];
        const float r0 = __fsub_rn(__fmul_rn(x0, c), __fmul_rn(x1, sn));
        const float r1 = __fadd_rn(__fmul_rn(x0, sn), __fmul_rn(x1, c));
        if (is_q) {
            qt[h * hd + 2 * pr] = r0;
            qt[h * hd + 2 * pr + 1] = r1;
        } else {
            *(__half2 *)(kv.k_pool + cell + hk * hd + 2 * pr) = __halves2half2(__float2half_rn(r0), __float2half_rn(r1));
        }
    }
    for (int i = threadIdx.x; i < hd; i += blockDim.x) kv.v_pool[cell + hk * hd + i] = __float2half_rn(vt[i]);
    __threadfence_block();
    __syncthreads(); // this CTA's q rows and cache cell are written: the attention below reads them back (L2 / L1 of this SM)
    attention_group<G>(qt, p + 1, kv.page_table + (int64_t)sl * kv.max_pages, kv, hk, hd, scale, nullptr, smem, threadIdx.x, ATT_THREADS, x16, t,
                       ((int64_t)hk * G) * hd, n_pad);
}

template <int G>
static bool try_launch_attention_decode_fused(float *q, const float *k, const float *v, const int32_t *pos, const int32_t *slot, int n_tok,
                                              int n_head, int n_head_kv, int hd, const float *rope_table, const KVLayout &kv, int max_ctx,
                                              void *x16, int n_pad, cudaStream_t s) {
    const size_t smem = attention_group_smem_bytes(G, hd, max_ctx, ATT_THREADS);
    if (smem > 96 * 1024) return false;
    const float scale = 1.0f / sqrtf((float)hd);
    // 4 CTAs per SM where the registers allow (G <= 4): a 64-stream step of 8 KV heads is 512 CTAs = one wave of 148 x 4.
    // Measured on Llama-3-8B, 64 streams: 4.61 ms per step against 4.76 at 3 CTAs/SM; 8 V rows in flight per warp and
    // scores by lane groups (coalesced K rows) were slower (4.70 / 4.93 ms: registers, hence fewer resident CTAs).
    constexpr int MINB = G <= 4 ? 4 : 2; // (G = 8 at 80 registers spills 364 B)
    static PerDeviceOnce configured;
    configured.run([&] { B2_CUDA(cudaFuncSetAttribute(attention_decode_fused_kernel<G, MINB>, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024)); });
    launch_pdl(attention_decode_fused_kernel<G, MINB>, dim3(n_head_kv, n_tok), dim3(ATT_THREADS), smem, s, q, k, v, pos, slot, n_head, n_head_kv, hd,
               rope_table, kv, scale, (__half *)x16, n_pad);
    g_launch_count++;
    B2_CUDA(cudaGetLastError());
    return true;
}

bool launch_attention_decode_fused(float *q, const float *k, const float *v, const int32_t *pos, const int32_t *slot, int n_tok, int n_head,
                                   int n_head_kv, int hd, const float *rope_table, const KVLayout &kv, int max_ctx, void *x_tiles, int n_pad,
                                   cudaStream_t s) {
    if (n_tok <= 0 || n_head % n_head_kv != 0 || hd > 128 || hd % 32 != 0) return false;
    const int G = n_head / n_head_kv;
    if (G == 2) return try_launch_attention_decode_fused<2>(q, k, v, pos, slot, n_tok, n_head, n_head_kv, hd, rope_table, kv, max_ctx, x_tiles, n_pad, s);
    if (G == 4) return try_launch_attention_decode_fused<4>(q, k, v, pos, slot, n_tok, n_head, n_head_kv, hd, rope_table, kv, max_ctx, x_tiles, n_pad, s);
    if (G == 8) return try_launch_attention_decode_fused<8>(q, k, v, pos, slot, n_tok, n_head, n_head_kv, hd, rope_table, kv, max_ctx, x_tiles, n_pad, s);
    return false;
}

template <int G>
static bool try_launch_attention_group(const float *q, const int32_t *pos, const int32_t *slot, int n_tok, int n_head, int n_head_kv,
                                       int hd, const KVLayout &kv, float *out, int max_ctx, cudaStream_t s) {
    const size_t smem = attention_group_smem_bytes(G, hd, max_ctx, ATT_THREADS);
    if (smem > 96 * 1024) return false; // keep at least two CTAs per SM; longer contexts use the per-head kernel
    static PerDeviceOnce configured;
    configured.run([&] { B2_CUDA(cudaFuncSetAttribute(attention_group_kernel<G>, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024)); });
    const float scale = 1.0f / sqrtf((float)hd);
    launch_pdl(attention_group_kernel<G>, dim3(n_head_kv, n_tok), dim3(ATT_THREADS), smem, s, q, pos, slot, n_head, hd, kv, out, scale);
    g_launch_count++;
    B2_CUDA(cudaGetLastError());
    return true;
}

void launch_attention(const float *q, const int32_t *pos, const int32_t *slot, int n_tok, int n_head, int n_head_kv, int hd,
                      const KVLayout &kv, float *out, int max_ctx, cudaStream_t s) {
    if (n_tok <= 0) return;
    B2_CHECK(hd % 32 == 0 && hd <= 256, B2_EUNSUPPORTED, "attention: head_dim must be a multiple of 32, at most 256");
    // B200LLM_NO_GQA_ATT=1 (A/B tests) is read at every call on purpose: the bit-identity test flips it inside one process
    const char *no_gqa = getenv("B200LLM_NO_GQA_ATT");
    if (n_head % n_head_kv == 0 && hd <= 128 && !(no_gqa && no_gqa[0] == '1')) {
        const int G = n_head / n_head_kv;
        if (G == 2 && try_launch_attention_group<2>(q, pos, slot, n_tok, n_head, n_head_kv, hd, kv, out, max_ctx, s)) return;
        if (G == 4 && try_launch_attention_group<4>(q, pos, slot, n_tok, n_head, n_head_kv, hd, kv, out, max_ctx, s)) return;
        if (G == 8 && try_launch_attention_group<8>(q, pos, slot, n_tok, n_head, n_head_kv, hd, kv, out, max_ctx, s)) return;
    }
    const size_t smem = attention_smem_bytes(hd, max_ctx, ATT_THREADS);
    B2_CHECK(smem <= 220 * 1024, B2_EUNSUPPORTED, "attention: context too long for the two-pass kernel");
    static PerDeviceOnce configured;
    if (smem > 48 * 1024)
        configured.run([&] { B2_CUDA(cudaFuncSetAttribute(attention_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024)); });
    const float scale = 1.0f / sqrtf((float)hd);
    launch_pdl(attention_kernel, dim3(n_head, n_tok), dim3(ATT_THREADS), smem, s, q, pos, slot, n_head, n_head_kv, hd, kv, out, scale);
    g_launch_count++;
    B2_CUDA(cudaGetLastError());
}

// ------------------------------------------------------------------------------------------ greedy sampling
// First maximal element, like std::max_element in llama.cpp's greedy sampler.  A row (128 256 logits for Llama-3) is
// split over ARGMAX_SPLIT CTAs; each folds its slice into one 64-bit key (order-preserving bits of the value in the
// high word, ~index in the low word, so a plain max picks the largest value and, among equals, the smallest index)
// and merges it with atomicMax -- order independent, hence deterministic.  The last CTA of the row to finish writes
// the token (and the device-side decode loop's next inputs) and re-zeroes the scratch for the next launch.
constexpr int ARGMAX_SPLIT = 32;
constexpr int ARGMAX_MAX_ROWS = 4096;
static __device__ unsigned long long g_argmax_key[ARGMAX_MAX_ROWS];
static __device__ unsigned int g_argmax_done[ARGMAX_MAX_ROWS];

__device__ __forceinline__ unsigned long long argmax_pack(float v, int i) {
    uint32_t b = __float_as_uint(v);
    b = (b & 0x80000000u) ? ~b : (b | 0x80000000u); // monotone float -> unsigned
    return ((unsigned long long)b << 32) | (uint32_t)(~(uint32_t)i);
}

__global__ void __launch_bounds__(256)
argmax_kernel(const float *__restrict__ logits, int64_t stride, int n_vocab, int32_t *__restrict__ out,
              int32_t *__restrict__ feed_tok, int32_t *__restrict__ feed_pos) {
    B2_PDL_TRIGGER();
    B2_PDL_WAIT();
    const int row = blockIdx.y;
    const float *x = logits + (int64_t)row * stride;
    const int per = (n_vocab + ARGMAX_SPLIT - 1) / ARGMAX_SPLIT;
    const int i0 = blockIdx.x * per, i1 = min(n_vocab, i0 + per);
    unsigned long long best = 0ull; // below every real key (NaNs aside): a real key has ~i >= 0x80000000 for i < 2^31
    for (int i = i0 + threadIdx.x; i < i1; i += blockDim.x) {
        const float v = x[i];
        if (v == v) best = max(best, argmax_pack(v, i));
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) best = max(best, __shfl_xor_sync(0xffffffffu, best, o));
    __shared__ unsigned long long sv[8];
    if ((threadIdx.x & 31) == 0) sv[threadIdx.x >> 5] = best;
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < 8; w++) best = max(best, sv[w]);
        atomicMax(&g_argmax_key[row], best);
        __threadfence();
        if (atomicAdd(&g_argmax_done[row], 1u) == (unsigned)gridDim.x - 1) { // last CTA of the row
            __threadfence();
            const unsigned long long k = atomicExch(&g_argmax_key[row], 0ull);
            g_argmax_done[row] = 0u;
            const int tok = k == 0ull ? 0 : (int)(~(uint32_t)(k & 0xFFFFFFFFull));
            out[row] = tok;
            if (feed_tok) feed_tok[row] = tok;
            if (feed_pos) feed_pos[row] += 1;
        }
    }
}

void launch_argmax(const float *logits, int64_t stride, int n_rows, int n_vocab, int32_t *out, int32_t *feed_tok,
                   int32_t *feed_pos, cudaStream_t s) {
    if (n_rows <= 0) return;
    B2_CHECK(n_rows <= ARGMAX_MAX_ROWS, B2_EINVAL, "argmax: too many rows");
    launch_pdl(argmax_kernel, dim3(ARGMAX_SPLIT, n_rows), dim3(256), 0, s, logits, stride, n_vocab, out, feed_tok, feed_pos);
    g_launch_count++;
    B2_CUDA(cudaGetLastError());
}

} // namespace b2
