// kernels_bert.cu -- the element-wise, normalisation, attention and pooling kernels of the BERT encoder (bert.h).
// The mat-muls are the tcgen05 GEMM of kernels_gemm.cu; everything here is HBM/L2-bound fp32 row work.
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <math_constants.h>

#include <cstdlib>

#include "bert.h"

namespace b2 {

namespace {

constexpr int LN_THREADS = 256;

__device__ __forceinline__ float block_sum(float v, float *red) { // red: 32 floats of shared memory
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
    __syncthreads(); // red may still be read from the previous reduction
    if (lane == 0) red[warp] = v;
    __syncthreads();
    float t = lane < nw ? red[lane] : 0.f;
    for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
    return t;
}

// ggml_norm: mean, then the variance of the centred values, y = (x - mean) / sqrt(var + eps); then * w + b
__device__ __forceinline__ void layernorm_row(float *row, const float *w, const float *b, int d, float eps, float *red) {
    float s = 0.f;
    for (int i = threadIdx.x; i < d; i += blockDim.x) s += row[i];
    const float mean = block_sum(s, red) / (float)d;
    float s2 = 0.f;
    for (int i = threadIdx.x; i < d; i += blockDim.x) {
        const float c = row[i] - mean;
        s2 += c * c;
    }
    const float inv = rsqrtf(block_sum(s2, red) / (float)d + eps);
    for (int i = threadIdx.x; i < d; i += blockDim.x) row[i] = (row[i] - mean) * inv * w[i] + b[i];
}

__global__ void __launch_bounds__(LN_THREADS) bert_embed_ln_kernel(float *x, const float *pos_embd, const float *type0, const int32_t *pos,
                                                                   const float *w, const float *b, int d, float eps) {
    __shared__ float red[32];
    float *row = x + (size_t)blockIdx.x * d;
    const float *pe = pos_embd + (size_t)pos[blockIdx.x] * d;
    for (int i = threadIdx.x; i < d; i += blockDim.x) row[i] = row[i] + type0[i] + pe[i];
    __syncthreads();
    layernorm_row(row, w, b, d, eps, red);
}

__global__ void __launch_bounds__(LN_THREADS) bert_bias_residual_ln_kernel(float *x, const float *h, const float *bias, const float *w,
                                                                           const float *b, int d, float eps) {
    __shared__ float red[32];
    float *row = x + (size_t)blockIdx.x * d;
    const float *hr = h + (size_t)blockIdx.x * d;
    for (int i = threadIdx.x; i < d; i += blockDim.x) row[i] = row[i] + (hr[i] + bias[i]);
    __syncthreads();
    layernorm_row(row, w, b, d, eps, red);
}

// One CTA = one head of one sequence x QB queries; K/V go through shared memory in tiles of KT keys, the softmax is
// the running (flash) form, exact in fp32.  Lane l owns keys l, l + 32 of a tile for the scores and features
// l, l + 32 of the output.
constexpr int ATT_QB = 32, ATT_KT = 64, ATT_WARPS = 4;

template <int HD>
__global__ void __launch_bounds__(ATT_WARPS * 32) bert_attention_kernel(const float *__restrict__ q, const float *__restrict__ k,
                                                                        const float *__restrict__ v, const float *__restrict__ bq,
                                                                        const float *__restrict__ bk, const float *__restrict__ bv,
                                                                        const int32_t *__restrict__ offs, int n_head, float *__restrict__ out) {
    constexpr int NC = HD / 32, QW = ATT_QB / ATT_WARPS;
    __shared__ float Qs[ATT_QB][HD];
    __shared__ float Ks[ATT_KT][HD + 1];
    __shared__ float Vs[ATT_KT][HD];
    const int seq = blockIdx.z, head = blockIdx.y, t0 = offs[seq], len = offs[seq + 1] - t0;
    const int q0 = blockIdx.x * ATT_QB;
    if (q0 >= len) return;
    const int d = n_head * HD, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const float scale = rsqrtf((float)HD);
    for (int i = tid; i < ATT_QB * HD; i += ATT_WARPS * 32) {
        const int r = i / HD, c = i % HD;
        Qs[r][c] = q0 + r < len ? (q[(size_t)(t0 + q0 + r) * d + head * HD + c] + bq[head * HD + c]) * scale : 0.f;
    }
    float m[QW], l[QW], acc[QW][NC];
#pragma unroll
    for (int i = 0; i < QW; i++) {
        m[i] = -CUDART_INF_F;
        l[i] = 0.f;
#pragma unroll
        for (int c = 0; c < NC; c++) acc[i][c] = 0.f;
    }
    for (int k0 = 0; k0 < len; k0 += ATT_KT) {
        __syncthreads(); // the previous tile is consumed (and Qs is written, first time round)
        for (int i = tid; i < ATT_KT * HD; i += ATT_WARPS * 32) {
            const int r = i / HD, c = i % HD;
            const bool ok = k0 + r < len;
            const size_t g = (size_t)(t0 + k0 + (ok ? r : 0)) * d + head * HD + c;
            Ks[r][c] = ok ? k[g] + bk[head * HD + c] : 0.f;
            Vs[r][c] = ok ? v[g] + bv[head * HD + c] : 0.f;
        }
        __syncthreads();
        const bool ok0 = k0 + lane < len, ok1 = k0 + lane + 32 < len;
#pragma unroll
        for (int i = 0; i < QW; i++) {
            const int qi = warp * QW + i;
            if (q0 + qi >= len) continue; // warp-uniform
            float s0 = 0.f, s1 = 0.f;
#pragma unroll 8
            for (int c = 0; c < HD; c++) {
                const float qv = Qs[qi][c];
                s0 += qv * Ks[lane][c];
                s1 += qv * Ks[lane + 32][c];
            }
            s0 = ok0 ? s0 : -CUDART_INF_F;
            s1 = ok1 ? s1 : -CUDART_INF_F;
            float mt = fmaxf(s0, s1);
            for (int o = 16; o > 0; o >>= 1) mt = fmaxf(mt, __shfl_xor_sync(0xffffffffu, mt, o));
            const float mn = fmaxf(m[i], mt); // finite: key k0 of every tile exists
            const float sc = __expf(m[i] - mn);
            const float p0 = __expf(s0 - mn), p1 = __expf(s1 - mn);
            m[i] = mn;
            l[i] = l[i] * sc + (p0 + p1);
#pragma unroll
            for (int c = 0; c < NC; c++) acc[i][c] *= sc;
#pragma unroll 8
            for (int j = 0; j < 32; j++) {
                const float pa = __shfl_sync(0xffffffffu, p0, j), pb = __shfl_sync(0xffffffffu, p1, j);
#pragma unroll
                for (int c = 0; c < NC; c++) acc[i][c] += pa * Vs[j][lane + 32 * c] + pb * Vs[j + 32][lane + 32 * c];
            }
        }
    }
#pragma unroll
    for (int i = 0; i < QW; i++) {
        const int qi = warp * QW + i;
        if (q0 + qi >= len) continue;
        float lt = l[i];
        for (int o = 16; o > 0; o >>= 1) lt += __shfl_xor_sync(0xffffffffu, lt, o);
        const float inv = 1.f / lt;
#pragma unroll
        for (int c = 0; c < NC; c++) out[(size_t)(t0 + q0 + qi) * d + head * HD + lane + 32 * c] = acc[i][c] * inv;
    }
}

// The same attention on the warp-level tensor-core path: S = Q K^T and O = P V as m16n8k16 bf16 MMAs with fp32
// accumulators, the running softmax in fp32 registers in between (the accumulator fragment of S is the A fragment of
// the next product, so P never leaves registers).  One warp = 16 queries, one CTA = 8 warps = 128 queries of one head.
// K and V are staged as bf16 [key][dim] rows padded by 8 elements, so that the 32 lanes of a K fragment load (8 rows x
// 4 words) and the 16 row addresses of a transposing ldmatrix on V fall into different banks.
constexpr int MMA_WARPS = 8, MMA_QB = 16 * MMA_WARPS, MMA_KT = 64, MMA_THREADS = 32 * MMA_WARPS;

__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi) {
    const __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<const uint32_t *>(&v);
}
__device__ __forceinline__ void mma_16816(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

template <int HD>
__global__ void __launch_bounds__(MMA_THREADS) bert_attention_mma_kernel(const float *__restrict__ q, const float *__restrict__ k,
                                                                 const float *__restrict__ v, const float *__restrict__ bq,
                                                                 const float *__restrict__ bk, const float *__restrict__ bv,
                                                                 const int32_t *__restrict__ offs, int n_head, float *__restrict__ out) {
    constexpr int KS = HD / 16, ND = HD / 8, NB = MMA_KT / 8, PAD = 8;
    __shared__ __align__(16) __nv_bfloat16 Kb[MMA_KT][HD + PAD];
    __shared__ __align__(16) __nv_bfloat16 Vb[MMA_KT][HD + PAD];
    const int seq = blockIdx.z, head = blockIdx.y, t0 = offs[seq], len = offs[seq + 1] - t0;
    const int q0 = blockIdx.x * MMA_QB;
    if (q0 >= len) return;
    const int d = n_head * HD, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int g = lane >> 2, tq = lane & 3; // row inside the fragment, position inside the quad
    const float scale = rsqrtf((float)HD);
    const int r0 = q0 + warp * 16 + g, r1 = r0 + 8; // the two query rows this thread holds parts of
    uint32_t qa[KS][4];
#pragma unroll
    for (int ks = 0; ks < KS; ks++) {
#pragma unroll
        for (int h = 0; h < 2; h++) { // h: columns +0 / +8 of the k-step
            const int c = ks * 16 + h * 8 + tq * 2;
            const float b0 = bq[head * HD + c], b1 = bq[head * HD + c + 1];
            const float *p0 = q + (size_t)(t0 + r0) * d + head * HD + c, *p1 = q + (size_t)(t0 + r1) * d + head * HD + c;
            qa[ks][2 * h + 0] = r0 < len ? pack_bf16((p0[0] + b0) * scale, (p0[1] + b1) * scale) : 0u;
            qa[ks][2 * h + 1] = r1 < len ? pack_bf16((p1[0] + b0) * scale, (p1[1] + b1) * scale) : 0u;
        }
    }
    float o[ND][4];
#pragma unroll
    for (int i = 0; i < ND; i++) o[i][0] = o[i][1] = o[i][2] = o[i][3] = 0.f;
    float m0 = -CUDART_INF_F, m1 = -CUDART_INF_F, l0 = 0.f, l1 = 0.f;
    for (int k0 = 0; k0 < len; k0 += MMA_KT) {
        __syncthreads();
        for (int i = tid; i < MMA_KT * (HD / 4); i += MMA_THREADS) { // four features per thread: 16-byte loads, 8-byte stores
            const int r = i / (HD / 4), c = (i % (HD / 4)) * 4;
            const bool ok = k0 + r < len;
            const size_t gi = (size_t)(t0 + k0 + (ok ? r : 0)) * d + head * HD + c;
            const float4 kv = *reinterpret_cast<const float4 *>(k + gi), vv = *reinterpret_cast<const float4 *>(v + gi);
            const float4 kb = *reinterpret_cast<const float4 *>(bk + head * HD + c), vb = *reinterpret_cast<const float4 *>(bv + head * HD + c);
            uint2 pk, pv;
            pk.x = ok ? pack_bf16(kv.x + kb.x, kv.y + kb.y) : 0u;
            pk.y = ok ? pack_bf16(kv.z + kb.z, kv.w + kb.w) : 0u;
            pv.x = ok ? pack_bf16(vv.x + vb.x, vv.y + vb.y) : 0u;
            pv.y = ok ? pack_bf16(vv.z + vb.z, vv.w + vb.w) : 0u;
            *reinterpret_cast<uint2 *>(&Kb[r][c]) = pk;
            *reinterpret_cast<uint2 *>(&Vb[r][c]) = pv;
        }
        __syncthreads();
        float s[NB][4];
#pragma unroll
        for (int nb = 0; nb < NB; nb++) {
            s[nb][0] = s[nb][1] = s[nb][2] = s[nb][3] = 0.f;
#pragma unroll
            for (int ks = 0; ks < KS; ks++) {
                const uint32_t b0 = *reinterpret_cast<const uint32_t *>(&Kb[nb * 8 + g][ks * 16 + tq * 2]);
                const uint32_t b1 = *reinterpret_cast<const uint32_t *>(&Kb[nb * 8 + g][ks * 16 + 8 + tq * 2]);
                mma_16816(s[nb], qa[ks], b0, b1);
            }
        }
        float mt0 = -CUDART_INF_F, mt1 = -CUDART_INF_F;
#pragma unroll
        for (int nb = 0; nb < NB; nb++) {
            const int key = k0 + nb * 8 + tq * 2;
            if (key >= len) s[nb][0] = s[nb][2] = -CUDART_INF_F;
            if (key + 1 >= len) s[nb][1] = s[nb][3] = -CUDART_INF_F;
            mt0 = fmaxf(mt0, fmaxf(s[nb][0], s[nb][1]));
            mt1 = fmaxf(mt1, fmaxf(s[nb][2], s[nb][3]));
        }
        mt0 = fmaxf(mt0, __shfl_xor_sync(0xffffffffu, mt0, 1));
        mt0 = fmaxf(mt0, __shfl_xor_sync(0xffffffffu, mt0, 2));
        mt1 = fmaxf(mt1, __shfl_xor_sync(0xffffffffu, mt1, 1));
        mt1 = fmaxf(mt1, __shfl_xor_sync(0xffffffffu, mt1, 2));
        const float mn0 = fmaxf(m0, mt0), mn1 = fmaxf(m1, mt1); // finite: key k0 of every tile exists
        const float sc0 = __expf(m0 - mn0), sc1 = __expf(m1 - mn1);
        m0 = mn0;
        m1 = mn1;
        float ps0 = 0.f, ps1 = 0.f;
#pragma unroll
        for (int nb = 0; nb < NB; nb++) {
            s[nb][0] = __expf(s[nb][0] - mn0);
            s[nb][1] = __expf(s[nb][1] - mn0);
            s[nb][2] = __expf(s[nb][2] - mn1);
            s[nb][3] = __expf(s[nb][3] - mn1);
            ps0 += s[nb][0] + s[nb][1];
            ps1 += s[nb][2] + s[nb][3];
        }
        l0 = l0 * sc0 + ps0; // per-thread part of the row sum; the quad is summed at the end
        l1 = l1 * sc1 + ps1;
#pragma unroll
        for (int nd = 0; nd < ND; nd++) {
            o[nd][0] *= sc0;
            o[nd][1] *= sc0;
            o[nd][2] *= sc1;
            o[nd][3] *= sc1;
        }
#pragma unroll
        for (int kk = 0; kk < MMA_KT / 16; kk++) {
            uint32_t pa[4];
            pa[0] = pack_bf16(s[2 * kk][0], s[2 * kk][1]);
            pa[1] = pack_bf16(s[2 * kk][2], s[2 * kk][3]);
            pa[2] = pack_bf16(s[2 * kk + 1][0], s[2 * kk + 1][1]);
            pa[3] = pack_bf16(s[2 * kk + 1][2], s[2 * kk + 1][3]);
#pragma unroll
            for (int nd = 0; nd < ND; nd++) {
                // B fragment = V[16 keys][8 features] read transposed: lanes 0-15 name the 16 key rows
                uint32_t b0, b1;
                const uint32_t addr = (uint32_t)__cvta_generic_to_shared(&Vb[kk * 16 + (lane & 15)][nd * 8]);
                asm volatile("ldmatrix.sync.aligned.m8n8.x2.trans.shared.b16 {%0, %1}, [%2];" : "=r"(b0), "=r"(b1) : "r"(addr));
                mma_16816(o[nd], pa, b0, b1);
            }
        }
    }
    l0 += __shfl_xor_sync(0xffffffffu, l0, 1);
    l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
    l1 += __shfl_xor_sync(0xffffffffu, l1, 1);
    l1 += __shfl_xor_sync(0xffffffffu, l1, 2);
    const float i0 = 1.f / l0, i1 = 1.f / l1;
#pragma unroll
    for (int nd = 0; nd < ND; nd++) {
        const int c = head * HD + nd * 8 + tq * 2;
        if (r0 < len) *reinterpret_cast<float2 *>(out + (size_t)(t0 + r0) * d + c) = make_float2(o[nd][0] * i0, o[nd][1] * i0);
        if (r1 < len) *reinterpret_cast<float2 *>(out + (size_t)(t0 + r1) * d + c) = make_float2(o[nd][2] * i1, o[nd][3] * i1);
    }
}

// ---- the same attention on fp16 rows (q | k | v as the projection GEMM's epilogue wrote them, biases included), for
// head size 64.  K and V tiles of 64 keys arrive by cp.async into a two-stage ring (the next tile loads while this one is
// multiplied), fragments come out of shared memory with ldmatrix.x4 (two k-steps of K, two feature groups of V per
// instruction), the 1/sqrt(hd) scale is folded into the exponent (exp2(s c - m c), c = scale log2 e), and the result goes
// straight into attn_output's operand tiles: [token tile][k tile = head][128 x 64] in core-matrix order, the layout
// convert_x_tiles_kernel would have produced from fp32 rows.
constexpr int F16_ROW = 64 + 8; // halves per staged row: 144 B, the 8 rows of an 8 x 8 matrix fall into different banks

__device__ __forceinline__ void cp_async16(void *dst, const void *src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(dst)), "l"(src) : "memory");
}
__device__ __forceinline__ void mma_16816_f16(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ uint32_t pack_f16(float lo, float hi) {
    const __half2 v = __floats2half2_rn(lo, hi);
    return *reinterpret_cast<const uint32_t *>(&v);
}

__global__ void __launch_bounds__(MMA_THREADS, 2)
bert_attention_f16_kernel(const __half *__restrict__ q, const __half *__restrict__ k, const __half *__restrict__ v,
                          const int32_t *__restrict__ offs, int n_head, __half *__restrict__ x_tiles) {
    constexpr int HD = 64, KS = HD / 16, ND = HD / 8, NB = MMA_KT / 8;
    __shared__ __align__(16) __half Kb[2][MMA_KT][F16_ROW];
    __shared__ __align__(16) __half Vb[2][MMA_KT][F16_ROW];
    const int seq = blockIdx.z, head = blockIdx.y, t0 = offs[seq], len = offs[seq + 1] - t0;
    const int q0 = blockIdx.x * MMA_QB;
    if (q0 >= len) return;
    const int d = n_head * HD, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int g = lane >> 2, tq = lane & 3;
    const int r0 = q0 + warp * 16 + g, r1 = r0 + 8;
    const int n_kt = (len + MMA_KT - 1) / MMA_KT;
    auto load_tile = [&](int kt, int buf) { // 64 keys x 128 B of K and of V: 2 + 2 copies of 16 B per thread; keys past the end re-read the last one
#pragma unroll
        for (int i = tid; i < MMA_KT * 8; i += MMA_THREADS) {
            const int r = i >> 3, c = (i & 7) * 8;
            const size_t gi = (size_t)(t0 + min(kt * MMA_KT + r, len - 1)) * d + head * HD + c;
            cp_async16(&Kb[buf][r][c], k + gi);
            cp_async16(&Vb[buf][r][c], v + gi);
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };
    load_tile(0, 0);
    uint32_t qa[KS][4];
#pragma unroll
    for (int ks = 0; ks < KS; ks++) {
#pragma unroll
        for (int h = 0; h < 2; h++) {
            const int c = head * HD + ks * 16 + h * 8 + tq * 2;
            qa[ks][2 * h + 0] = r0 < len ? *reinterpret_cast<const uint32_t *>(q + (size_t)(t0 + r0) * d + c) : 0u;
            qa[ks][2 * h + 1] = r1 < len ? *reinterpret_cast<const uint32_t *>(q + (size_t)(t0 + r1) * d + c) : 0u;
        }
    }
    const float cexp = rsqrtf((float)HD) * 1.4426950408889634f;
    float o[ND][4];
#pragma unroll
    for (int i = 0; i < ND; i++) o[i][0] = o[i][1] = o[i][2] = o[i][3] = 0.f;
    float m0 = -CUDART_INF_F, m1 = -CUDART_INF_F, l0 = 0.f, l1 = 0.f;
    for (int kt = 0; kt < n_kt; kt++) {
        const int buf = kt & 1, k0 = kt * MMA_KT;
        asm volatile("cp.async.wait_group 0;" ::: "memory");
        __syncthreads(); // tile kt is visible to everyone, and everyone is done with tile kt - 1
        if (kt + 1 < n_kt) load_tile(kt + 1, buf ^ 1);
        float s[NB][4];
#pragma unroll
        for (int nb = 0; nb < NB; nb++) {
            s[nb][0] = s[nb][1] = s[nb][2] = s[nb][3] = 0.f;
#pragma unroll
            for (int kp = 0; kp < KS / 2; kp++) { // matrices: (k-step 2 kp: dims +0, +8), (k-step 2 kp + 1: dims +0, +8)
                uint32_t b[4];
                const uint32_t addr = (uint32_t)__cvta_generic_to_shared(&Kb[buf][nb * 8 + (lane & 7)][kp * 32 + (lane >> 3) * 8]);
                asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0, %1, %2, %3}, [%4];" : "=r"(b[0]), "=r"(b[1]), "=r"(b[2]), "=r"(b[3]) : "r"(addr));
                mma_16816_f16(s[nb], qa[2 * kp], b[0], b[1]);
                mma_16816_f16(s[nb], qa[2 * kp + 1], b[2], b[3]);
            }
        }
        float mt0 = -CUDART_INF_F, mt1 = -CUDART_INF_F;
#pragma unroll
        for (int nb = 0; nb < NB; nb++) {
            const int key = k0 + nb * 8 + tq * 2;
            if (key >= len) s[nb][0] = s[nb][2] = -CUDART_INF_F;
            if (key + 1 >= len) s[nb][1] = s[nb][3] = -CUDART_INF_F;
            mt0 = fmaxf(mt0, fmaxf(s[nb][0], s[nb][1]));
            mt1 = fmaxf(mt1, fmaxf(s[nb][2], s[nb][3]));
        }
        mt0 = fmaxf(mt0, __shfl_xor_sync(0xffffffffu, mt0, 1));
        mt0 = fmaxf(mt0, __shfl_xor_sync(0xffffffffu, mt0, 2));
        mt1 = fmaxf(mt1, __shfl_xor_sync(0xffffffffu, mt1, 1));
        mt1 = fmaxf(mt1, __shfl_xor_sync(0xffffffffu, mt1, 2));
        const float mn0 = fmaxf(m0, mt0), mn1 = fmaxf(m1, mt1); // finite: key k0 of every tile exists
        const float sc0 = exp2f((m0 - mn0) * cexp), sc1 = exp2f((m1 - mn1) * cexp);
        m0 = mn0;
        m1 = mn1;
        const float off0 = mn0 * cexp, off1 = mn1 * cexp;
        float ps0 = 0.f, ps1 = 0.f;
#pragma unroll
        for (int nb = 0; nb < NB; nb++) {
            s[nb][0] = exp2f(fmaf(s[nb][0], cexp, -off0));
            s[nb][1] = exp2f(fmaf(s[nb][1], cexp, -off0));
            s[nb][2] = exp2f(fmaf(s[nb][2], cexp, -off1));
            s[nb][3] = exp2f(fmaf(s[nb][3], cexp, -off1));
            ps0 += s[nb][0] + s[nb][1];
            ps1 += s[nb][2] + s[nb][3];
        }
        l0 = l0 * sc0 + ps0;
        l1 = l1 * sc1 + ps1;
#pragma unroll
        for (int nd = 0; nd < ND; nd++) {
            o[nd][0] *= sc0;
            o[nd][1] *= sc0;
            o[nd][2] *= sc1;
            o[nd][3] *= sc1;
        }
#pragma unroll
        for (int kk = 0; kk < MMA_KT / 16; kk++) {
            uint32_t pa[4];
            pa[0] = pack_f16(s[2 * kk][0], s[2 * kk][1]);
            pa[1] = pack_f16(s[2 * kk][2], s[2 * kk][3]);
            pa[2] = pack_f16(s[2 * kk + 1][0], s[2 * kk + 1][1]);
            pa[3] = pack_f16(s[2 * kk + 1][2], s[2 * kk + 1][3]);
#pragma unroll
            for (int np = 0; np < ND / 2; np++) { // V[16 keys][8 features] transposed, for feature groups 2 np and 2 np + 1
                uint32_t b[4];
                const uint32_t addr = (uint32_t)__cvta_generic_to_shared(&Vb[buf][kk * 16 + (lane & 15)][(2 * np + (lane >> 4)) * 8]);
                asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0, %1, %2, %3}, [%4];" : "=r"(b[0]), "=r"(b[1]), "=r"(b[2]), "=r"(b[3]) : "r"(addr));
                mma_16816_f16(o[2 * np], pa, b[0], b[1]);
                mma_16816_f16(o[2 * np + 1], pa, b[2], b[3]);
            }
        }
    }
    l0 += __shfl_xor_sync(0xffffffffu, l0, 1);
    l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
    l1 += __shfl_xor_sync(0xffffffffu, l1, 1);
    l1 += __shfl_xor_sync(0xffffffffu, l1, 2);
    const float i0 = 1.f / l0, i1 = 1.f / l1;
    // operand tile of (token tile, head): element (r, c) at (r >> 3) * 512 + (c >> 3) * 64 + (r & 7) * 8 + (c & 7)
    const int64_t k_tiles = n_head;
#pragma unroll
    for (int h = 0; h < 2; h++) {
        const int r = h ? r1 : r0;
        if (r >= len) continue;
        const int64_t row = t0 + r;
        __half *tile = x_tiles + ((row >> 7) * k_tiles + head) * (int64_t)(128 * 64) + (int64_t)((row & 127) >> 3) * 512 + (row & 7) * 8 + tq * 2;
        const float inv = h ? i1 : i0;
#pragma unroll
        for (int nd = 0; nd < ND; nd++) *reinterpret_cast<__half2 *>(tile + nd * 64) = __floats2half2_rn(o[nd][2 * h] * inv, o[nd][2 * h + 1] * inv);
    }
}

// mean over the tokens of a sequence (bert.cpp: mul_mat with a vector of 1/N): one CTA per (sequence, 32 features),
// 8 warps stride over the tokens; then x / sqrt(sum x^2) per sequence
__global__ void __launch_bounds__(256) bert_mean_pool_kernel(const float *x, const int32_t *offs, int d, float *out) {
    __shared__ float part[8][32];
    const int t0 = offs[blockIdx.x], n = offs[blockIdx.x + 1] - t0;
    const int f = blockIdx.y * 32 + (threadIdx.x & 31), w = threadIdx.x >> 5;
    float s = 0.f;
    if (f < d)
        for (int t = w; t < n; t += 8) s += x[(size_t)(t0 + t) * d + f];
    part[w][threadIdx.x & 31] = s;
    __syncthreads();
    if (w == 0 && f < d) {
        float tot = 0.f;
#pragma unroll
        for (int i = 0; i < 8; i++) tot += part[i][threadIdx.x];
        out[(size_t)blockIdx.x * d + f] = tot / (float)n;
    }
}

__global__ void __launch_bounds__(LN_THREADS) bert_l2_normalize_kernel(float *out, int d) {
    __shared__ float red[32];
    float *o = out + (size_t)blockIdx.x * d;
    float ss = 0.f;
    for (int i = threadIdx.x; i < d; i += blockDim.x) ss += o[i] * o[i];
    const float inv = rsqrtf(block_sum(ss, red));
    for (int i = threadIdx.x; i < d; i += blockDim.x) o[i] *= inv;
}

} // namespace

void launch_bert_embed_ln(float *x, const float *pos_embd, const float *type0, const int32_t *pos, const float *w, const float *b,
                          int n_tok, int d, float eps, cudaStream_t s) {
    bert_embed_ln_kernel<<<n_tok, LN_THREADS, 0, s>>>(x, pos_embd, type0, pos, w, b, d, eps);
    g_launch_count++;
    B2_CUDA(cudaGetLastError());
}

void launch_bert_bias_residual_ln(float *x, const float *h, const float *bias, const float *w, const float *b, int n_tok, int d,
                                  float eps, cudaStream_t s) {
    bert_bias_residual_ln_kernel<<<n_tok, LN_THREADS, 0, s>>>(x, h, bias, w, b, d, eps);
    g_launch_count++;
    B2_CUDA(cudaGetLastError());
}

void launch_bert_attention_f16(const void *q, const void *k, const void *v, const int32_t *offs, int n_seq, int max_len, int n_head,
                               void *x_tiles, cudaStream_t s) {
    const dim3 grid((max_len + MMA_QB - 1) / MMA_QB, n_head, n_seq);
    bert_attention_f16_kernel<<<grid, MMA_THREADS, 0, s>>>((const __half *)q, (const __half *)k, (const __half *)v, offs, n_head, (__half *)x_tiles);
    g_launch_count++;
    B2_CUDA(cudaGetLastError());
}

void launch_bert_attention(const float *q, const float *k, const float *v, const float *bq, const float *bk, const float *bv,
                           const int32_t *offs, int n_seq, int max_len, int n_head, int hd, float *out, cudaStream_t s) {
    const bool simple = getenv("B200LLM_BERT_ATT_SIMPLE") && atoi(getenv("B200LLM_BERT_ATT_SIMPLE")) != 0;
    if (!simple && (hd == 64 || hd == 32)) { // tensor-core attention; the fp32 CUDA-core kernel below is kept for A/B checks
        const dim3 grid((max_len + MMA_QB - 1) / MMA_QB, n_head, n_seq);
        if (hd == 64) bert_attention_mma_kernel<64><<<grid, MMA_THREADS, 0, s>>>(q, k, v, bq, bk, bv, offs, n_head, out);
        else bert_attention_mma_kernel<32><<<grid, MMA_THREADS, 0, s>>>(q, k, v, bq, bk, bv, offs, n_head, out);
        g_launch_count++;
        B2_CUDA(cudaGetLastError());
        return;
    }
    const dim3 grid((max_len + ATT_QB - 1) / ATT_QB, n_head, n_seq);
    if (hd == 64) bert_attention_kernel<64><<<grid, ATT_WARPS * 32, 0, s>>>(q, k, v, bq, bk, bv, offs, n_head, out);
    else if (hd == 32) bert_attention_kernel<32><<<grid, ATT_WARPS * 32, 0, s>>>(q, k, v, bq, bk, bv, offs, n_head, out);
    else throw Error(B2_EUNSUPPORTED, "bert: head size must be 32 or 64");
    g_launch_count++;
    B2_CUDA(cudaGetLastError());
}

void launch_bert_mean_pool_l2(const float *x, const int32_t *offs, int n_seq, int d, float *out, cudaStream_t s) {
    bert_mean_pool_kernel<<<dim3(n_seq, (d + 31) / 32), 256, 0, s>>>(x, offs, d, out);
    bert_l2_normalize_kernel<<<n_seq, LN_THREADS, 0, s>>>(out, d);
    g_launch_count += 2;
    B2_CUDA(cudaGetLastError());
}

} // namespace b2
