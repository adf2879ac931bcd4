// attention.cuh -- paged-KV attention for one (head, query token), callable by any group of `nthreads` threads
// (a whole CTA in attention_kernel, the CTA of the decode mega-kernel).  See kernels_misc.cu for the description.
#pragma once
#include <cuda_fp16.h>
#include "kernels.h"
#include "quant.cuh"

namespace b2 {

constexpr int ATT_THREADS = 256;

__host__ __device__ inline size_t attention_smem_bytes(int hd, int max_ctx, int nthreads) {
    // q[hd] | red[32] | warp partials [nthreads/32][hd] | scores[max_ctx] | page bases [max_ctx/page + 1] (as int)
    return sizeof(float) * ((size_t)hd + 32 + (size_t)(nthreads / 32) * hd + (size_t)max_ctx + (size_t)max_ctx / 8 + 8);
}

// All `nthreads` threads must call this (contains __syncthreads-equivalent named barrier 1 over nthreads threads).
__device__ __forceinline__ void att_sync(int nthreads) { asm volatile("bar.sync 1, %0;" ::"r"(nthreads)); }

__device__ __forceinline__ void attention_head(const float *__restrict__ qrow, int n_kv, const int32_t *__restrict__ pt,
                                               const KVLayout &kv, int hk, int hd, float scale, float *__restrict__ orow,
                                               float *smem, int tid, int nthreads) {
    const int lane = tid & 31, warp = tid >> 5, nwarps = nthreads >> 5;
    float *qs = smem;                       // [hd]
    float *red = qs + hd;                   // [32]
    float *part = red + 32;                 // [nwarps][hd]
    float *sc = part + (size_t)nwarps * hd; // [n_kv]
    int *pbase = (int *)(sc + ((n_kv + 3) & ~3)); // page index per 32-position page
    const int n_pages = (n_kv + kv.page_size - 1) / kv.page_size;
    for (int i = tid; i < hd; i += nthreads) qs[i] = __half2float(__float2half_rn(__ldcg(qrow + i))); // L2: q may come from another SM
    for (int i = tid; i < n_pages; i += nthreads) pbase[i] = pt[i];
    att_sync(nthreads);

    // ---- pass 1: scores; thread j-strided, one position per thread
    float lmax = -INFINITY;
    for (int j = tid; j < n_kv; j += nthreads) {
        const __half *kr = kv.k_pool + ((int64_t)pbase[j / kv.page_size] * kv.page_size + j % kv.page_size) * kv.kvd + hk * hd;
        float a = 0.f;
        for (int i0 = 0; i0 < hd; i0 += 128) { // up to 16 x 16-byte loads in flight: one round trip per 128 dims
            int4 kq[16];
#pragma unroll
            for (int u = 0; u < 16; u++) kq[u] = *(const int4 *)(kr + min(i0 + u * 8, hd - 8)); // clamped, unconditional
#pragma unroll
            for (int u = 0; u < 16; u++) {
                if (i0 + u * 8 < hd) {
                    const __half2 *kh = (const __half2 *)&kq[u];
#pragma unroll
                    for (int k = 0; k < 4; k++) {
                        const float2 kk = __half22float2(kh[k]);
                        a += kk.x * qs[i0 + u * 8 + 2 * k];
                        a += kk.y * qs[i0 + u * 8 + 2 * k + 1];
                    }
                }
            }
        }
        a *= scale;
        sc[j] = a;
        lmax = fmaxf(lmax, a);
    }
    lmax = warp_max(lmax);
    if (lane == 0) red[warp] = lmax;
    att_sync(nthreads);
    float mx = -INFINITY;
    for (int i = 0; i < nwarps; i++) mx = fmaxf(mx, red[i]);
    att_sync(nthreads);
    float lsum = 0.f;
    for (int j = tid; j < n_kv; j += nthreads) {
        const float e = expf(sc[j] - mx);
        sc[j] = e;
        lsum += e;
    }
    lsum = warp_sum(lsum);
    if (lane == 0) red[warp] = lsum;
    att_sync(nthreads);
    double tot = 0.0;
    for (int i = 0; i < nwarps; i++) tot += (double)red[i];
    const float inv = (float)(1.0 / tot);
    for (int j = tid; j < n_kv; j += nthreads) sc[j] = __half2float(__float2half_rn(__fmul_rn(sc[j], inv)));
    att_sync(nthreads);

    // ---- pass 2: out = sum_j p_j v_j ; lane owns dims [lane*dpl, lane*dpl + dpl), warps split positions
    const int dpl = hd / 32; // 2, 4 or 8 halves per lane
    float acc[8];
#pragma unroll
    for (int i = 0; i < 8; i++) acc[i] = 0.f;
    for (int j0 = warp * 4; j0 < n_kv; j0 += nwarps * 4) {
        uint32_t raw[4][4];
        float p[4];
#pragma unroll
        for (int u = 0; u < 4; u++) {
            const int j = j0 + u;
            const bool ok = j < n_kv;
            p[u] = ok ? sc[j] : 0.f;
            const int jj = ok ? j : 0;
            const __half *vr = kv.v_pool + ((int64_t)pbase[jj / kv.page_size] * kv.page_size + jj % kv.page_size) * kv.kvd + hk * hd +
                               lane * dpl;
            if (dpl == 8) { const int4 v = *(const int4 *)vr; raw[u][0] = v.x; raw[u][1] = v.y; raw[u][2] = v.z; raw[u][3] = v.w; }
            else if (dpl == 4) { const int2 v = *(const int2 *)vr; raw[u][0] = v.x; raw[u][1] = v.y; }
            else { raw[u][0] = *(const uint32_t *)vr; }
        }
#pragma unroll
        for (int u = 0; u < 4; u++) {
#pragma unroll
            for (int k = 0; k < 4; k++) {
                if (2 * k < dpl) {
                    const float2 vv = __half22float2(*(const __half2 *)&raw[u][k]);
                    acc[2 * k] += vv.x * p[u];
                    acc[2 * k + 1] += vv.y * p[u];
                }
            }
        }
    }
#pragma unroll
    for (int i = 0; i < 8; i++)
        if (i < dpl) part[(size_t)warp * hd + lane * dpl + i] = acc[i];
    att_sync(nthreads);
    for (int i = tid; i < hd; i += nthreads) {
        float a = 0.f;
        for (int w = 0; w < nwarps; w++) a += part[(size_t)w * hd + i];
        orow[i] = a;
    }
    att_sync(nthreads);
}

// ---- grouped-query variant: ONE CTA serves the G query heads that share a KV head (token t), so every K and V row is
// fetched once instead of G times and a batched decode step launches n_head_kv x n_tok CTAs instead of n_head x n_tok.
// Per head the arithmetic and its order are those of attention_head (bit-identical results).
// smem: q[G][hd] | red[G][32] | part[nwarps][G][hd] | scores[G][n_pad] | page bases
__host__ __device__ inline size_t attention_group_smem_bytes(int G, int hd, int max_ctx, int nthreads) {
    const size_t n_pad = ((size_t)max_ctx + 3) & ~(size_t)3;
    return sizeof(float) * ((size_t)G * hd + (size_t)G * 32 + (size_t)(nthreads / 32) * G * hd + (size_t)G * n_pad + (size_t)max_ctx / 8 + 8);
}

// Element offset of (token r, feature k) in the fp16 activation tiles of the batched-decode GEMM (kernels_qgemm.cu):
// [k / 256][r / 8][(k % 256) / 8][r % 8][k % 8], and inside every group of four k the positions hold elements (0, 2, 1, 3).
__host__ __device__ inline int64_t qg_x_offset(int r, int64_t k, int n_pad) {
    return (k >> 8) * ((int64_t)n_pad * 256) + (int64_t)(r >> 3) * 2048 + (int64_t)((k & 255) >> 3) * 64 + (r & 7) * 8 + (k & 7);
}
__host__ __device__ inline int64_t qg_x_offset_of_element(int r, int64_t k, int n_pad) { // k = the element's true index
    return qg_x_offset(r, (k & ~(int64_t)3) | ((k & 1) << 1) | ((k >> 1) & 1), n_pad);
}

// x16 (optional): write the G output rows as fp16 into the GEMM's activation tiles (token x16_tok, features x16_k0 ..)
// instead of fp32 rows -- the fused decode-attention kernel feeds attn_output's GEMM directly.
template <int G>
__device__ __forceinline__ void attention_group(const float *__restrict__ qrows, int n_kv, const int32_t *__restrict__ pt,
                                                const KVLayout &kv, int hk, int hd, float scale, float *__restrict__ orows,
                                                float *smem, int tid, int nthreads, __half *__restrict__ x16 = nullptr,
                                                int x16_tok = 0, int64_t x16_k0 = 0, int x16_n_pad = 0) {
    const int lane = tid & 31, warp = tid >> 5, nwarps = nthreads >> 5;
    const int n_pad = (n_kv + 3) & ~3;
    float *qs = smem;                            // [G][hd]   (the G heads are consecutive rows of q)
    float *red = qs + G * hd;                    // [G][32]
    float *part = red + G * 32;                  // [nwarps][G][hd]
    float *sc = part + (size_t)nwarps * G * hd;  // [G][n_pad]
    int *pbase = (int *)(sc + (size_t)G * n_pad);
    const int n_pages = (n_kv + kv.page_size - 1) / kv.page_size;
    for (int i = tid; i < G * hd; i += nthreads) qs[i] = __half2float(__float2half_rn(__ldcg(qrows + i)));
    for (int i = tid; i < n_pages; i += nthreads) pbase[i] = pt[i];
    att_sync(nthreads);

    // ---- pass 1: scores of the G heads; one position per thread, its K row loaded once
    float lmax[G];
#pragma unroll
    for (int g = 0; g < G; g++) lmax[g] = -INFINITY;
    for (int j = tid; j < n_kv; j += nthreads) {
        const __half *kr = kv.k_pool + ((int64_t)pbase[j / kv.page_size] * kv.page_size + j % kv.page_size) * kv.kvd + hk * hd;
        float a[G];
#pragma unroll
        for (int g = 0; g < G; g++) a[g] = 0.f;
        for (int i0 = 0; i0 < hd; i0 += 64) { // 8 x 16-byte loads in flight per 64 dims
            int4 kq[8];
#pragma unroll
            for (int u = 0; u < 8; u++) kq[u] = *(const int4 *)(kr + min(i0 + u * 8, hd - 8)); // clamped, unconditional
#pragma unroll
            for (int u = 0; u < 8; u++) {
                if (i0 + u * 8 < hd) {
                    const __half2 *kh = (const __half2 *)&kq[u];
#pragma unroll
                    for (int k = 0; k < 4; k++) {
                        const float2 kk = __half22float2(kh[k]);
#pragma unroll
                        for (int g = 0; g < G; g++) {
                            a[g] += kk.x * qs[g * hd + i0 + u * 8 + 2 * k];
                            a[g] += kk.y * qs[g * hd + i0 + u * 8 + 2 * k + 1];
                        }
                    }
                }
            }
        }
#pragma unroll
        for (int g = 0; g < G; g++) {
            const float v = a[g] * scale;
            sc[g * n_pad + j] = v;
            lmax[g] = fmaxf(lmax[g], v);
        }
    }
#pragma unroll
    for (int g = 0; g < G; g++) {
        const float m = warp_max(lmax[g]);
        if (lane == 0) red[g * 32 + warp] = m;
    }
    att_sync(nthreads);
    float mx[G], lsum[G];
#pragma unroll
    for (int g = 0; g < G; g++) {
        mx[g] = -INFINITY;
        for (int i = 0; i < nwarps; i++) mx[g] = fmaxf(mx[g], red[g * 32 + i]);
        lsum[g] = 0.f;
    }
    att_sync(nthreads);
    for (int j = tid; j < n_kv; j += nthreads) {
#pragma unroll
        for (int g = 0; g < G; g++) {
            const float e = expf(sc[g * n_pad + j] - mx[g]);
            sc[g * n_pad + j] = e;
            lsum[g] += e;
        }
    }
#pragma unroll
    for (int g = 0; g < G; g++) {
        const float t = warp_sum(lsum[g]);
        if (lane == 0) red[g * 32 + warp] = t;
    }
    att_sync(nthreads);
    float inv[G];
#pragma unroll
    for (int g = 0; g < G; g++) {
        double tot = 0.0;
        for (int i = 0; i < nwarps; i++) tot += (double)red[g * 32 + i];
        inv[g] = (float)(1.0 / tot);
    }
    for (int j = tid; j < n_kv; j += nthreads) {
#pragma unroll
        for (int g = 0; g < G; g++) sc[g * n_pad + j] = __half2float(__float2half_rn(__fmul_rn(sc[g * n_pad + j], inv[g])));
    }
    att_sync(nthreads);
    // ---- pass 2: out_g = sum_j p_gj v_j ; lane owns dims [lane*dpl, +dpl), warps split positions, V row loaded once
    const int dpl = hd / 32; // 2, 4 or 8 halves per lane
    float acc[G][8];
#pragma unroll
    for (int g = 0; g < G; g++)
#pragma unroll
        for (int i = 0; i < 8; i++) acc[g][i] = 0.f;
    constexpr int PP = 4; // V rows in flight per warp
    for (int j0 = warp * PP; j0 < n_kv; j0 += nwarps * PP) {
        uint32_t raw[PP][4];
        int jj[PP];
#pragma unroll
        for (int u = 0; u < PP; u++) {
            const int j = j0 + u;
            jj[u] = j < n_kv ? j : -1;
            const int jc = j < n_kv ? j : 0;
            const __half *vr = kv.v_pool + ((int64_t)pbase[jc / kv.page_size] * kv.page_size + jc % kv.page_size) * kv.kvd + hk * hd +
                               lane * dpl;
            if (dpl == 8) { const int4 v = *(const int4 *)vr; raw[u][0] = v.x; raw[u][1] = v.y; raw[u][2] = v.z; raw[u][3] = v.w; }
            else if (dpl == 4) { const int2 v = *(const int2 *)vr; raw[u][0] = v.x; raw[u][1] = v.y; raw[u][2] = 0; raw[u][3] = 0; }
            else { raw[u][0] = *(const uint32_t *)vr; raw[u][1] = 0; raw[u][2] = 0; raw[u][3] = 0; }
        }
#pragma unroll
        for (int u = 0; u < PP; u++) {
#pragma unroll
            for (int g = 0; g < G; g++) {
                const float p = jj[u] >= 0 ? sc[g * n_pad + jj[u]] : 0.f;
#pragma unroll
                for (int k = 0; k < 4; k++) {
                    if (2 * k < dpl) {
                        const float2 vv = __half22float2(*(const __half2 *)&raw[u][k]);
                        acc[g][2 * k] += vv.x * p;
                        acc[g][2 * k + 1] += vv.y * p;
                    }
                }
            }
        }
    }
#pragma unroll
    for (int g = 0; g < G; g++)
#pragma unroll
        for (int i = 0; i < 8; i++)
            if (i < dpl) part[((size_t)warp * G + g) * hd + lane * dpl + i] = acc[g][i];
    att_sync(nthreads);
    for (int i = tid; i < G * hd; i += nthreads) { // i = g * hd + dim; the G output rows are consecutive
        float o = 0.f;
        for (int w = 0; w < nwarps; w++) o += part[(size_t)w * G * hd + i];
        if (x16) x16[qg_x_offset_of_element(x16_tok, x16_k0 + i, x16_n_pad)] = __float2half_rn(fminf(fmaxf(o, -65504.f), 65504.f));
        else orows[i] = o;
    }
    att_sync(nthreads);
}

} // namespace b2
