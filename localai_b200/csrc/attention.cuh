// attention.cuh -- paged-KV attention for one (head, query token), callable by any group of `nthreads` threads
// (a whole CTA in attention_kernel, the CTA of the decode mega-kernel).  See kernels_misc.cu for the description.
#pragma once
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

} // namespace b2
