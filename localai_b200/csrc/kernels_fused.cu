// kernels_fused.cu -- the decode-path mat-vec with everything around it fused in.
//
// One launch = [prologue] RMSNorm(x)*w (optional) -> activation quantisation (Q8_K / Q8_0 / f16, bit-exact with the
// stand-alone kernels) into shared memory, done redundantly by every CTA (x is KBs, L2-resident)
//            [main]     streaming dequant mat-vec over up to three weight matrices that share the activation;
//                       a warp owns TWO rows at a time (shared activation loads, 2x loads in flight) with a
//                       register double buffer that prefetches the next block group while the current one is reduced
//            [epilogue] plain (+residual) | SwiGLU (gate row r, up row r -> silu(g)*u) |
//                       QKV: RoPE on adjacent row pairs, q -> fp32, k/v -> F16 paged KV cache.
// A decode layer is then 5 launches: qkv, attention, wo, gate/up, down.  HBM-bound; algorithmic bytes per launch =
// sum(rows x row_bytes) of its matrices.
#include "act_quant.cuh"
#include "kernels.h"

namespace b2 {

constexpr int FW = 8;         // warps per CTA
constexpr int FUSED_MAX_NT = 4;

// ------------------------------------------------------------------------------------------------ two-row dots
// Each returns this lane's partial sums acc[r][t]; the caller warp-reduces.

template <int NT>
__device__ __forceinline__ void rows2_q4_K(const uint8_t *r0, const uint8_t *r1, int nb, int64_t K, const uint8_t *xs, int64_t xrb,
                                           int lane, float (&acc)[2][NT]) {
    const int p = lane & 7, j = p >> 1, off = (p & 1) * 16;
    const uint8_t *rows[2] = {r0, r1};
    int ib = lane >> 3;
    int4 hdr[2], qv[2];
    if (ib < nb) {
#pragma unroll
        for (int r = 0; r < 2; r++) {
            hdr[r] = ld_stream(rows[r] + (size_t)ib * 144);
            qv[r] = ld_stream(rows[r] + (size_t)ib * 144 + 16 + p * 16);
        }
    }
    for (; ib < nb; ib += 4) {
        int4 nh[2], nq[2];
        const int nib = ib + 4;
        if (nib < nb) {
#pragma unroll
            for (int r = 0; r < 2; r++) {
                nh[r] = ld_stream(rows[r] + (size_t)nib * 144);
                nq[r] = ld_stream(rows[r] + (size_t)nib * 144 + 16 + p * 16);
            }
        }
        float d[2], dmin[2];
        int sc_lo[2], sc_hi[2], m_p[2];
#pragma unroll
        for (int r = 0; r < 2; r++) {
            int m_lo, m_hi, sc_p;
            d[r] = h2f((uint16_t)(hdr[r].x & 0xFFFF));
            dmin[r] = h2f((uint16_t)((uint32_t)hdr[r].x >> 16));
            get_scale_min_k4_w(2 * j, hdr[r].y, hdr[r].z, hdr[r].w, sc_lo[r], m_lo);
            get_scale_min_k4_w(2 * j + 1, hdr[r].y, hdr[r].z, hdr[r].w, sc_hi[r], m_hi);
            get_scale_min_k4_w(p, hdr[r].y, hdr[r].z, hdr[r].w, sc_p, m_p[r]);
        }
#pragma unroll
        for (int t = 0; t < NT; t++) {
            const uint8_t *xt = xs + t * xrb;
            const int8_t *xq = (const int8_t *)xt + (size_t)ib * 256 + j * 64 + off;
            const int4 xl = *(const int4 *)xq, xh = *(const int4 *)(xq + 32);
            const int xlo[4] = {xl.x, xl.y, xl.z, xl.w}, xhi[4] = {xh.x, xh.y, xh.z, xh.w};
            const float dx = ((const float *)(xt + K))[ib];
            const int16_t *bs = (const int16_t *)(xt + K + (K >> 6)) + ib * 16;
            const int bsum = (int)bs[2 * p] + (int)bs[2 * p + 1];
#pragma unroll
            for (int r = 0; r < 2; r++) {
                const uint32_t q[4] = {(uint32_t)qv[r].x, (uint32_t)qv[r].y, (uint32_t)qv[r].z, (uint32_t)qv[r].w};
                int s_lo = 0, s_hi = 0;
#pragma unroll
                for (int i = 0; i < 4; i++) {
                    s_lo = __dp4a((int)(q[i] & 0x0F0F0F0Fu), xlo[i], s_lo);
                    s_hi = __dp4a((int)((q[i] >> 4) & 0x0F0F0F0Fu), xhi[i], s_hi);
                }
                acc[r][t] += (d[r] * dx) * (float)(sc_lo[r] * s_lo + sc_hi[r] * s_hi) - (dmin[r] * dx) * (float)(m_p[r] * bsum);
            }
        }
#pragma unroll
        for (int r = 0; r < 2; r++) { hdr[r] = nh[r]; qv[r] = nq[r]; }
    }
}

template <int NT>
__device__ __forceinline__ void rows2_q5_K(const uint8_t *r0, const uint8_t *r1, int nb, int64_t K, const uint8_t *xs, int64_t xrb,
                                           int lane, float (&acc)[2][NT]) {
    const int p = lane & 7, j = p >> 1, off = (p & 1) * 16;
    const uint8_t *rows[2] = {r0, r1};
    for (int ib = lane >> 3; ib < nb; ib += 4) {
        int4 hdr[2], hv[2], qv[2];
#pragma unroll
        for (int r = 0; r < 2; r++) {
            const uint8_t *blk = rows[r] + (size_t)ib * 176;
            hdr[r] = ld_stream(blk);
            hv[r] = ld_stream(blk + 16 + off);
            qv[r] = ld_stream(blk + 48 + p * 16);
        }
        float d[2], dmin[2];
        int sc_lo[2], sc_hi[2], m_p[2];
        uint32_t wl[2][4], wh[2][4];
#pragma unroll
        for (int r = 0; r < 2; r++) {
            int m_lo, m_hi, sc_p;
            d[r] = h2f((uint16_t)(hdr[r].x & 0xFFFF));
            dmin[r] = h2f((uint16_t)((uint32_t)hdr[r].x >> 16));
            get_scale_min_k4_w(2 * j, hdr[r].y, hdr[r].z, hdr[r].w, sc_lo[r], m_lo);
            get_scale_min_k4_w(2 * j + 1, hdr[r].y, hdr[r].z, hdr[r].w, sc_hi[r], m_hi);
            get_scale_min_k4_w(p, hdr[r].y, hdr[r].z, hdr[r].w, sc_p, m_p[r]);
            const uint32_t q[4] = {(uint32_t)qv[r].x, (uint32_t)qv[r].y, (uint32_t)qv[r].z, (uint32_t)qv[r].w};
            const uint32_t h[4] = {(uint32_t)hv[r].x, (uint32_t)hv[r].y, (uint32_t)hv[r].z, (uint32_t)hv[r].w};
#pragma unroll
            for (int i = 0; i < 4; i++) {
                wl[r][i] = (q[i] & 0x0F0F0F0Fu) | (((h[i] >> (2 * j)) & 0x01010101u) << 4);
                wh[r][i] = ((q[i] >> 4) & 0x0F0F0F0Fu) | (((h[i] >> (2 * j + 1)) & 0x01010101u) << 4);
            }
        }
#pragma unroll
        for (int t = 0; t < NT; t++) {
            const uint8_t *xt = xs + t * xrb;
            const int8_t *xq = (const int8_t *)xt + (size_t)ib * 256 + j * 64 + off;
            const int4 xl = *(const int4 *)xq, xh = *(const int4 *)(xq + 32);
            const int xlo[4] = {xl.x, xl.y, xl.z, xl.w}, xhi[4] = {xh.x, xh.y, xh.z, xh.w};
            const float dx = ((const float *)(xt + K))[ib];
            const int16_t *bs = (const int16_t *)(xt + K + (K >> 6)) + ib * 16;
            const int bsum = (int)bs[2 * p] + (int)bs[2 * p + 1];
#pragma unroll
            for (int r = 0; r < 2; r++) {
                int s_lo = 0, s_hi = 0;
#pragma unroll
                for (int i = 0; i < 4; i++) {
                    s_lo = __dp4a((int)wl[r][i], xlo[i], s_lo);
                    s_hi = __dp4a((int)wh[r][i], xhi[i], s_hi);
                }
                acc[r][t] += (d[r] * dx) * (float)(sc_lo[r] * s_lo + sc_hi[r] * s_hi) - (dmin[r] * dx) * (float)(m_p[r] * bsum);
            }
        }
    }
}

template <int NT>
__device__ __forceinline__ void rows2_q6_K(const uint8_t *r0, const uint8_t *r1, int nb, int64_t K, const uint8_t *xs, int64_t xrb,
                                           int lane, float (&acc)[2][NT]) {
    const Q6Row w[2] = {Q6Row(r0, nb), Q6Row(r1, nb)};
    const int p = lane & 7, n = p >> 2, c = p & 3;
    const int sh = 2 * (c >> 1);
    int ib = lane >> 3;
    int4 lv[2], hv[2];
    uint32_t scw[2]; // the two int8 scales this lane needs, packed (lo | hi << 8)
    uint16_t dh[2];
    auto load = [&](int b, int4(&l)[2], int4(&h)[2], uint32_t(&s)[2], uint16_t(&dd)[2]) {
#pragma unroll
        for (int r = 0; r < 2; r++) {
            l[r] = ld_stream(w[r].ql + (size_t)b * 128 + p * 16);
            h[r] = ld_stream(w[r].qh + (size_t)b * 64 + n * 32 + (c & 1) * 16);
            const int8_t *sc = w[r].sc + b * 16 + 8 * n + c;
            s[r] = (uint32_t)(uint8_t)sc[0] | ((uint32_t)(uint8_t)sc[4] << 8);
            dd[r] = w[r].d[b];
        }
    };
    if (ib < nb) load(ib, lv, hv, scw, dh);
    for (; ib < nb; ib += 4) {
        int4 nl[2], nh[2];
        uint32_t ns[2];
        uint16_t nd[2];
        if (ib + 4 < nb) load(ib + 4, nl, nh, ns, nd);
        uint32_t wl[2][4], wh[2][4];
        float d[2];
        int sc_lo[2], sc_hi[2];
#pragma unroll
        for (int r = 0; r < 2; r++) {
            const uint32_t l[4] = {(uint32_t)lv[r].x, (uint32_t)lv[r].y, (uint32_t)lv[r].z, (uint32_t)lv[r].w};
            const uint32_t h[4] = {(uint32_t)hv[r].x, (uint32_t)hv[r].y, (uint32_t)hv[r].z, (uint32_t)hv[r].w};
#pragma unroll
            for (int i = 0; i < 4; i++) {
                wl[r][i] = (l[i] & 0x0F0F0F0Fu) | (((h[i] >> sh) & 0x03030303u) << 4);
                wh[r][i] = ((l[i] >> 4) & 0x0F0F0F0Fu) | (((h[i] >> (sh + 4)) & 0x03030303u) << 4);
            }
            d[r] = h2f(dh[r]);
            sc_lo[r] = (int)(int8_t)(scw[r] & 0xFF);
            sc_hi[r] = (int)(int8_t)((scw[r] >> 8) & 0xFF);
        }
#pragma unroll
        for (int t = 0; t < NT; t++) {
            const uint8_t *xt = xs + t * xrb;
            const int8_t *xq = (const int8_t *)xt + (size_t)ib * 256 + 128 * n + 16 * c;
            const int4 xl = *(const int4 *)xq, xh = *(const int4 *)(xq + 64);
            const int xlo[4] = {xl.x, xl.y, xl.z, xl.w}, xhi[4] = {xh.x, xh.y, xh.z, xh.w};
            const float dx = ((const float *)(xt + K))[ib];
            const int16_t *bs = (const int16_t *)(xt + K + (K >> 6)) + ib * 16;
            const int b_lo = 32 * (int)bs[8 * n + c], b_hi = 32 * (int)bs[8 * n + 4 + c];
#pragma unroll
            for (int r = 0; r < 2; r++) {
                int s_lo = 0, s_hi = 0;
#pragma unroll
                for (int i = 0; i < 4; i++) {
                    s_lo = __dp4a((int)wl[r][i], xlo[i], s_lo);
                    s_hi = __dp4a((int)wh[r][i], xhi[i], s_hi);
                }
                acc[r][t] += (d[r] * dx) * (float)(sc_lo[r] * (s_lo - b_lo) + sc_hi[r] * (s_hi - b_hi));
            }
        }
#pragma unroll
        for (int r = 0; r < 2; r++) { lv[r] = nl[r]; hv[r] = nh[r]; scw[r] = ns[r]; dh[r] = nd[r]; }
    }
}

template <int NT>
__device__ __forceinline__ void rows2_q8_0(const uint8_t *r0, const uint8_t *r1, int64_t K, const uint8_t *xs, int64_t xrb, int lane,
                                           float (&acc)[2][NT]) {
    const Q80Row w[2] = {Q80Row(r0, K), Q80Row(r1, K)};
    const int n16 = (int)(K / 16);
    for (int i0 = 0; i0 < n16; i0 += 64) { // two 16-byte pieces per lane per trip; warp-uniform trip count
        int4 wv[2][2];
        float dw[2][2];
        bool valid[2];
#pragma unroll
        for (int u = 0; u < 2; u++) {
            const int i = i0 + u * 32 + lane;
            valid[u] = i < n16;
#pragma unroll
            for (int r = 0; r < 2; r++) {
                wv[r][u] = valid[u] ? ld_stream(w[r].qs + (size_t)i * 16) : make_int4(0, 0, 0, 0);
                dw[r][u] = valid[u] ? h2f(w[r].d[i >> 1]) : 0.f;
            }
        }
#pragma unroll
        for (int u = 0; u < 2; u++) {
            const int i = i0 + u * 32 + lane;
#pragma unroll
            for (int t = 0; t < NT; t++) {
                const uint8_t *xt = xs + t * xrb;
                const int4 xv = valid[u] ? *(const int4 *)(xt + (size_t)i * 16) : make_int4(0, 0, 0, 0);
                const float dx = valid[u] ? ((const float *)(xt + K))[i >> 1] : 0.f;
#pragma unroll
                for (int r = 0; r < 2; r++) {
                    int s = __dp4a(wv[r][u].x, xv.x, 0);
                    s = __dp4a(wv[r][u].y, xv.y, s);
                    s = __dp4a(wv[r][u].z, xv.z, s);
                    s = __dp4a(wv[r][u].w, xv.w, s);
                    s += __shfl_xor_sync(0xffffffffu, s, 1); // whole 32-element block, as on the CPU
                    if ((lane & 1) == 0) acc[r][t] += (float)s * (dw[r][u] * dx);
                }
            }
        }
    }
}

template <int NT>
__device__ __forceinline__ void rows2_f16(const uint8_t *r0, const uint8_t *r1, int64_t K, const uint8_t *xs, int64_t xrb, int lane,
                                          float (&acc)[2][NT]) {
    const uint8_t *rows[2] = {r0, r1};
    const int n8 = (int)(K / 8);
    for (int i = lane; i < n8; i += 32) {
        int4 wv[2];
#pragma unroll
        for (int r = 0; r < 2; r++) wv[r] = ld_stream(rows[r] + (size_t)i * 16);
#pragma unroll
        for (int t = 0; t < NT; t++) {
            const int4 xv = *(const int4 *)(xs + t * xrb + (size_t)i * 16);
            const __half2 *xh = (const __half2 *)&xv;
#pragma unroll
            for (int r = 0; r < 2; r++) {
                const __half2 *wh = (const __half2 *)&wv[r];
#pragma unroll
                for (int k = 0; k < 4; k++) {
                    const float2 a = __half22float2(wh[k]), b = __half22float2(xh[k]);
                    acc[r][t] += a.x * b.x;
                    acc[r][t] += a.y * b.y;
                }
            }
        }
    }
}

template <int NT>
__device__ __forceinline__ void rows2_f32(const uint8_t *r0, const uint8_t *r1, int64_t K, const uint8_t *xs, int64_t xrb, int lane,
                                          float (&acc)[2][NT]) {
    const uint8_t *rows[2] = {r0, r1};
    const int n4 = (int)(K / 4);
    for (int i = lane; i < n4; i += 32) {
#pragma unroll
        for (int t = 0; t < NT; t++) {
            const float4 xv = *(const float4 *)(xs + t * xrb + (size_t)i * 16);
#pragma unroll
            for (int r = 0; r < 2; r++) {
                const int4 wv = ld_stream(rows[r] + (size_t)i * 16);
                const float *wf = (const float *)&wv;
                acc[r][t] += wf[0] * xv.x;
                acc[r][t] += wf[1] * xv.y;
                acc[r][t] += wf[2] * xv.z;
                acc[r][t] += wf[3] * xv.w;
            }
        }
    }
}

// ------------------------------------------------------------------------------------------------ the kernel
template <int NT>
__global__ void __launch_bounds__(FW * 32)
fused_gemv_kernel(const FusedGemv a) {
    extern __shared__ __align__(16) uint8_t xs[];
    __shared__ double s_part[FW];
    __shared__ float s_scale;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t K = a.K;
    const int nb = (int)(K / 256);

    // ---- prologue: (RMSNorm ->) quantise the NT activation rows into shared memory
    if (a.xq_pre) {
        const int n16 = (int)(a.xrb * NT / 16);
        const int4 *src = (const int4 *)a.xq_pre;
        int4 *dst = (int4 *)xs;
        for (int i = threadIdx.x; i < n16; i += blockDim.x) dst[i] = src[i];
    } else {
        for (int t = 0; t < NT; t++) {
            const float *xr = a.x + (int64_t)(a.x_rows ? a.x_rows[t] : t) * a.x_stride;
            float scale = 1.f;
            if (a.norm_w) {
                double sum = 0.0;
                for (int i = threadIdx.x; i < (int)K; i += blockDim.x) {
                    const float v = xr[i];
                    sum += (double)__fmul_rn(v, v);
                }
                sum = warp_sum_d(sum);
                if (lane == 0) s_part[warp] = sum;
                __syncthreads();
                if (threadIdx.x == 0) {
                    double tot = 0.0;
                    for (int i = 0; i < FW; i++) tot += s_part[i];
                    const float mean = (float)(tot / (double)K);
                    s_scale = __fdiv_rn(1.0f, __fsqrt_rn(__fadd_rn(mean, a.eps)));
                }
                __syncthreads();
                scale = s_scale;
            }
            uint8_t *row = xs + t * a.xrb;
            for (int b = warp; b < nb; b += FW) {
                const float *xb = xr + b * 256 + lane * 8;
                const float4 a0 = *(const float4 *)xb, a1 = *(const float4 *)(xb + 4);
                float v[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
                if (a.norm_w) {
                    const float *wb = a.norm_w + b * 256 + lane * 8;
                    const float4 w0 = *(const float4 *)wb, w1 = *(const float4 *)(wb + 4);
                    const float w[8] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w};
#pragma unroll
                    for (int i = 0; i < 8; i++) v[i] = __fmul_rn(__fmul_rn(v[i], scale), w[i]);
                    if (a.xn_out && blockIdx.x == 0) { // the normalised row itself (embeddings / debugging)
                        float *o = a.xn_out + (int64_t)t * K + b * 256 + lane * 8;
                        *(float4 *)o = make_float4(v[0], v[1], v[2], v[3]);
                        *(float4 *)(o + 4) = make_float4(v[4], v[5], v[6], v[7]);
                    }
                }
                quant_block(a.fmt, v, lane, row, K, b);
            }
        }
    }
    __syncthreads();

    // ---- main loop over row-pair tasks
    const int n_tasks = a.n_tasks;
    for (int task = blockIdx.x * FW + warp; task < n_tasks; task += gridDim.x * FW) {
        int jb = 0, r_local;
        const uint8_t *r0, *r1;
        if (a.epi == EPI_SWIGLU) {
            r_local = task;
            r0 = a.job[0].W + (int64_t)task * a.job[0].row_stride;
            r1 = a.job[1].W + (int64_t)task * a.job[1].row_stride;
        } else {
            int row = task * 2;
            while (jb + 1 < a.n_jobs && row >= a.job[jb].rows) { row -= a.job[jb].rows; jb++; }
            r_local = row;
            r0 = a.job[jb].W + (int64_t)row * a.job[jb].row_stride;
            r1 = (row + 1 < a.job[jb].rows) ? r0 + a.job[jb].row_stride : r0;
        }
        float acc[2][NT];
#pragma unroll
        for (int r = 0; r < 2; r++)
#pragma unroll
            for (int t = 0; t < NT; t++) acc[r][t] = 0.f;
        switch (a.job[jb].type) {
            case B2_T_Q4_K: rows2_q4_K<NT>(r0, r1, nb, K, xs, a.xrb, lane, acc); break;
            case B2_T_Q5_K: rows2_q5_K<NT>(r0, r1, nb, K, xs, a.xrb, lane, acc); break;
            case B2_T_Q6_K: rows2_q6_K<NT>(r0, r1, nb, K, xs, a.xrb, lane, acc); break;
            case B2_T_Q8_0: rows2_q8_0<NT>(r0, r1, K, xs, a.xrb, lane, acc); break;
            case B2_T_F16: rows2_f16<NT>(r0, r1, K, xs, a.xrb, lane, acc); break;
            default: rows2_f32<NT>(r0, r1, K, xs, a.xrb, lane, acc); break;
        }
#pragma unroll
        for (int t = 0; t < NT; t++) {
            const float v0 = warp_sum(acc[0][t]), v1 = warp_sum(acc[1][t]);
            if (lane != 0) continue;
            if (a.epi == EPI_PLAIN) {
                float *o = a.out + (int64_t)t * a.out_stride + r_local;
                const float *res = a.residual ? a.residual + (int64_t)t * a.res_stride + r_local : nullptr;
                o[0] = res ? v0 + res[0] : v0;
                if (r_local + 1 < a.job[0].rows) o[1] = res ? v1 + res[1] : v1;
            } else if (a.epi == EPI_SWIGLU) {
                const float s = __fdiv_rn(v0, __fadd_rn(1.0f, expf(-v0)));
                a.out[(int64_t)t * a.out_stride + r_local] = __fmul_rn(s, v1);
            } else { // EPI_QKV: jb 0 = q, 1 = k, 2 = v; (r_local, r_local + 1) is a rotary pair inside one head
                const int pos = a.pos[t], sl = a.slot[t];
                const int page = a.kv.page_table[(int64_t)sl * a.kv.max_pages + pos / a.kv.page_size];
                const int64_t cell = ((int64_t)page * a.kv.page_size + pos % a.kv.page_size) * a.kv.kvd;
                if (jb == 2) {
                    *(__half2 *)(a.kv.v_pool + cell + r_local) = __halves2half2(__float2half_rn(v0), __float2half_rn(v1));
                } else {
                    const float *cs = a.rope_table + (int64_t)pos * a.hd + (r_local % a.hd);
                    const float c = cs[0], sn = cs[1];
                    const float o0 = __fsub_rn(__fmul_rn(v0, c), __fmul_rn(v1, sn));
                    const float o1 = __fadd_rn(__fmul_rn(v0, sn), __fmul_rn(v1, c));
                    if (jb == 0) {
                        float *q = a.out + (int64_t)t * a.out_stride + r_local;
                        q[0] = o0;
                        q[1] = o1;
                    } else {
                        *(__half2 *)(a.kv.k_pool + cell + r_local) = __halves2half2(__float2half_rn(o0), __float2half_rn(o1));
                    }
                }
            }
        }
    }
}

void launch_fused_gemv(FusedGemv a, int n_tok, cudaStream_t s) {
    if (n_tok <= 0) return;
    B2_CHECK(a.K % 256 == 0, B2_EINVAL, "fused gemv: K must be a multiple of 256");
    a.xrb = act_row_bytes((ActFmt)a.fmt, a.K);
    int64_t rows = 0;
    for (int j = 0; j < a.n_jobs; j++) {
        B2_CHECK(act_format_for(a.job[j].type) == (ActFmt)a.fmt, B2_EINVAL, "fused gemv: jobs need one activation format");
        B2_CHECK(a.epi == EPI_SWIGLU || a.job[j].rows % 2 == 0 || a.n_jobs == 1, B2_EINVAL, "fused gemv: odd row count in a multi-matrix job");
        rows += a.job[j].rows;
    }
    if (a.epi == EPI_SWIGLU) {
        B2_CHECK(a.n_jobs == 2 && a.job[0].rows == a.job[1].rows && a.job[0].type == a.job[1].type, B2_EINVAL, "fused gemv: bad SwiGLU jobs");
        a.n_tasks = (int)a.job[0].rows;
    } else {
        a.n_tasks = (int)((rows + 1) / 2);
    }
    int ctas = (a.n_tasks + FW - 1) / FW;
    ctas = std::min(ctas, 148 * 4);
    const float *x0 = a.x;
    const int32_t *xr0 = a.x_rows;
    const uint8_t *xq0 = a.xq_pre;
    float *out0 = a.out, *xn0 = a.xn_out;
    const float *res0 = a.residual;
    const int32_t *pos0 = a.pos, *slot0 = a.slot;
    for (int t0 = 0; t0 < n_tok; t0 += FUSED_MAX_NT) {
        const int nt = std::min(FUSED_MAX_NT, n_tok - t0);
        const size_t smem = (size_t)a.xrb * nt;
        B2_CHECK(smem <= 200 * 1024, B2_EUNSUPPORTED, "fused gemv: activation rows exceed shared memory");
        a.x = (x0 && !xr0) ? x0 + (int64_t)t0 * a.x_stride : x0;
        a.x_rows = xr0 ? xr0 + t0 : nullptr;
        a.xq_pre = xq0 ? xq0 + (int64_t)t0 * a.xrb : nullptr;
        a.out = out0 + (int64_t)t0 * a.out_stride;
        a.xn_out = xn0 ? xn0 + (int64_t)t0 * a.K : nullptr;
        a.residual = res0 ? res0 + (int64_t)t0 * a.res_stride : nullptr;
        a.pos = pos0 ? pos0 + t0 : nullptr;
        a.slot = slot0 ? slot0 + t0 : nullptr;
#define B2_FUSED_CASE(N)                                                                                                  \
    case N: {                                                                                                             \
        auto kfn = fused_gemv_kernel<N>;                                                                                  \
        if (smem > 48 * 1024) B2_CUDA(cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024)); \
        kfn<<<ctas, FW * 32, smem, s>>>(a);                                                                               \
    } break;
        switch (nt) {
            B2_FUSED_CASE(1)
            B2_FUSED_CASE(2)
            B2_FUSED_CASE(3)
            B2_FUSED_CASE(4)
        }
#undef B2_FUSED_CASE
        g_launch_count++;
        B2_CUDA(cudaGetLastError());
    }
}

} // namespace b2
