// gemv_rows2.cuh -- "one warp x TWO weight rows x NT activation rows" block dot products over a block range
// [b0, b1) of the rows.  Two rows share every activation load from shared memory and keep twice the weight loads in
// flight.  Arithmetic = ggml_vec_dot_{q4_K,q5_K,q6_K}_q8_K / q8_0_q8_0 / f16 (integer dots per block with dp4a,
// fp32 scaling); only the fp32 summation ORDER differs from the scalar reference.
//
// Instruction budget: at the HBM roofline an SM must retire ~23 weight bytes per clock, i.e. it can spend at most
// ~5.6 lane-instructions per weight byte.  The K-quant routines are therefore written around FOUR lanes per
// 256-weight super-block (64 weights = 32 B of nibbles per lane and row: the per-lane scale/min unpacking is
// amortised over 16 dp4a), masked-nibble dp4a (high nibbles are used in place as 16*q and the sum is shifted
// once), and activation rows whose int8 section is chunk-swizzled in shared memory (act_swz) so the stride-64-byte
// reads are bank-conflict free.  ~2 lane-instructions per weight byte.
//
// LD selects where the weight rows live: LdGlobal streams them from HBM (128-bit, L1 no-allocate), LdShared reads
// rows that a bulk-async copy (TMA) has staged in shared memory (decode mega-kernel).
#pragma once
#include "act_quant.cuh"
#include "common.h"
#include "quant.cuh"

namespace b2 {

struct LdGlobal {
    static constexpr bool kPrefetch = true;  // keep the next block group's loads in flight in registers
    static __device__ __forceinline__ int4 v4(const void *p) { return ld_stream(p); }
    static __device__ __forceinline__ int2 v2(const void *p) { return *(const int2 *)p; }
    static __device__ __forceinline__ uint32_t u16(const void *p) { return *(const uint16_t *)p; }
};
struct LdShared {
    static constexpr bool kPrefetch = false; // shared-memory latency is short: no register double buffer
    static __device__ __forceinline__ int4 v4(const void *p) {
        int4 r;
        asm volatile("ld.shared.v4.s32 {%0,%1,%2,%3}, [%4];"
                     : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                     : "r"((uint32_t)__cvta_generic_to_shared(p)));
        return r;
    }
    static __device__ __forceinline__ int2 v2(const void *p) { return *(const int2 *)p; }
    static __device__ __forceinline__ uint32_t u16(const void *p) { return *(const uint16_t *)p; }
};

// 16-byte chunk `c` of the (swizzled) int8 section of an activation row in shared memory
__device__ __forceinline__ int4 x_chunk(const uint8_t *xt, int c) { return *(const int4 *)(xt + ((size_t)act_swz(c) << 4)); }

__device__ __forceinline__ int dp4a_us(uint32_t a, int b, int c) { // 4 x (u8 * s8) + c
    int d;
    asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}

// Each function adds this lane's partial sums into acc[r][t]; the caller warp-reduces.

// ---------------------------------------------------------------------------------------------------- Q4_K
// lane = (block slot bs = lane >> 2, 64-weight group g = lane & 3); 8 blocks per warp pass.
template <class LD, int R, int NT>
__device__ __forceinline__ void rows2_q4_K(const uint8_t *r0, const uint8_t *r1, int b0, int b1, int64_t K, const uint8_t *xs,
                                           int64_t xrb, int lane, float (&acc)[R][NT]) {
    const int g = lane & 3;
    const uint8_t *rows[2] = {r0, r1}; // rows[1] unused when R == 1
    int ib = b0 + (lane >> 2);
    int4 hdr[2], qa[2], qb[2];
    if (LD::kPrefetch && ib < b1) {
#pragma unroll
        for (int r = 0; r < R; r++) {
            const uint8_t *blk = rows[r] + (size_t)ib * 144;
            hdr[r] = LD::v4(blk);
            qa[r] = LD::v4(blk + 16 + g * 32);
            qb[r] = LD::v4(blk + 32 + g * 32);
        }
    }
    for (; ib < b1; ib += 8) {
        int4 nh[2], na[2], nq[2];
        if (!LD::kPrefetch) {
#pragma unroll
            for (int r = 0; r < R; r++) {
                const uint8_t *blk = rows[r] + (size_t)ib * 144;
                hdr[r] = LD::v4(blk);
                qa[r] = LD::v4(blk + 16 + g * 32);
                qb[r] = LD::v4(blk + 32 + g * 32);
            }
        } else if (ib + 8 < b1) {
#pragma unroll
            for (int r = 0; r < R; r++) {
                const uint8_t *blk = rows[r] + (size_t)(ib + 8) * 144;
                nh[r] = LD::v4(blk);
                na[r] = LD::v4(blk + 16 + g * 32);
                nq[r] = LD::v4(blk + 32 + g * 32);
            }
        }
        // scales / mins of sub-blocks 2g, 2g+1 (ggml's kmask unpacking of the 12 packed bytes)
        float dl[2], dm[2];
        int sc0[2], sc1[2], m0[2], m1[2];
#pragma unroll
        for (int r = 0; r < R; r++) {
            const uint32_t u0 = hdr[r].y, u1 = hdr[r].z, u2 = hdr[r].w;
            const uint32_t s_lo = u0 & 0x3f3f3f3fu, m_lo = u1 & 0x3f3f3f3fu;
            const uint32_t s_hi = (u2 & 0x0f0f0f0fu) | (((u0 >> 6) & 0x03030303u) << 4);
            const uint32_t m_hi = ((u2 >> 4) & 0x0f0f0f0fu) | (((u1 >> 6) & 0x03030303u) << 4);
            const uint32_t sw = (g & 2) ? s_hi : s_lo, mw = (g & 2) ? m_hi : m_lo;
            const int sh = (g & 1) * 16;
            sc0[r] = (sw >> sh) & 0xFF;
            sc1[r] = (sw >> (sh + 8)) & 0xFF;
            m0[r] = (mw >> sh) & 0xFF;
            m1[r] = (mw >> (sh + 8)) & 0xFF;
            const float2 dd = __half22float2(*(const __half2 *)&hdr[r].x);
            dl[r] = dd.x;
            dm[r] = dd.y;
        }
#pragma unroll
        for (int t = 0; t < NT; t++) {
            const uint8_t *xt = xs + t * xrb;
            const int c = ib * 16 + g * 4; // first 16-byte chunk of this lane's 64 activations
            const int4 x0 = x_chunk(xt, c), x1 = x_chunk(xt, c + 1), x2 = x_chunk(xt, c + 2), x3 = x_chunk(xt, c + 3);
            const int xl[8] = {x0.x, x0.y, x0.z, x0.w, x1.x, x1.y, x1.z, x1.w};
            const int xh[8] = {x2.x, x2.y, x2.z, x2.w, x3.x, x3.y, x3.z, x3.w};
            const float dx = ((const float *)(xt + K))[ib];
            const int *bsp = (const int *)((const int16_t *)(xt + K + (K >> 6)) + ib * 16 + g * 4); // 4 x int16 (4-byte aligned)
            const int bx = bsp[0], by = bsp[1];
            const int bs0 = (int)(int16_t)(bx & 0xFFFF) + (bx >> 16), bs1 = (int)(int16_t)(by & 0xFFFF) + (by >> 16);
#pragma unroll
            for (int r = 0; r < R; r++) {
                const uint32_t q[8] = {(uint32_t)qa[r].x, (uint32_t)qa[r].y, (uint32_t)qa[r].z, (uint32_t)qa[r].w,
                                       (uint32_t)qb[r].x, (uint32_t)qb[r].y, (uint32_t)qb[r].z, (uint32_t)qb[r].w};
                int s_lo = 0, s_hi = 0;
#pragma unroll
                for (int i = 0; i < 8; i++) {
                    s_lo = dp4a_us(q[i] & 0x0F0F0F0Fu, xl[i], s_lo);
                    s_hi = dp4a_us(q[i] & 0xF0F0F0F0u, xh[i], s_hi); // 16 * (high nibbles . x)
                }
                const int isum = sc0[r] * s_lo + sc1[r] * (s_hi >> 4);
                const int imin = m0[r] * bs0 + m1[r] * bs1;
                acc[r][t] += (dl[r] * dx) * (float)isum - (dm[r] * dx) * (float)imin;
            }
        }
        if (LD::kPrefetch) {
#pragma unroll
            for (int r = 0; r < R; r++) { hdr[r] = nh[r]; qa[r] = na[r]; qb[r] = nq[r]; }
        }
    }
}

// ---------------------------------------------------------------------------------------------------- Q5_K
// 8 lanes per block (16 bytes of nibbles each), as Q4_K used to be; Q5_K only appears in 70B attn_v.
template <class LD, int R, int NT>
__device__ __forceinline__ void rows2_q5_K(const uint8_t *r0, const uint8_t *r1, int b0, int b1, int64_t K, const uint8_t *xs,
                                           int64_t xrb, int lane, float (&acc)[R][NT]) {
    const int p = lane & 7, j = p >> 1, off = (p & 1) * 16;
    const uint8_t *rows[2] = {r0, r1}; // rows[1] unused when R == 1
    for (int ib = b0 + (lane >> 3); ib < b1; ib += 4) {
        int4 hdr[2], hv[2], qv[2];
#pragma unroll
        for (int r = 0; r < R; r++) {
            const uint8_t *blk = rows[r] + (size_t)ib * 176;
            hdr[r] = LD::v4(blk);
            hv[r] = LD::v4(blk + 16 + off);
            qv[r] = LD::v4(blk + 48 + p * 16);
        }
        float d[2], dmin[2];
        int sc_lo[2], sc_hi[2], m_p[2];
        uint32_t wl[2][4], wh[2][4];
#pragma unroll
        for (int r = 0; r < R; r++) {
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
            const int c = ib * 16 + j * 4 + (p & 1);
            const int4 xl = x_chunk(xt, c), xh = x_chunk(xt, c + 2);
            const int xlo[4] = {xl.x, xl.y, xl.z, xl.w}, xhi[4] = {xh.x, xh.y, xh.z, xh.w};
            const float dx = ((const float *)(xt + K))[ib];
            const int16_t *bs = (const int16_t *)(xt + K + (K >> 6)) + ib * 16;
            const int bsum = (int)bs[2 * p] + (int)bs[2 * p + 1];
#pragma unroll
            for (int r = 0; r < R; r++) {
                int s_lo = 0, s_hi = 0;
#pragma unroll
                for (int i = 0; i < 4; i++) {
                    s_lo = dp4a_us(wl[r][i], xlo[i], s_lo);
                    s_hi = dp4a_us(wh[r][i], xhi[i], s_hi);
                }
                acc[r][t] += (d[r] * dx) * (float)(sc_lo[r] * s_lo + sc_hi[r] * s_hi) - (dmin[r] * dx) * (float)(m_p[r] * bsum);
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------------- Q6_K
// Device row layout: nb blocks of 208 B [ql 128][qh 64][scales 16], then d nb*2 (quant.cuh Q6Row); nb_row locates d.
// lane = (block slot lane >> 2, half n = (lane >> 1) & 1, column half lh = lane & 1): 64 weights per lane =
// the four 16-weight sub-blocks {128n + 32c + 16lh + [0,16)} c = 0..3, which share one 16-byte piece of qh.
template <class LD, int R, int NT>
__device__ __forceinline__ void rows2_q6_K(const uint8_t *r0, const uint8_t *r1, int nb_row, int b0, int b1, int64_t K,
                                           const uint8_t *xs, int64_t xrb, int lane, float (&acc)[R][NT]) {
    const Q6Row w[2] = {Q6Row(r0, nb_row), Q6Row(r1, nb_row)}; // w[1] unused when R == 1
    const int n = (lane >> 1) & 1, lh = lane & 1;
    int ib = b0 + (lane >> 2);
    int4 l0[2], l1[2], hq[2];
    int2 scv[2];
    uint32_t dh[2];
    auto load = [&](int b, int4(&a0)[2], int4(&a1)[2], int4(&h)[2], int2(&s)[2], uint32_t(&dd)[2]) {
#pragma unroll
        for (int r = 0; r < R; r++) {
            a0[r] = LD::v4(w[r].ql(b) + 64 * n + 16 * lh);      // c = 0 (low nibbles), c = 2 (high)
            a1[r] = LD::v4(w[r].ql(b) + 64 * n + 32 + 16 * lh); // c = 1 (low nibbles), c = 3 (high)
            h[r] = LD::v4(w[r].qh(b) + 32 * n + 16 * lh);
            s[r] = LD::v2(w[r].sc(b) + 8 * n);                  // 8 int8 scales of this half
            dd[r] = LD::u16(w[r].d + b);
        }
    };
    if (LD::kPrefetch && ib < b1) load(ib, l0, l1, hq, scv, dh);
    for (; ib < b1; ib += 8) {
        int4 n0[2], n1[2], nh[2];
        int2 ns[2];
        uint32_t nd[2];
        if (!LD::kPrefetch) load(ib, l0, l1, hq, scv, dh);
        else if (ib + 8 < b1) load(ib + 8, n0, n1, nh, ns, nd);
#pragma unroll
        for (int t = 0; t < NT; t++) {
            const uint8_t *xt = xs + t * xrb;
            const int c0 = ib * 16 + 8 * n + lh; // chunk of sub-block c: c0 + 2c
            const int4 xa = x_chunk(xt, c0), xb = x_chunk(xt, c0 + 2), xc = x_chunk(xt, c0 + 4), xd = x_chunk(xt, c0 + 6);
            const int xv[4][4] = {{xa.x, xa.y, xa.z, xa.w}, {xb.x, xb.y, xb.z, xb.w}, {xc.x, xc.y, xc.z, xc.w}, {xd.x, xd.y, xd.z, xd.w}};
            const float dx = ((const float *)(xt + K))[ib];
            const int16_t *bs = (const int16_t *)(xt + K + (K >> 6)) + ib * 16 + 8 * n + lh;
            const int bsum[4] = {bs[0], bs[2], bs[4], bs[6]};
#pragma unroll
            for (int r = 0; r < R; r++) {
                const uint32_t a0[4] = {(uint32_t)l0[r].x, (uint32_t)l0[r].y, (uint32_t)l0[r].z, (uint32_t)l0[r].w};
                const uint32_t a1[4] = {(uint32_t)l1[r].x, (uint32_t)l1[r].y, (uint32_t)l1[r].z, (uint32_t)l1[r].w};
                const uint32_t hh[4] = {(uint32_t)hq[r].x, (uint32_t)hq[r].y, (uint32_t)hq[r].z, (uint32_t)hq[r].w};
                int s[4] = {0, 0, 0, 0};
#pragma unroll
                for (int i = 0; i < 4; i++) {
                    const uint32_t h = hh[i];
                    s[0] = dp4a_us((a0[i] & 0x0F0F0F0Fu) | ((h << 4) & 0x30303030u), xv[0][i], s[0]);
                    s[1] = dp4a_us((a1[i] & 0x0F0F0F0Fu) | ((h << 2) & 0x30303030u), xv[1][i], s[1]);
                    s[2] = dp4a_us(((a0[i] >> 4) & 0x0F0F0F0Fu) | (h & 0x30303030u), xv[2][i], s[2]);
                    s[3] = dp4a_us(((a1[i] >> 4) & 0x0F0F0F0Fu) | ((h >> 2) & 0x30303030u), xv[3][i], s[3]);
                }
                // scales of sub-blocks 8n + 2c + lh: bytes lh, lh+2 of scv.x and scv.y
                const uint32_t sx = (uint32_t)scv[r].x >> (8 * lh), sy = (uint32_t)scv[r].y >> (8 * lh);
                const int k0 = (int)(int8_t)(sx & 0xFF), k1 = (int)(int8_t)((sx >> 16) & 0xFF);
                const int k2 = (int)(int8_t)(sy & 0xFF), k3 = (int)(int8_t)((sy >> 16) & 0xFF);
                // stored quants are q+32: subtract 32 * sum(x) of each 16-weight sub-block (exact, integer)
                const int isum = k0 * (s[0] - 32 * bsum[0]) + k1 * (s[1] - 32 * bsum[1]) + k2 * (s[2] - 32 * bsum[2]) +
                                 k3 * (s[3] - 32 * bsum[3]);
                acc[r][t] += (h2f((uint16_t)dh[r]) * dx) * (float)isum;
            }
        }
        if (LD::kPrefetch) {
#pragma unroll
            for (int r = 0; r < R; r++) { l0[r] = n0[r]; l1[r] = n1[r]; hq[r] = nh[r]; scv[r] = ns[r]; dh[r] = nd[r]; }
        }
    }
}

// ---------------------------------------------------------------------------------------------------- Q8_0
// rows are [qs: K x int8][d: K/32 x f16]; pieces of 16 bytes, [i0, i1) = piece range (2 pieces per block)
template <class LD, int R, int NT>
__device__ __forceinline__ void rows2_q8_0(const uint8_t *r0, const uint8_t *r1, int i0, int i1, int64_t K, const uint8_t *xs,
                                           int64_t xrb, int lane, float (&acc)[R][NT]) {
    const Q80Row w[2] = {Q80Row(r0, K), Q80Row(r1, K)}; // w[1] unused when R == 1
    for (int base = i0; base < i1; base += 64) { // two 16-byte pieces per lane per trip; warp-uniform trip count
        int4 wv[2][2];
        float dw[2][2];
        bool valid[2];
#pragma unroll
        for (int u = 0; u < 2; u++) {
            const int i = base + u * 32 + lane;
            valid[u] = i < i1;
#pragma unroll
            for (int r = 0; r < R; r++) {
                wv[r][u] = valid[u] ? LD::v4(w[r].qs + (size_t)i * 16) : make_int4(0, 0, 0, 0);
                dw[r][u] = valid[u] ? h2f((uint16_t)LD::u16(w[r].d + (i >> 1))) : 0.f;
            }
        }
#pragma unroll
        for (int u = 0; u < 2; u++) {
            const int i = base + u * 32 + lane;
#pragma unroll
            for (int t = 0; t < NT; t++) {
                const uint8_t *xt = xs + t * xrb;
                const int4 xv = valid[u] ? x_chunk(xt, i) : make_int4(0, 0, 0, 0);
                const float dx = valid[u] ? ((const float *)(xt + K))[i >> 1] : 0.f;
#pragma unroll
                for (int r = 0; r < R; r++) {
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

// F16 rows: pieces of 8 halves; [i0, i1) = piece range.  (f16 / f32 activation rows are not swizzled.)
template <class LD, int R, int NT>
__device__ __forceinline__ void rows2_f16(const uint8_t *r0, const uint8_t *r1, int i0, int i1, const uint8_t *xs, int64_t xrb,
                                          int lane, float (&acc)[R][NT]) {
    const uint8_t *rows[2] = {r0, r1}; // rows[1] unused when R == 1
    for (int i = i0 + lane; i < i1; i += 32) {
        int4 wv[2];
#pragma unroll
        for (int r = 0; r < R; r++) wv[r] = LD::v4(rows[r] + (size_t)i * 16);
#pragma unroll
        for (int t = 0; t < NT; t++) {
            const int4 xv = *(const int4 *)(xs + t * xrb + (size_t)i * 16);
            const __half2 *xh = (const __half2 *)&xv;
#pragma unroll
            for (int r = 0; r < R; r++) {
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

// F32 rows: pieces of 4 floats
template <class LD, int R, int NT>
__device__ __forceinline__ void rows2_f32(const uint8_t *r0, const uint8_t *r1, int i0, int i1, const uint8_t *xs, int64_t xrb,
                                          int lane, float (&acc)[R][NT]) {
    const uint8_t *rows[2] = {r0, r1}; // rows[1] unused when R == 1
    for (int i = i0 + lane; i < i1; i += 32) {
#pragma unroll
        for (int t = 0; t < NT; t++) {
            const float4 xv = *(const float4 *)(xs + t * xrb + (size_t)i * 16);
#pragma unroll
            for (int r = 0; r < R; r++) {
                const int4 wv = LD::v4(rows[r] + (size_t)i * 16);
                const float *wf = (const float *)&wv;
                acc[r][t] += wf[0] * xv.x;
                acc[r][t] += wf[1] * xv.y;
                acc[r][t] += wf[2] * xv.z;
                acc[r][t] += wf[3] * xv.w;
            }
        }
    }
}

// Dispatch on the ggml type.  `seg`/`n_seg` split the row's blocks into n_seg contiguous parts (the mega-kernel
// spreads one long row over several warps); pass 0/1 for the whole row.
template <class LD, int R, int NT>
__device__ __forceinline__ void rows2_dot(int type, const uint8_t *r0, const uint8_t *r1, int64_t K, int seg, int n_seg,
                                          const uint8_t *xs, int64_t xrb, int lane, float (&acc)[R][NT]) {
    const int nb = (int)(K / 256);
    switch (type) {
        case B2_T_Q4_K: rows2_q4_K<LD, R, NT>(r0, r1, nb * seg / n_seg, nb * (seg + 1) / n_seg, K, xs, xrb, lane, acc); break;
        case B2_T_Q5_K: rows2_q5_K<LD, R, NT>(r0, r1, nb * seg / n_seg, nb * (seg + 1) / n_seg, K, xs, xrb, lane, acc); break;
        case B2_T_Q6_K: rows2_q6_K<LD, R, NT>(r0, r1, nb, nb * seg / n_seg, nb * (seg + 1) / n_seg, K, xs, xrb, lane, acc); break;
        case B2_T_Q8_0: { // split on whole 32-element blocks (2 pieces)
            const int n32 = (int)(K / 32);
            rows2_q8_0<LD, R, NT>(r0, r1, 2 * (n32 * seg / n_seg), 2 * (n32 * (seg + 1) / n_seg), K, xs, xrb, lane, acc);
        } break;
        case B2_T_F16: {
            const int n8 = (int)(K / 8);
            rows2_f16<LD, R, NT>(r0, r1, n8 * seg / n_seg, n8 * (seg + 1) / n_seg, xs, xrb, lane, acc);
        } break;
        default: {
            const int n4 = (int)(K / 4);
            rows2_f32<LD, R, NT>(r0, r1, n4 * seg / n_seg, n4 * (seg + 1) / n_seg, xs, xrb, lane, acc);
        } break;
    }
}

} // namespace b2
