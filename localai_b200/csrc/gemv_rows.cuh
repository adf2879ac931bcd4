// gemv_rows.cuh -- per-type "one warp x one weight row x NT activation rows" dot products shared by the GEMV kernels.
// Arithmetic = ggml_vec_dot_{q4_K,q5_K,q6_K}_q8_K / q8_0_q8_0 / f16: integer dots per block (dp4a), fp32 scaling;
// only the fp32 summation ORDER differs from the scalar reference.  Activations come from shared memory (SoA rows,
// see common.h act_row_bytes); weights are streamed with 128-bit no-allocate loads.
#pragma once
#include "quant.cuh"

namespace b2 {

// ---- per-type row dot products. Each returns this lane's partial sums acc[t]; caller warp-reduces. ----

template <int NT>
__device__ __forceinline__ void row_dot_q4_K(const uint8_t *__restrict__ row, int nb, int64_t K, const uint8_t *__restrict__ xs,
                                             int64_t xrb, int lane, float (&acc)[NT]) {
    const int p = lane & 7;      // 16-byte piece of the block's 128 B of nibbles
    const int j = p >> 1;        // 64-element group
    const int off = (p & 1) * 16; // element offset inside the 32-element sub-block
    for (int ib = lane >> 3; ib < nb; ib += 4) {
        const uint8_t *blk = row + (size_t)ib * 144;
        const int4 hdr = ld_stream(blk);
        const int4 qv = ld_stream(blk + 16 + p * 16);
        const float d = h2f((uint16_t)(hdr.x & 0xFFFF)), dmin = h2f((uint16_t)((uint32_t)hdr.x >> 16));
        int sc_lo, m_lo, sc_hi, m_hi, sc_p, m_p;
        get_scale_min_k4_w(2 * j, hdr.y, hdr.z, hdr.w, sc_lo, m_lo);
        get_scale_min_k4_w(2 * j + 1, hdr.y, hdr.z, hdr.w, sc_hi, m_hi);
        get_scale_min_k4_w(p, hdr.y, hdr.z, hdr.w, sc_p, m_p); // lane p also owns the `min` term of sub-block p
        const uint32_t q[4] = {(uint32_t)qv.x, (uint32_t)qv.y, (uint32_t)qv.z, (uint32_t)qv.w};
#pragma unroll
        for (int t = 0; t < NT; t++) {
            const uint8_t *xt = xs + t * xrb;
            const int8_t *xq = (const int8_t *)xt + (size_t)ib * 256 + j * 64 + off;
            const int4 xl = *(const int4 *)xq, xh = *(const int4 *)(xq + 32);
            const int xlo[4] = {xl.x, xl.y, xl.z, xl.w}, xhi[4] = {xh.x, xh.y, xh.z, xh.w};
            int s_lo = 0, s_hi = 0;
#pragma unroll
            for (int i = 0; i < 4; i++) {
                s_lo = __dp4a((int)(q[i] & 0x0F0F0F0Fu), xlo[i], s_lo);
                s_hi = __dp4a((int)((q[i] >> 4) & 0x0F0F0F0Fu), xhi[i], s_hi);
            }
            const float dx = ((const float *)(xt + K))[ib];
            const int16_t *bs = (const int16_t *)(xt + K + (K >> 6)) + ib * 16;
            const int isum = sc_lo * s_lo + sc_hi * s_hi;
            const int imin = m_p * ((int)bs[2 * p] + (int)bs[2 * p + 1]);
            acc[t] += (d * dx) * (float)isum - (dmin * dx) * (float)imin;
        }
    }
}

template <int NT>
__device__ __forceinline__ void row_dot_q5_K(const uint8_t *__restrict__ row, int nb, int64_t K, const uint8_t *__restrict__ xs,
                                             int64_t xrb, int lane, float (&acc)[NT]) {
    const int p = lane & 7, j = p >> 1, off = (p & 1) * 16;
    for (int ib = lane >> 3; ib < nb; ib += 4) {
        const uint8_t *blk = row + (size_t)ib * 176;
        const int4 hdr = ld_stream(blk);
        const int4 hv = ld_stream(blk + 16 + off);      // qh bytes for elements off..off+15 of every sub-block
        const int4 qv = ld_stream(blk + 48 + p * 16);
        const float d = h2f((uint16_t)(hdr.x & 0xFFFF)), dmin = h2f((uint16_t)((uint32_t)hdr.x >> 16));
        int sc_lo, m_lo, sc_hi, m_hi, sc_p, m_p;
        get_scale_min_k4_w(2 * j, hdr.y, hdr.z, hdr.w, sc_lo, m_lo);
        get_scale_min_k4_w(2 * j + 1, hdr.y, hdr.z, hdr.w, sc_hi, m_hi);
        get_scale_min_k4_w(p, hdr.y, hdr.z, hdr.w, sc_p, m_p);
        const uint32_t q[4] = {(uint32_t)qv.x, (uint32_t)qv.y, (uint32_t)qv.z, (uint32_t)qv.w};
        const uint32_t h[4] = {(uint32_t)hv.x, (uint32_t)hv.y, (uint32_t)hv.z, (uint32_t)hv.w};
        uint32_t wl[4], wh[4];
#pragma unroll
        for (int i = 0; i < 4; i++) {
            wl[i] = (q[i] & 0x0F0F0F0Fu) | (((h[i] >> (2 * j)) & 0x01010101u) << 4);
            wh[i] = ((q[i] >> 4) & 0x0F0F0F0Fu) | (((h[i] >> (2 * j + 1)) & 0x01010101u) << 4);
        }
#pragma unroll
        for (int t = 0; t < NT; t++) {
            const uint8_t *xt = xs + t * xrb;
            const int8_t *xq = (const int8_t *)xt + (size_t)ib * 256 + j * 64 + off;
            const int4 xl = *(const int4 *)xq, xh = *(const int4 *)(xq + 32);
            const int xlo[4] = {xl.x, xl.y, xl.z, xl.w}, xhi[4] = {xh.x, xh.y, xh.z, xh.w};
            int s_lo = 0, s_hi = 0;
#pragma unroll
            for (int i = 0; i < 4; i++) {
                s_lo = __dp4a((int)wl[i], xlo[i], s_lo);
                s_hi = __dp4a((int)wh[i], xhi[i], s_hi);
            }
            const float dx = ((const float *)(xt + K))[ib];
            const int16_t *bs = (const int16_t *)(xt + K + (K >> 6)) + ib * 16;
            const int isum = sc_lo * s_lo + sc_hi * s_hi;
            const int imin = m_p * ((int)bs[2 * p] + (int)bs[2 * p + 1]);
            acc[t] += (d * dx) * (float)isum - (dmin * dx) * (float)imin;
        }
    }
}

template <int NT>
__device__ __forceinline__ void row_dot_q6_K(const uint8_t *__restrict__ row, int nb, int64_t K, const uint8_t *__restrict__ xs,
                                             int64_t xrb, int lane, float (&acc)[NT]) {
    const Q6Row w(row, nb);
    const int p = lane & 7, n = p >> 2, c = p & 3; // 16-byte piece p of ql: half n, 16-element column c
    const int sh = 2 * (c >> 1);                    // qh bit position for the low-nibble elements
    for (int ib = lane >> 3; ib < nb; ib += 4) {
        const int4 lv = ld_stream(w.ql(ib) + p * 16);
        const int4 hv = ld_stream(w.qh(ib) + n * 32 + (c & 1) * 16);
        const int sc_lo = w.sc(ib)[8 * n + c], sc_hi = w.sc(ib)[8 * n + 4 + c];
        const float d = h2f(w.d[ib]);
        const uint32_t l[4] = {(uint32_t)lv.x, (uint32_t)lv.y, (uint32_t)lv.z, (uint32_t)lv.w};
        const uint32_t h[4] = {(uint32_t)hv.x, (uint32_t)hv.y, (uint32_t)hv.z, (uint32_t)hv.w};
        uint32_t wl[4], wh[4];
#pragma unroll
        for (int i = 0; i < 4; i++) {
            wl[i] = (l[i] & 0x0F0F0F0Fu) | (((h[i] >> sh) & 0x03030303u) << 4);
            wh[i] = ((l[i] >> 4) & 0x0F0F0F0Fu) | (((h[i] >> (sh + 4)) & 0x03030303u) << 4);
        }
#pragma unroll
        for (int t = 0; t < NT; t++) {
            const uint8_t *xt = xs + t * xrb;
            const int8_t *xq = (const int8_t *)xt + (size_t)ib * 256 + 128 * n + 16 * c;
            const int4 xl = *(const int4 *)xq, xh = *(const int4 *)(xq + 64);
            const int xlo[4] = {xl.x, xl.y, xl.z, xl.w}, xhi[4] = {xh.x, xh.y, xh.z, xh.w};
            int s_lo = 0, s_hi = 0;
#pragma unroll
            for (int i = 0; i < 4; i++) {
                s_lo = __dp4a((int)wl[i], xlo[i], s_lo);
                s_hi = __dp4a((int)wh[i], xhi[i], s_hi);
            }
            const float dx = ((const float *)(xt + K))[ib];
            const int16_t *bs = (const int16_t *)(xt + K + (K >> 6)) + ib * 16;
            // stored quants are q+32: subtract 32 * sum(x) of the 16-element sub-block (exact, integer)
            s_lo -= 32 * (int)bs[8 * n + c];
            s_hi -= 32 * (int)bs[8 * n + 4 + c];
            acc[t] += (d * dx) * (float)(sc_lo * s_lo + sc_hi * s_hi);
        }
    }
}

template <int NT>
__device__ __forceinline__ void row_dot_q8_0(const uint8_t *__restrict__ row, int64_t K, const uint8_t *__restrict__ xs,
                                             int64_t xrb, int lane, float (&acc)[NT]) {
    const Q80Row w(row, K);
    const int n16 = (int)(K / 16); // 16-byte pieces; a block is 2 pieces -> lanes (2i, 2i+1) share a block
    for (int i0 = 0; i0 < n16; i0 += 32) { // warp-uniform trip count: the shuffle below needs all lanes
        const int i = i0 + lane;
        const bool valid = i < n16;
        const int4 wv = valid ? ld_stream(w.qs + (size_t)i * 16) : make_int4(0, 0, 0, 0);
        const float dw = valid ? h2f(w.d[i >> 1]) : 0.f;
#pragma unroll
        for (int t = 0; t < NT; t++) {
            const uint8_t *xt = xs + t * xrb;
            const int4 xv = valid ? *(const int4 *)(xt + (size_t)i * 16) : make_int4(0, 0, 0, 0);
            int s = __dp4a(wv.x, xv.x, 0);
            s = __dp4a(wv.y, xv.y, s);
            s = __dp4a(wv.z, xv.z, s);
            s = __dp4a(wv.w, xv.w, s);
            s += __shfl_xor_sync(0xffffffffu, s, 1); // whole 32-element block, as on the CPU
            const float dx = valid ? ((const float *)(xt + K))[i >> 1] : 0.f;
            if ((lane & 1) == 0) acc[t] += (float)s * (dw * dx);
        }
    }
}

template <int NT>
__device__ __forceinline__ void row_dot_f16(const uint8_t *__restrict__ row, int64_t K, const uint8_t *__restrict__ xs,
                                            int64_t xrb, int lane, float (&acc)[NT]) {
    const int n8 = (int)(K / 8);
    for (int i = lane; i < n8; i += 32) {
        const int4 wv = ld_stream(row + (size_t)i * 16);
        const __half2 *wh = (const __half2 *)&wv;
#pragma unroll
        for (int t = 0; t < NT; t++) {
            const int4 xv = *(const int4 *)(xs + t * xrb + (size_t)i * 16);
            const __half2 *xh = (const __half2 *)&xv;
#pragma unroll
            for (int k = 0; k < 4; k++) {
                const float2 a = __half22float2(wh[k]), b = __half22float2(xh[k]);
                acc[t] += a.x * b.x;
                acc[t] += a.y * b.y;
            }
        }
    }
}

template <int NT>
__device__ __forceinline__ void row_dot_f32(const uint8_t *__restrict__ row, int64_t K, const uint8_t *__restrict__ xs,
                                            int64_t xrb, int lane, float (&acc)[NT]) {
    const int n4 = (int)(K / 4);
    for (int i = lane; i < n4; i += 32) {
        const int4 wv = ld_stream(row + (size_t)i * 16);
        const float *wf = (const float *)&wv;
#pragma unroll
        for (int t = 0; t < NT; t++) {
            const float4 xv = *(const float4 *)(xs + t * xrb + (size_t)i * 16);
            acc[t] += wf[0] * xv.x;
            acc[t] += wf[1] * xv.y;
            acc[t] += wf[2] * xv.z;
            acc[t] += wf[3] * xv.w;
        }
    }
}

} // namespace b2
