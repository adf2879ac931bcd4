// quant.cuh -- device-side views of the GGUF block formats (device layout, see common.h) and helpers.
#pragma once
#include <cuda_fp16.h>
#include <cstdint>

namespace b2 {

#define QK_K 256
#define QK8_0 32

__device__ __forceinline__ float h2f(uint16_t h) { return __half2float(__ushort_as_half(h)); }

// 6-bit scale / min of sub-block j (0..7) from the 12 packed bytes of a Q4_K / Q5_K block
// (ggml-quants.c get_scale_min_k4)
__device__ __forceinline__ void get_scale_min_k4(int j, const uint8_t *q, int &sc, int &m) {
    if (j < 4) {
        sc = q[j] & 63;
        m = q[j + 4] & 63;
    } else {
        sc = (q[j + 4] & 0xF) | ((q[j - 4] >> 6) << 4);
        m = (q[j + 4] >> 4) | ((q[j] >> 6) << 4);
    }
}

// same, from the three 32-bit words of the packed scales held in registers
__device__ __forceinline__ void get_scale_min_k4_w(int j, uint32_t w0, uint32_t w1, uint32_t w2, int &sc, int &m) {
    // bytes 0..3 = w0, 4..7 = w1, 8..11 = w2
    const int jj = j & 3;
    const uint32_t b0 = (w0 >> (8 * jj)) & 0xFF, b1 = (w1 >> (8 * jj)) & 0xFF, b2 = (w2 >> (8 * jj)) & 0xFF;
    if (j < 4) {
        sc = b0 & 63;
        m = b1 & 63;
    } else {
        sc = (b2 & 0xF) | ((b0 >> 6) << 4);
        m = (b2 >> 4) | ((b1 >> 6) << 4);
    }
}

// the 16 levels of IQ4_NL / IQ4_XS (ggml-quants.c kvalues_iq4nl), packed one per byte
__device__ __forceinline__ int iq4nl_value(int idx) {
    const unsigned long long t = idx < 8 ? 0xF6EADDCFBFAD9881ull : 0x7159453526190D01ull;
    return (int)(int8_t)(t >> (8 * (idx & 7)));
}
// 32-element blocks: Q4_0 Q4_1 Q5_0 Q5_1 IQ4_NL
__device__ __forceinline__ bool is_legacy_type(int t) { return t == 2 || t == 3 || t == 6 || t == 7 || t == 20; }
__device__ __forceinline__ int legacy_block_bytes(int t) { return (t == 2 || t == 20) ? 18 : t == 3 ? 20 : t == 6 ? 22 : 24; }
// element l (0..31) of a legacy block (GGUF layout, 2-byte aligned), with dequantize_row_q4_0 / q4_1 / q5_0 / q5_1's
// arithmetic and rounding (unfused multiply / add)
__device__ __forceinline__ float dequant_legacy_elem(int type, const uint8_t *blk, int l) {
    const float d = h2f(*(const uint16_t *)blk);
    const int j = l & 15, hi = l >> 4;
    if (type == 2 /* Q4_0 */) {
        const uint8_t b = blk[2 + j];
        return __fmul_rn((float)((hi ? (b >> 4) : (b & 0x0F)) - 8), d);
    }
    if (type == 20 /* IQ4_NL: d * level[nibble] */) {
        const uint8_t b = blk[2 + j];
        return __fmul_rn(d, (float)iq4nl_value(hi ? (b >> 4) : (b & 0x0F)));
    }
    if (type == 3 /* Q4_1 */) {
        const float m = h2f(*(const uint16_t *)(blk + 2));
        const uint8_t b = blk[4 + j];
        return __fadd_rn(__fmul_rn((float)(hi ? (b >> 4) : (b & 0x0F)), d), m);
    }
    const int off = type == 6 /* Q5_0 */ ? 2 : 4; // qh[4] then qs[16]
    const uint32_t qh = (uint32_t)blk[off] | ((uint32_t)blk[off + 1] << 8) | ((uint32_t)blk[off + 2] << 16) | ((uint32_t)blk[off + 3] << 24);
    const uint8_t b = blk[off + 4 + j];
    const int q = (hi ? (b >> 4) : (b & 0x0F)) | (((qh >> l) & 1) << 4);
    if (type == 6) return __fmul_rn((float)(q - 16), d);
    return __fadd_rn(__fmul_rn((float)q, d), h2f(*(const uint16_t *)(blk + 2)));
}

// element e (0..255) of a Q2_K (84 B: scales[16] qs[64] d dmin) or Q3_K (110 B: hmask[32] qs[64] scales[12] d) block in
// GGUF layout, with dequantize_row_q2_K / q3_K's arithmetic and rounding
// (also IQ4_XS, 136 B: d, scales_h, scales_l[4], qs[128] -- dequantize_row_iq4_xs)
__device__ __forceinline__ bool is_small_k_type(int t) { return t == 10 || t == 11 || t == 23; }
__device__ __forceinline__ int small_k_block_bytes(int t) { return t == 10 ? 84 : t == 11 ? 110 : 136; }
__device__ __forceinline__ float dequant_small_k_elem(int type, const uint8_t *blk, int e) {
    if (type == 23) {
        const int ib = e >> 5, l = e & 31;
        const uint32_t scales_h = *(const uint16_t *)(blk + 2);
        const int ls = ((blk[4 + ib / 2] >> (4 * (ib & 1))) & 0xF) | (((scales_h >> (2 * ib)) & 3) << 4);
        const float dl = __fmul_rn(h2f(*(const uint16_t *)blk), (float)(ls - 32));
        const uint8_t b = blk[8 + 16 * ib + (l & 15)];
        return __fmul_rn(dl, (float)iq4nl_value(l < 16 ? (b & 0x0F) : (b >> 4)));
    }
    const int n = e >> 7, j = (e & 127) >> 5, half = (e & 31) >> 4, l = (e & 15) + 16 * half;
    const int is = n * 8 + 2 * j + half;
    if (type == 10) {
        const uint8_t sc = blk[is];
        const float d = h2f(*(const uint16_t *)(blk + 80)), mn = h2f(*(const uint16_t *)(blk + 82));
        const float dl = __fmul_rn(d, (float)(sc & 0xF)), ml = __fmul_rn(mn, (float)(sc >> 4));
        const int q = (blk[16 + 32 * n + l] >> (2 * j)) & 3;
        return __fsub_rn(__fmul_rn(dl, (float)q), ml);
    }
    // Q3_K: 6-bit scale `is` = 4 low bits from scales[0..7] nibbles, 2 high bits from scales[8..11]
    const uint8_t *sp = blk + 96;
    const int lo = is < 8 ? (sp[is] & 0xF) : (sp[is - 8] >> 4);
    const int hi = (sp[8 + (is & 3)] >> (2 * (is >> 2))) & 3;
    const int scale = (int)(int8_t)(lo | (hi << 4)) - 32;
    const float dl = __fmul_rn(h2f(*(const uint16_t *)(blk + 108)), (float)scale);
    const int q = ((blk[32 + 32 * n + l] >> (2 * j)) & 3) - ((blk[l] & (1 << (4 * n + j))) ? 0 : 4);
    return __fmul_rn(dl, (float)q);
}

// one Q6_K row in device layout: nb 208-byte blocks [ql 128][qh 64][scales 16] (16-byte aligned; 13 x 16 B, an odd
// stride, so lanes that each read a different block of a staged row hit different shared-memory banks), then the nb
// fp16 super-block scales
constexpr int Q6_BLK = 208;
struct Q6Row {
    const uint8_t *base;
    const uint16_t *d;
    __device__ __forceinline__ Q6Row(const uint8_t *row, int nb) : base(row), d((const uint16_t *)(row + (size_t)nb * Q6_BLK)) {}
    __device__ __forceinline__ const uint8_t *ql(int b) const { return base + (size_t)b * Q6_BLK; }
    __device__ __forceinline__ const uint8_t *qh(int b) const { return base + (size_t)b * Q6_BLK + 128; }
    __device__ __forceinline__ const int8_t *sc(int b) const { return (const int8_t *)(base + (size_t)b * Q6_BLK + 192); }
};
// Q8_0 row in device (SoA) layout
struct Q80Row {
    const int8_t *qs;
    const uint16_t *d;
    __device__ __forceinline__ Q80Row(const uint8_t *row, int64_t cols)
        : qs((const int8_t *)row), d((const uint16_t *)(row + cols)) {}
};

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ double warp_sum_d(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

// 128-bit streaming load of weights that are read exactly once per step (no L1 allocation)
__device__ __forceinline__ int4 ld_stream(const void *p) {
    int4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.s32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "l"(p));
    return r;
}

} // namespace b2
