// act_quant.cuh -- warp-level activation quantisation of one 256-element block (lane l owns elements 8l..8l+7),
// bit-exact with ggml's quantize_row_q8_K_ref / quantize_row_q8_0_ref.  `row` is the destination activation row in
// the SoA layout of common.h (global or shared memory), `b` the 256-element block index inside it.
//
// SWZ: the int8 section of a row that lives in SHARED memory for the mat-vec kernels is stored with its 16-byte
// chunks permuted (act_swz) so that the kernels' stride-64-byte reads are bank-conflict free.  Global-memory rows
// (and the values themselves) are never swizzled.
#pragma once
#include "quant.cuh"

namespace b2 {

// chunk c (16 bytes) of the int8 section -> physical chunk; G = 64-byte group, k = chunk inside the group
__host__ __device__ __forceinline__ int act_swz(int c) { return (c & ~3) | ((c & 3) ^ ((c >> 3) & 3)); }

template <bool SWZ>
__device__ __forceinline__ int8_t *act_q8_ptr(uint8_t *row, int64_t off) { // off: byte offset, multiple of 8
    if (!SWZ) return (int8_t *)row + off;
    const int c = (int)(off >> 4);
    return (int8_t *)row + ((int64_t)act_swz(c) << 4) + (off & 15);
}

__device__ __forceinline__ void store_q8x8(int8_t *dst, const int (&q)[8]) {
    int2 packed;
    packed.x = (q[0] & 0xFF) | ((q[1] & 0xFF) << 8) | ((q[2] & 0xFF) << 16) | ((q[3] & 0xFF) << 24);
    packed.y = (q[4] & 0xFF) | ((q[5] & 0xFF) << 8) | ((q[6] & 0xFF) << 16) | ((q[7] & 0xFF) << 24);
    *(int2 *)dst = packed;
}

// Q8_K: [qs: K x int8][d: K/256 x f32][bsums: K/16 x int16]
template <bool SWZ>
__device__ __forceinline__ void quant_block_q8_K(const float (&v)[8], int lane, uint8_t *row, int64_t K, int64_t b) {
    const int64_t nb = K / 256;
    // first element (in index order) with the largest |x|
    float amax = 0.f, mx = 0.f;
    int idx = 0x7fffffff;
#pragma unroll
    for (int i = 0; i < 8; i++) {
        const float ax = fabsf(v[i]);
        if (ax > amax) { amax = ax; mx = v[i]; idx = lane * 8 + i; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const float oa = __shfl_xor_sync(0xffffffffu, amax, o);
        const float om = __shfl_xor_sync(0xffffffffu, mx, o);
        const int oi = __shfl_xor_sync(0xffffffffu, idx, o);
        if (oa > amax || (oa == amax && oi < idx)) { amax = oa; mx = om; idx = oi; }
    }
    int8_t *qs = act_q8_ptr<SWZ>(row, b * 256 + lane * 8);
    float *dptr = (float *)(row + K) + b;
    int16_t *bs = (int16_t *)(row + K + nb * 4) + b * 16;
    if (amax == 0.f) {
        *(int2 *)qs = make_int2(0, 0);
        if (lane == 0) *dptr = 0.f;
        if ((lane & 1) == 0) bs[lane >> 1] = 0;
        return;
    }
    const float iscale = __fdiv_rn(-127.f, mx);
    int sum = 0;
    int q[8];
#pragma unroll
    for (int i = 0; i < 8; i++) {
        int vi = __float2int_rn(__fmul_rn(iscale, v[i])); // nearest_int(): round-half-even
        vi = min(127, vi);
        q[i] = vi;
        sum += vi;
    }
    store_q8x8(qs, q);
    sum += __shfl_xor_sync(0xffffffffu, sum, 1);
    if ((lane & 1) == 0) bs[lane >> 1] = (int16_t)sum;
    if (lane == 0) *dptr = __fdiv_rn(1.f, iscale);
}

// Q8_0: [qs: K x int8][d: K/32 x f32 (value of the fp16-rounded scale)]; a 32-element block = 4 lanes
template <bool SWZ>
__device__ __forceinline__ void quant_block_q8_0(const float (&v)[8], int lane, uint8_t *row, int64_t K, int64_t b) {
    float amax = 0.f;
#pragma unroll
    for (int i = 0; i < 8; i++) amax = fmaxf(amax, fabsf(v[i]));
    amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 1));
    amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 2));
    const float d = __fdiv_rn(amax, 127.f);
    const float id = d != 0.f ? __fdiv_rn(1.f, d) : 0.f;
    int q[8];
#pragma unroll
    for (int i = 0; i < 8; i++) q[i] = (int)roundf(__fmul_rn(v[i], id));
    store_q8x8(act_q8_ptr<SWZ>(row, b * 256 + lane * 8), q);
    if ((lane & 3) == 0) ((float *)(row + K))[b * 8 + (lane >> 2)] = __half2float(__float2half_rn(d));
}

__device__ __forceinline__ void quant_block_f16(const float (&v)[8], int lane, uint8_t *row, int64_t b) {
    __half2 h[4];
#pragma unroll
    for (int i = 0; i < 4; i++) h[i] = __halves2half2(__float2half_rn(v[2 * i]), __float2half_rn(v[2 * i + 1]));
    *(int4 *)((__half *)row + b * 256 + lane * 8) = *(const int4 *)h;
}

__device__ __forceinline__ void quant_block_f32(const float (&v)[8], int lane, uint8_t *row, int64_t b) {
    float *dst = (float *)row + b * 256 + lane * 8;
    *(float4 *)dst = make_float4(v[0], v[1], v[2], v[3]);
    *(float4 *)(dst + 4) = make_float4(v[4], v[5], v[6], v[7]);
}

// fmt: ActFmt value (common.h)
template <bool SWZ>
__device__ __forceinline__ void quant_block(int fmt, const float (&v)[8], int lane, uint8_t *row, int64_t K, int64_t b) {
    if (fmt == 3) quant_block_q8_K<SWZ>(v, lane, row, K, b);
    else if (fmt == 2) quant_block_q8_0<SWZ>(v, lane, row, K, b);
    else if (fmt == 1) quant_block_f16(v, lane, row, b);
    else quant_block_f32(v, lane, row, b);
}

} // namespace b2
