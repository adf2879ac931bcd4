// kernels_quant.cu -- load-time repack, on-device dequantisation, activation quantisation.
//
// Numerics contract (tests/test_parity_gpu.py):
//   * dequantisation is BIT-EXACT with ggml's dequantize_row_* / gguf-py: every multiply and subtract is
//     an explicit round-to-nearest intrinsic so nvcc cannot fuse `d*q - m` into an FMA.
//   * activation quantisation is BIT-EXACT with quantize_row_q8_0_ref / quantize_row_q8_K_ref.
#include "kernels.h"
#include "quant.cuh"

namespace b2 {

thread_local int64_t g_launch_count = 0;

// ------------------------------------------------------------------------------------------ repack
// One thread per 2-byte unit of the source tensor (GGUF blocks are 2-byte aligned at worst).
__global__ void repack_q8_0_kernel(const uint16_t *__restrict__ src, uint8_t *__restrict__ dst, int64_t rows, int64_t cols,
                                   int64_t dst_stride) {
    const int64_t nb = cols / 32;
    const int64_t total = rows * nb * 17;
    for (int64_t u = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; u < total; u += (int64_t)gridDim.x * blockDim.x) {
        const int64_t blk = u / 17;
        const int w = (int)(u % 17);
        const int64_t r = blk / nb, b = blk % nb;
        uint8_t *row = dst + r * dst_stride;
        const uint16_t v = src[u];
        if (w == 0) ((uint16_t *)(row + cols))[b] = v;
        else ((uint16_t *)row)[b * 16 + (w - 1)] = v;
    }
}

__global__ void repack_q6_k_kernel(const uint16_t *__restrict__ src, uint8_t *__restrict__ dst, int64_t rows, int64_t cols,
                                   int64_t dst_stride) {
    const int64_t nb = cols / 256;
    const int64_t total = rows * nb * 105;
    for (int64_t u = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; u < total; u += (int64_t)gridDim.x * blockDim.x) {
        const int64_t blk = u / 105;
        const int w = (int)(u % 105);
        const int64_t r = blk / nb, b = blk % nb;
        uint8_t *row = dst + r * dst_stride;
        const uint16_t v = src[u];
        if (w < 104) ((uint16_t *)(row + b * Q6_BLK))[w] = v;   // ql | qh | scales keep their order inside the block
        else ((uint16_t *)(row + nb * Q6_BLK))[b] = v;           // d
    }
}

__global__ void repack_copy_kernel(const uint8_t *__restrict__ src, uint8_t *__restrict__ dst, int64_t rows,
                                   int64_t src_row_bytes, int64_t dst_stride) {
    // src rows are at least 2-byte aligned; copy 2 bytes per thread
    const int64_t per_row = src_row_bytes / 2;
    const int64_t total = rows * per_row;
    for (int64_t u = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; u < total; u += (int64_t)gridDim.x * blockDim.x) {
        const int64_t r = u / per_row, w = u % per_row;
        ((uint16_t *)(dst + r * dst_stride))[w] = ((const uint16_t *)src)[u];
    }
}

void launch_repack(int type, const uint8_t *src, uint8_t *dst, int64_t rows, int64_t cols, cudaStream_t s) {
    const int64_t stride = device_row_stride(type, cols);
    const int threads = 256;
    const int blocks = 148 * 8;
    if (type == B2_T_Q8_0) {
        repack_q8_0_kernel<<<blocks, threads, 0, s>>>((const uint16_t *)src, dst, rows, cols, stride);
    } else if (type == B2_T_Q6_K) {
        repack_q6_k_kernel<<<blocks, threads, 0, s>>>((const uint16_t *)src, dst, rows, cols, stride);
    } else {
        const int64_t rb = gguf_row_bytes(type, cols);
        if (rb == stride) {
            B2_CUDA(cudaMemcpyAsync(dst, src, rows * rb, cudaMemcpyDeviceToDevice, s));
            return;
        }
        repack_copy_kernel<<<blocks, threads, 0, s>>>(src, dst, rows, rb, stride);
    }
    g_launch_count++;
    B2_CUDA(cudaGetLastError());
}

// ------------------------------------------------------------------------------------------ dequant
// One thread per 32 output elements.
__global__ void dequant_rows_kernel(const uint8_t *__restrict__ W, int type, int64_t cols, int64_t row_stride,
                                    const int32_t *__restrict__ row_ids, int64_t row0, int n_rows, float *__restrict__ out,
                                    int64_t out_stride) {
    const int64_t groups_per_row = cols / 32;
    const int64_t total = (int64_t)n_rows * groups_per_row;
    for (int64_t g = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; g < total; g += (int64_t)gridDim.x * blockDim.x) {
        const int64_t i = g / groups_per_row;
        const int64_t grp = g % groups_per_row;
        const int64_t r = row_ids ? (int64_t)row_ids[i] : row0 + i;
        const uint8_t *row = W + r * row_stride;
        float *y = out + i * out_stride + grp * 32;
        if (type == B2_T_F32) {
            const float *x = (const float *)row + grp * 32;
            for (int l = 0; l < 32; l++) y[l] = x[l];
        } else if (type == B2_T_F16) {
            const uint16_t *x = (const uint16_t *)row + grp * 32;
            for (int l = 0; l < 32; l++) y[l] = h2f(x[l]);
        } else if (type == B2_T_Q8_0) {
            Q80Row q(row, cols);
            const float d = h2f(q.d[grp]);
            const int8_t *qs = q.qs + grp * 32;
            for (int l = 0; l < 32; l++) y[l] = __fmul_rn((float)qs[l], d);
        } else if (is_small_k_type(type)) {
            const uint8_t *blk = row + (grp / 8) * small_k_block_bytes(type);
            for (int l = 0; l < 32; l++) y[l] = dequant_small_k_elem(type, blk, (int)(grp % 8) * 32 + l);
        } else if (is_legacy_type(type)) {
            const uint8_t *blk = row + grp * legacy_block_bytes(type);
            for (int l = 0; l < 32; l++) y[l] = dequant_legacy_elem(type, blk, l);
        } else if (type == B2_T_Q4_K || type == B2_T_Q5_K) {
            const int bb = type == B2_T_Q4_K ? 144 : 176;
            const int64_t ib = grp / 8;
            const int j = (int)(grp % 8);
            const uint8_t *blk = row + ib * bb;
            const float d = h2f(*(const uint16_t *)blk), dmin = h2f(*(const uint16_t *)(blk + 2));
            int sc, m;
            get_scale_min_k4(j, blk + 4, sc, m);
            const float d1 = __fmul_rn(d, (float)sc), m1 = __fmul_rn(dmin, (float)m);
            const uint8_t *qh = blk + 16;
            const uint8_t *qs = blk + (type == B2_T_Q4_K ? 16 : 48) + (j / 2) * 32;
            for (int l = 0; l < 32; l++) {
                int q = (j & 1) ? (qs[l] >> 4) : (qs[l] & 0xF);
                if (type == B2_T_Q5_K) q += ((qh[l] >> j) & 1) ? 16 : 0;
                y[l] = __fsub_rn(__fmul_rn(d1, (float)q), m1);
            }
        } else if (type == B2_T_Q6_K) {
            const int nb = (int)(cols / 256);
            Q6Row q(row, nb);
            const int64_t ib = grp / 8;
            const int j = (int)(grp % 8), n = j / 4, c = j % 4;
            const float d = h2f(q.d[ib]);
            const uint8_t *ql = q.ql(ib) + 64 * n + ((c & 1) ? 32 : 0);
            const uint8_t *qh = q.qh(ib) + 32 * n;
            const int8_t *sc = q.sc(ib) + 8 * n + 2 * c;
            for (int l = 0; l < 32; l++) {
                const int lo = (c < 2) ? (ql[l] & 0xF) : (ql[l] >> 4);
                const int hi = (qh[l] >> (2 * c)) & 3;
                const int v = (int)(int8_t)(lo | (hi << 4)) - 32;
                y[l] = __fmul_rn(__fmul_rn(d, (float)sc[l / 16]), (float)v);
            }
        }
    }
}

void launch_dequant_rows(const DTensor &W, const int32_t *row_ids, int64_t row0, int n_rows, float *out, int64_t out_stride,
                         cudaStream_t s) {
    if (n_rows <= 0) return;
    const int64_t total = (int64_t)n_rows * (W.cols / 32);
    const int threads = 128;
    const int blocks = (int)std::min<int64_t>((total + threads - 1) / threads, 148 * 16);
    dequant_rows_kernel<<<blocks, threads, 0, s>>>(W.data, W.type, W.cols, W.row_stride, row_ids, row0, n_rows, out, out_stride);
    g_launch_count++;
    B2_CUDA(cudaGetLastError());
}

// ------------------------------------------------------------------------------------------ activation quantisation
// Q8_K: one warp per 256-element block; lane l owns elements 8l..8l+7.
__global__ void quantize_q8_K_kernel(const float *__restrict__ x, int64_t x_stride, int n_tok, int64_t K,
                                     uint8_t *__restrict__ out, int64_t row_bytes) {
    const int lane = threadIdx.x & 31;
    const int64_t nb = K / 256;
    const int64_t warp = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
    const int64_t n_warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t wb = warp; wb < (int64_t)n_tok * nb; wb += n_warps) {
        const int64_t t = wb / nb, b = wb % nb;
        const float *xb = x + t * x_stride + b * 256 + lane * 8;
        float v[8];
        const float4 a0 = *(const float4 *)xb, a1 = *(const float4 *)(xb + 4);
        v[0] = a0.x; v[1] = a0.y; v[2] = a0.z; v[3] = a0.w; v[4] = a1.x; v[5] = a1.y; v[6] = a1.z; v[7] = a1.w;
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
        uint8_t *row = out + t * row_bytes;
        int8_t *qs = (int8_t *)row + b * 256 + lane * 8;
        float *dptr = (float *)(row + K) + b;
        int16_t *bs = (int16_t *)(row + K + nb * 4) + b * 16;
        if (amax == 0.f) {
            *(int2 *)qs = make_int2(0, 0);
            if (lane == 0) *dptr = 0.f;
            if ((lane & 1) == 0) bs[lane >> 1] = 0;
            continue;
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
        int2 packed;
        packed.x = (q[0] & 0xFF) | ((q[1] & 0xFF) << 8) | ((q[2] & 0xFF) << 16) | ((q[3] & 0xFF) << 24);
        packed.y = (q[4] & 0xFF) | ((q[5] & 0xFF) << 8) | ((q[6] & 0xFF) << 16) | ((q[7] & 0xFF) << 24);
        *(int2 *)qs = packed;
        sum += __shfl_xor_sync(0xffffffffu, sum, 1);
        if ((lane & 1) == 0) bs[lane >> 1] = (int16_t)sum;
        if (lane == 0) *dptr = __fdiv_rn(1.f, iscale);
    }
}

// Q8_0: lane owns 8 consecutive elements; a 32-element block = 4 lanes; a warp covers 8 blocks.
__global__ void quantize_q8_0_kernel(const float *__restrict__ x, int64_t x_stride, int n_tok, int64_t K,
                                     uint8_t *__restrict__ out, int64_t row_bytes) {
    const int lane = threadIdx.x & 31;
    const int64_t nw = K / 256; // warps-worth of work per row (K % 256 == 0 is guaranteed by the model loader)
    const int64_t warp = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
    const int64_t n_warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t wb = warp; wb < (int64_t)n_tok * nw; wb += n_warps) {
        const int64_t t = wb / nw, b = wb % nw;
        const float *xb = x + t * x_stride + b * 256 + lane * 8;
        float v[8];
        const float4 a0 = *(const float4 *)xb, a1 = *(const float4 *)(xb + 4);
        v[0] = a0.x; v[1] = a0.y; v[2] = a0.z; v[3] = a0.w; v[4] = a1.x; v[5] = a1.y; v[6] = a1.z; v[7] = a1.w;
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
        int2 packed;
        packed.x = (q[0] & 0xFF) | ((q[1] & 0xFF) << 8) | ((q[2] & 0xFF) << 16) | ((q[3] & 0xFF) << 24);
        packed.y = (q[4] & 0xFF) | ((q[5] & 0xFF) << 8) | ((q[6] & 0xFF) << 16) | ((q[7] & 0xFF) << 24);
        uint8_t *row = out + t * row_bytes;
        *(int2 *)((int8_t *)row + b * 256 + lane * 8) = packed;
        if ((lane & 3) == 0) ((float *)(row + K))[b * 8 + (lane >> 2)] = __half2float(__float2half_rn(d));
    }
}

__global__ void convert_f16_kernel(const float *__restrict__ x, int64_t x_stride, int n_tok, int64_t K, uint8_t *__restrict__ out,
                                   int64_t row_bytes) {
    const int64_t total = (int64_t)n_tok * K;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t t = i / K, c = i % K;
        ((__half *)(out + t * row_bytes))[c] = __float2half_rn(x[t * x_stride + c]);
    }
}

__global__ void copy_f32_kernel(const float *__restrict__ x, int64_t x_stride, int n_tok, int64_t K, uint8_t *__restrict__ out,
                                int64_t row_bytes) {
    const int64_t total = (int64_t)n_tok * K;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t t = i / K, c = i % K;
        ((float *)(out + t * row_bytes))[c] = x[t * x_stride + c];
    }
}

void launch_quantize_act(ActFmt f, const float *x, int64_t x_stride, int n_tok, int64_t K, uint8_t *out, cudaStream_t s) {
    if (n_tok <= 0) return;
    const int64_t rb = act_row_bytes(f, K);
    const int threads = 128;
    if (f == ACT_Q8_K || f == ACT_Q8_0) {
        const int64_t warps = (int64_t)n_tok * (K / 256);
        const int blocks = (int)std::min<int64_t>((warps * 32 + threads - 1) / threads, 148 * 16);
        if (f == ACT_Q8_K) quantize_q8_K_kernel<<<blocks, threads, 0, s>>>(x, x_stride, n_tok, K, out, rb);
        else quantize_q8_0_kernel<<<blocks, threads, 0, s>>>(x, x_stride, n_tok, K, out, rb);
    } else {
        const int64_t total = (int64_t)n_tok * K;
        const int blocks = (int)std::min<int64_t>((total + 255) / 256, 148 * 16);
        if (f == ACT_F16) convert_f16_kernel<<<blocks, 256, 0, s>>>(x, x_stride, n_tok, K, out, rb);
        else copy_f32_kernel<<<blocks, 256, 0, s>>>(x, x_stride, n_tok, K, out, rb);
    }
    g_launch_count++;
    B2_CUDA(cudaGetLastError());
}

} // namespace b2
