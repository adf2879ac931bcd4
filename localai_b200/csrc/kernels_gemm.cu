// kernels_gemm.cu -- prefill / batched-decode mat-mul on the 5th-generation tensor cores (tcgen05 + TMEM).
//
// out[M tokens][N rows] = X[M][K] . W[N][K]^T, a real dense contraction once M is in the tens or more.
//   1. dequant_tiles_kernel   quantised W (GGUF blocks in HBM) -> fp16, written straight into the UMMA "core matrix"
//                             tile layout (below), so that a whole 256 x 64 operand tile is ONE contiguous 32 KB chunk
//   2. convert_x_tiles_kernel fp32 activations -> fp16 in the same layout (128 x 64 tiles, rows zero-padded)
//   3. gemm_tcgen05_kernel    persistent, warp-specialised: warp 0 = producer (one cp.async.bulk per operand tile,
//                             mbarrier ring of 4 stages), warp 1 = MMA issuer (one elected lane issues
//                             tcgen05.mma.cta_group::1.kind::f16, 128 x 256 x 16, fp32 accumulators in TMEM, two
//                             accumulator stages = 512 columns), warp 2 = TMEM allocator, warps 4..7 = epilogue
//                             (tcgen05.ld 32x32b -> registers -> fp32 rows (+ residual) to global).
// Tile layout (no swizzle, K-major): a tile of R rows x 64 k-elements is stored as 8x8 core matrices (8 rows x 16 B,
// 128 B contiguous): byte offset = (r/8) * 1024 + (k/8) * 128 + (r%8) * 16 + (k%8) * 2.  The shared-memory matrix
// descriptor therefore has LBO (K direction) = 128 B and SBO (row direction) = 1024 B, and one K = 16 MMA step
// advances the start address by 256 B.  Because the tiles are laid out this way in global memory already, no TMA
// tensor map is needed.
// Tensor-bound: 2 M N K flops; roofline = MEASURED_PEAKS.json dense bf16 figure (1392 TF/s sustained; kind::f16 runs fp16 at the same rate).
#include <cuda_fp16.h>

#include "kernels.h"
#include "quant.cuh"

namespace b2 {

constexpr int GM_BM = 128, GM_BN = 256, GM_BK = 64;
constexpr int GM_STAGES = 4;
constexpr int GM_A_BYTES = GM_BM * GM_BK * 2; // 16 KB
constexpr int GM_B_BYTES = GM_BN * GM_BK * 2; // 32 KB
constexpr int GM_STAGE_BYTES = GM_A_BYTES + GM_B_BYTES;
constexpr int GM_THREADS = 384; // warp 0 producer, 1 MMA issuer, 2 TMEM allocation, 4 .. 11 epilogue (two per TMEM lane quarter)
constexpr int GM_EPI_WARPS = 8;
constexpr size_t GM_SMEM = (size_t)GM_STAGES * GM_STAGE_BYTES + 1024 + GM_EPI_WARPS * 4096; // ring, barriers, 4 KB of row staging per epilogue warp

// fp32 pair -> fp16 pair, saturated (an inf would poison a whole accumulator row; llama.cpp's CUDA path makes the same
// fp16 trade for these mat-muls)
__device__ __forceinline__ __half2 gm_pack2(float a, float b) {
    return __floats2half2_rn(fminf(fmaxf(a, -65504.f), 65504.f), fminf(fmaxf(b, -65504.f), 65504.f));
}

__host__ __device__ inline int64_t tile_elem_offset(int r, int k) { // element offset inside a tile (rows x 64)
    return (int64_t)(r >> 3) * 512 + (k >> 3) * 64 + (r & 7) * 8 + (k & 7);
}

// ------------------------------------------------------------------------------------------------ operand preparation
// one thread = 8 consecutive k of one row = one 16-byte core-matrix row
// half >= 0: the matrix is one half of an interleaved pair (ffn_gate | ffn_up): its rows [128 j, 128 j + 128) become rows
// [128 half, 128 half + 128) of the pair's 256-row tile j, so that one accumulator tile holds gate and up of the same 128
// features and the GEMM epilogue can apply SwiGLU itself (rows_pad is then a multiple of 128).
__global__ void dequant_tiles_kernel(const uint8_t *__restrict__ W, int type, int64_t rows, int64_t cols, int64_t row_stride,
                                     __half *__restrict__ out, int64_t rows_pad, int half) {
    const int64_t kc_per_row = cols / 8;
    const int64_t total = rows_pad * kc_per_row;
    const int64_t k_tiles = cols / GM_BK;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t r = i / kc_per_row;
        const int k0 = (int)(i % kc_per_row) * 8;
        float v[8];
        if (r >= rows) {
#pragma unroll
            for (int j = 0; j < 8; j++) v[j] = 0.f;
        } else {
            const uint8_t *row = W + r * row_stride;
            if (type == B2_T_F32) {
#pragma unroll
                for (int j = 0; j < 8; j++) v[j] = ((const float *)row)[k0 + j];
            } else if (type == B2_T_F16) {
#pragma unroll
                for (int j = 0; j < 8; j++) v[j] = h2f(((const uint16_t *)row)[k0 + j]);
            } else if (type == B2_T_Q8_0) {
                Q80Row q(row, cols);
                const float d = h2f(q.d[k0 / 32]);
#pragma unroll
                for (int j = 0; j < 8; j++) v[j] = (float)q.qs[k0 + j] * d;
            } else if (is_small_k_type(type)) {
                const uint8_t *blk = row + (size_t)(k0 / 256) * small_k_block_bytes(type);
#pragma unroll
                for (int j = 0; j < 8; j++) v[j] = dequant_small_k_elem(type, blk, k0 % 256 + j);
            } else if (is_legacy_type(type)) {
                const uint8_t *blk = row + (size_t)(k0 / 32) * legacy_block_bytes(type);
#pragma unroll
                for (int j = 0; j < 8; j++) v[j] = dequant_legacy_elem(type, blk, k0 % 32 + j);
            } else if (type == B2_T_Q4_K || type == B2_T_Q5_K) {
                const int bb = type == B2_T_Q4_K ? 144 : 176;
                const uint8_t *blk = row + (size_t)(k0 / 256) * bb;
                const int e = k0 % 256, sb = e / 32, l0 = e % 32; // sub-block, offset inside it
                const float d = h2f(*(const uint16_t *)blk), dmin = h2f(*(const uint16_t *)(blk + 2));
                int sc, m;
                get_scale_min_k4(sb, blk + 4, sc, m);
                const float d1 = d * (float)sc, m1 = dmin * (float)m;
                const uint8_t *qh = blk + 16;
                const uint8_t *qs = blk + (type == B2_T_Q4_K ? 16 : 48) + (sb / 2) * 32 + l0;
#pragma unroll
                for (int j = 0; j < 8; j++) {
                    int q = (sb & 1) ? (qs[j] >> 4) : (qs[j] & 0xF);
                    if (type == B2_T_Q5_K) q += ((qh[l0 + j] >> sb) & 1) ? 16 : 0;
                    v[j] = d1 * (float)q - m1;
                }
            } else { // Q6_K (device SoA layout)
                const int nb = (int)(cols / 256);
                Q6Row q(row, nb);
                const int64_t ib = k0 / 256;
                const int e = k0 % 256, n = e / 128, c = (e % 128) / 32, l0 = e % 32;
                const float d = h2f(q.d[ib]);
                const uint8_t *ql = q.ql(ib) + 64 * n + ((c & 1) ? 32 : 0) + l0;
                const uint8_t *qh = q.qh(ib) + 32 * n + l0;
                const float ds = d * (float)q.sc(ib)[8 * n + 2 * c + l0 / 16];
#pragma unroll
                for (int j = 0; j < 8; j++) {
                    const int lo = (c < 2) ? (ql[j] & 0xF) : (ql[j] >> 4);
                    const int hi = (qh[j] >> (2 * c)) & 3;
                    v[j] = ds * (float)((lo | (hi << 4)) - 32);
                }
            }
        }
        const int64_t nt = half < 0 ? r / GM_BN : r / 128, kt = k0 / GM_BK;
        const int rt = half < 0 ? (int)(r % GM_BN) : (int)(r % 128) + 128 * half;
        __half *dst = out + (nt * k_tiles + kt) * (int64_t)(GM_BN * GM_BK) + tile_elem_offset(rt, k0 % GM_BK);
        __half2 p[4];
#pragma unroll
        for (int j = 0; j < 4; j++) p[j] = gm_pack2(v[2 * j], v[2 * j + 1]);
        *(int4 *)dst = *(const int4 *)p;
    }
}

__global__ void convert_x_tiles_kernel(const float *__restrict__ x, int64_t x_stride, int n_tok, int64_t cols,
                                       __half *__restrict__ out, int64_t m_pad) {
    const int64_t kc_per_row = cols / 8;
    const int64_t total = m_pad * kc_per_row;
    const int64_t k_tiles = cols / GM_BK;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t r = i / kc_per_row;
        const int k0 = (int)(i % kc_per_row) * 8;
        float4 a = make_float4(0, 0, 0, 0), b = a;
        if (r < n_tok) {
            a = *(const float4 *)(x + r * x_stride + k0);
            b = *(const float4 *)(x + r * x_stride + k0 + 4);
        }
        const int64_t mt = r / GM_BM, kt = k0 / GM_BK;
        __half *dst = out + (mt * k_tiles + kt) * (int64_t)(GM_BM * GM_BK) + tile_elem_offset((int)(r % GM_BM), k0 % GM_BK);
        __half2 p[4] = {gm_pack2(a.x, a.y), gm_pack2(a.z, a.w), gm_pack2(b.x, b.y),
                               gm_pack2(b.z, b.w)};
        *(int4 *)dst = *(const int4 *)p;
    }
}

// RMSNorm (ggml arithmetic: double-precision sum, (x * scale) * w) fused with the fp16 tile conversion: one CTA per
// token row; rows m >= n_tok of the last tile are zeroed by the CTAs of the same tile index range.
__global__ void __launch_bounds__(256)
rmsnorm_x_tiles_kernel(const float *__restrict__ x, int64_t x_stride, const float *__restrict__ w, int n_tok, int64_t cols, float eps,
                       __half *__restrict__ out, float *__restrict__ xn_out) {
    const int r = blockIdx.x; // < m_pad
    const int64_t k_tiles = cols / GM_BK;
    const int64_t mt = r / GM_BM;
    __half *tile0 = out + mt * k_tiles * (int64_t)(GM_BM * GM_BK);
    __shared__ double part[8];
    __shared__ float s_scale;
    float scale = 0.f;
    const float *xr = x + (int64_t)r * x_stride;
    if (r < n_tok) {
        double sum = 0.0;
        for (int64_t i = threadIdx.x; i < cols; i += blockDim.x) {
            const float v = xr[i];
            sum += (double)__fmul_rn(v, v);
        }
        sum = warp_sum_d(sum);
        if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = sum;
        __syncthreads();
        if (threadIdx.x == 0) {
            double tot = 0.0;
            for (int i = 0; i < 8; i++) tot += part[i];
            const float mean = (float)(tot / (double)cols);
            s_scale = __fdiv_rn(1.0f, __fsqrt_rn(__fadd_rn(mean, eps)));
        }
        __syncthreads();
        scale = s_scale;
    }
    for (int64_t k0 = (int64_t)threadIdx.x * 8; k0 < cols; k0 += (int64_t)blockDim.x * 8) {
        float v[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        if (r < n_tok) {
            const float4 a = *(const float4 *)(xr + k0), b = *(const float4 *)(xr + k0 + 4);
            const float4 wa = *(const float4 *)(w + k0), wb = *(const float4 *)(w + k0 + 4);
            v[0] = __fmul_rn(__fmul_rn(a.x, scale), wa.x); v[1] = __fmul_rn(__fmul_rn(a.y, scale), wa.y);
            v[2] = __fmul_rn(__fmul_rn(a.z, scale), wa.z); v[3] = __fmul_rn(__fmul_rn(a.w, scale), wa.w);
            v[4] = __fmul_rn(__fmul_rn(b.x, scale), wb.x); v[5] = __fmul_rn(__fmul_rn(b.y, scale), wb.y);
            v[6] = __fmul_rn(__fmul_rn(b.z, scale), wb.z); v[7] = __fmul_rn(__fmul_rn(b.w, scale), wb.w);
            if (xn_out) {
                *(float4 *)(xn_out + (int64_t)r * cols + k0) = make_float4(v[0], v[1], v[2], v[3]);
                *(float4 *)(xn_out + (int64_t)r * cols + k0 + 4) = make_float4(v[4], v[5], v[6], v[7]);
            }
        }
        __half *dst = tile0 + (k0 / GM_BK) * (int64_t)(GM_BM * GM_BK) + tile_elem_offset(r % GM_BM, (int)(k0 % GM_BK));
        __half2 p4[4] = {gm_pack2(v[0], v[1]), gm_pack2(v[2], v[3]), gm_pack2(v[4], v[5]),
                                gm_pack2(v[6], v[7])};
        *(int4 *)dst = *(const int4 *)p4;
    }
}

// The same with one WARP per token row and the row held in registers (cols = 256 NCH, 8 NCH features per lane): one pass over
// x instead of two, and the four rows of a CTA fill neighbouring 16-byte pieces of the same 128-byte lines of the tile (a row
// per CTA writes 16 bytes into each of 512 different lines).  Used for long batches (prefill).
template <int NCH>
__global__ void __launch_bounds__(128)
rmsnorm_x_tiles_warp_kernel(const float *__restrict__ x, int64_t x_stride, const float *__restrict__ w, int n_tok, float eps,
                            __half *__restrict__ out, float *__restrict__ xn_out, int m_pad) {
    constexpr int64_t cols = 256 * NCH;
    const int lane = threadIdx.x & 31;
    const int r = blockIdx.x * 4 + (threadIdx.x >> 5);
    if (r >= m_pad) return;
    constexpr int64_t k_tiles = cols / GM_BK;
    __half *tile0 = out + (int64_t)(r / GM_BM) * k_tiles * (int64_t)(GM_BM * GM_BK);
    float v[NCH][8];
    if (r < n_tok) { // warp-uniform
        const float *xr = x + (int64_t)r * x_stride;
#pragma unroll
        for (int i = 0; i < NCH; i++) {
            const float4 a = *(const float4 *)(xr + 256 * i + 8 * lane), b = *(const float4 *)(xr + 256 * i + 8 * lane + 4);
            v[i][0] = a.x; v[i][1] = a.y; v[i][2] = a.z; v[i][3] = a.w; v[i][4] = b.x; v[i][5] = b.y; v[i][6] = b.z; v[i][7] = b.w;
        }
        double sum = 0.0;
#pragma unroll
        for (int i = 0; i < NCH; i++)
#pragma unroll
            for (int j = 0; j < 8; j++) sum += (double)__fmul_rn(v[i][j], v[i][j]);
        sum = warp_sum_d(sum);
        const float mean = (float)(sum / (double)cols);
        const float scale = __fdiv_rn(1.0f, __fsqrt_rn(__fadd_rn(mean, eps)));
#pragma unroll
        for (int i = 0; i < NCH; i++) {
            const int k0 = 256 * i + 8 * lane;
            const float4 wa = *(const float4 *)(w + k0), wb = *(const float4 *)(w + k0 + 4);
            const float ww[8] = {wa.x, wa.y, wa.z, wa.w, wb.x, wb.y, wb.z, wb.w};
#pragma unroll
            for (int j = 0; j < 8; j++) v[i][j] = __fmul_rn(__fmul_rn(v[i][j], scale), ww[j]);
            if (xn_out) {
                *(float4 *)(xn_out + (int64_t)r * cols + k0) = make_float4(v[i][0], v[i][1], v[i][2], v[i][3]);
                *(float4 *)(xn_out + (int64_t)r * cols + k0 + 4) = make_float4(v[i][4], v[i][5], v[i][6], v[i][7]);
            }
        }
    } else {
#pragma unroll
        for (int i = 0; i < NCH; i++)
#pragma unroll
            for (int j = 0; j < 8; j++) v[i][j] = 0.f;
    }
#pragma unroll
    for (int i = 0; i < NCH; i++) {
        const int k0 = 256 * i + 8 * lane;
        __half *dst = tile0 + (k0 / GM_BK) * (int64_t)(GM_BM * GM_BK) + tile_elem_offset(r % GM_BM, k0 % GM_BK);
        const __half2 p4[4] = {gm_pack2(v[i][0], v[i][1]), gm_pack2(v[i][2], v[i][3]), gm_pack2(v[i][4], v[i][5]), gm_pack2(v[i][6], v[i][7])};
        *(int4 *)dst = *(const int4 *)p4;
    }
}

// BERT: x = LayerNorm(x + h + bias) * w + b (ggml_norm: mean, variance of the centred values), the row written back as
// fp32 (it is the next residual) AND as fp16 operand tiles of the next GEMM.  One CTA per token row, the row lives in
// registers (8 features per thread, cols <= 2048); rows >= n_tok of the last tile get zero tiles.
__global__ void __launch_bounds__(256)
bias_residual_ln_x_tiles_kernel(float *__restrict__ x, const float *__restrict__ h, const float *__restrict__ bias,
                                const float *__restrict__ w, const float *__restrict__ b, int n_tok, int64_t cols, float eps,
                                __half *__restrict__ out) {
    const int r = blockIdx.x; // < m_pad
    const int64_t k_tiles = cols / GM_BK;
    __half *tile0 = out + (int64_t)(r / GM_BM) * k_tiles * (int64_t)(GM_BM * GM_BK);
    __shared__ float red[2][8];
    const int64_t k0 = (int64_t)threadIdx.x * 8;
    const bool mine = k0 < cols;
    float v[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    if (r < n_tok) { // block-uniform
        float s = 0.f;
        if (mine) {
            float *xr = x + (int64_t)r * cols + k0;
            const float *hr = h + (int64_t)r * cols + k0;
            const float4 xa = *(const float4 *)xr, xb = *(const float4 *)(xr + 4);
            v[0] = xa.x; v[1] = xa.y; v[2] = xa.z; v[3] = xa.w; v[4] = xb.x; v[5] = xb.y; v[6] = xb.z; v[7] = xb.w;
            if (h) { // (h == nullptr: the producing GEMM's epilogue already added its output and bias into x)
                const float4 ha = *(const float4 *)hr, hb = *(const float4 *)(hr + 4);
                const float4 ba = *(const float4 *)(bias + k0), bb = *(const float4 *)(bias + k0 + 4);
                v[0] = xa.x + (ha.x + ba.x); v[1] = xa.y + (ha.y + ba.y); v[2] = xa.z + (ha.z + ba.z); v[3] = xa.w + (ha.w + ba.w);
                v[4] = xb.x + (hb.x + bb.x); v[5] = xb.y + (hb.y + bb.y); v[6] = xb.z + (hb.z + bb.z); v[7] = xb.w + (hb.w + bb.w);
            }
#pragma unroll
            for (int j = 0; j < 8; j++) s += v[j];
        }
        for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
        if ((threadIdx.x & 31) == 0) red[0][threadIdx.x >> 5] = s;
        __syncthreads();
        float tot = 0.f;
#pragma unroll
        for (int i = 0; i < 8; i++) tot += red[0][i];
        const float mean = tot / (float)cols;
        float s2 = 0.f;
        if (mine) {
#pragma unroll
            for (int j = 0; j < 8; j++) {
                v[j] -= mean;
                s2 += v[j] * v[j];
            }
        }
        for (int o = 16; o > 0; o >>= 1) s2 += __shfl_xor_sync(0xffffffffu, s2, o);
        if ((threadIdx.x & 31) == 0) red[1][threadIdx.x >> 5] = s2;
        __syncthreads();
        float tot2 = 0.f;
#pragma unroll
        for (int i = 0; i < 8; i++) tot2 += red[1][i];
        const float inv = rsqrtf(tot2 / (float)cols + eps);
        if (mine) {
            const float4 wa = *(const float4 *)(w + k0), wb = *(const float4 *)(w + k0 + 4);
            const float4 ca = *(const float4 *)(b + k0), cb = *(const float4 *)(b + k0 + 4);
            v[0] = v[0] * inv * wa.x + ca.x; v[1] = v[1] * inv * wa.y + ca.y; v[2] = v[2] * inv * wa.z + ca.z; v[3] = v[3] * inv * wa.w + ca.w;
            v[4] = v[4] * inv * wb.x + cb.x; v[5] = v[5] * inv * wb.y + cb.y; v[6] = v[6] * inv * wb.z + cb.z; v[7] = v[7] * inv * wb.w + cb.w;
            float *xr = x + (int64_t)r * cols + k0;
            *(float4 *)xr = make_float4(v[0], v[1], v[2], v[3]);
            *(float4 *)(xr + 4) = make_float4(v[4], v[5], v[6], v[7]);
        }
    }
    if (mine) {
        __half *dst = tile0 + (k0 / GM_BK) * (int64_t)(GM_BM * GM_BK) + tile_elem_offset(r % GM_BM, (int)(k0 % GM_BK));
        __half2 p4[4] = {gm_pack2(v[0], v[1]), gm_pack2(v[2], v[3]), gm_pack2(v[4], v[5]),
                                gm_pack2(v[6], v[7])};
        *(int4 *)dst = *(const int4 *)p4;
    }
}

// The same for rows of at most 1024 features (BERT-base: 768): ONE WARP per row, 8 rows per CTA.  With a CTA per row only
// cols / 8 of its 256 threads hold data (96 for BERT-base) and both reductions cross the CTA; here a lane holds up to four
// groups of 8 features (group i = features [256 i + 8 lane, +8)) and the reductions are warp shuffles.
__global__ void __launch_bounds__(256)
bias_residual_ln_x_tiles_warp_kernel(float *__restrict__ x, const float *__restrict__ h, const float *__restrict__ bias,
                                     const float *__restrict__ w, const float *__restrict__ b, int n_tok, int64_t cols, float eps,
                                     __half *__restrict__ out, int m_pad) {
    const int lane = threadIdx.x & 31;
    const int r = blockIdx.x * 8 + (threadIdx.x >> 5); // < m_pad (a multiple of 128)
    if (r >= m_pad) return;
    const int64_t k_tiles = cols / GM_BK;
    __half *tile0 = out + (int64_t)(r / GM_BM) * k_tiles * (int64_t)(GM_BM * GM_BK);
    float v[4][8];
#pragma unroll
    for (int i = 0; i < 4; i++)
#pragma unroll
        for (int j = 0; j < 8; j++) v[i][j] = 0.f;
    if (r < n_tok) { // warp-uniform
        float s = 0.f;
#pragma unroll
        for (int i = 0; i < 4; i++) {
            const int64_t k0 = 256 * i + 8 * lane;
            if (k0 < cols) {
                const float *xr = x + (int64_t)r * cols + k0, *hr = h + (int64_t)r * cols + k0;
                const float4 xa = *(const float4 *)xr, xb = *(const float4 *)(xr + 4);
                v[i][0] = xa.x; v[i][1] = xa.y; v[i][2] = xa.z; v[i][3] = xa.w; v[i][4] = xb.x; v[i][5] = xb.y; v[i][6] = xb.z; v[i][7] = xb.w;
                if (h) { // (h == nullptr: the producing GEMM's epilogue already added its output and bias into x)
                    const float4 ha = *(const float4 *)hr, hb = *(const float4 *)(hr + 4);
                    const float4 ba = *(const float4 *)(bias + k0), bb = *(const float4 *)(bias + k0 + 4);
                    v[i][0] = xa.x + (ha.x + ba.x); v[i][1] = xa.y + (ha.y + ba.y); v[i][2] = xa.z + (ha.z + ba.z); v[i][3] = xa.w + (ha.w + ba.w);
                    v[i][4] = xb.x + (hb.x + bb.x); v[i][5] = xb.y + (hb.y + bb.y); v[i][6] = xb.z + (hb.z + bb.z); v[i][7] = xb.w + (hb.w + bb.w);
                }
#pragma unroll
                for (int j = 0; j < 8; j++) s += v[i][j];
            }
        }
        for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
        const float mean = s / (float)cols;
        float s2 = 0.f;
#pragma unroll
        for (int i = 0; i < 4; i++)
            if (256 * i + 8 * lane < cols) {
#pragma unroll
                for (int j = 0; j < 8; j++) {
                    v[i][j] -= mean;
                    s2 += v[i][j] * v[i][j];
                }
            }
        for (int o = 16; o > 0; o >>= 1) s2 += __shfl_xor_sync(0xffffffffu, s2, o);
        const float inv = rsqrtf(s2 / (float)cols + eps);
#pragma unroll
        for (int i = 0; i < 4; i++) {
            const int64_t k0 = 256 * i + 8 * lane;
            if (k0 < cols) {
                const float4 wa = *(const float4 *)(w + k0), wb = *(const float4 *)(w + k0 + 4);
                const float4 ca = *(const float4 *)(b + k0), cb = *(const float4 *)(b + k0 + 4);
                v[i][0] = v[i][0] * inv * wa.x + ca.x; v[i][1] = v[i][1] * inv * wa.y + ca.y; v[i][2] = v[i][2] * inv * wa.z + ca.z; v[i][3] = v[i][3] * inv * wa.w + ca.w;
                v[i][4] = v[i][4] * inv * wb.x + cb.x; v[i][5] = v[i][5] * inv * wb.y + cb.y; v[i][6] = v[i][6] * inv * wb.z + cb.z; v[i][7] = v[i][7] * inv * wb.w + cb.w;
                float *xr = x + (int64_t)r * cols + k0;
                *(float4 *)xr = make_float4(v[i][0], v[i][1], v[i][2], v[i][3]);
                *(float4 *)(xr + 4) = make_float4(v[i][4], v[i][5], v[i][6], v[i][7]);
            }
        }
    }
#pragma unroll
    for (int i = 0; i < 4; i++) {
        const int64_t k0 = 256 * i + 8 * lane;
        if (k0 < cols) {
            __half *dst = tile0 + (k0 / GM_BK) * (int64_t)(GM_BM * GM_BK) + tile_elem_offset(r % GM_BM, (int)(k0 % GM_BK));
            __half2 p4[4] = {gm_pack2(v[i][0], v[i][1]), gm_pack2(v[i][2], v[i][3]), gm_pack2(v[i][4], v[i][5]), gm_pack2(v[i][6], v[i][7])};
            *(int4 *)dst = *(const int4 *)p4;
        }
    }
}

// silu(gate) * up (ggml arithmetic) fused with the fp16 tile conversion
__global__ void swiglu_x_tiles_kernel(const float *__restrict__ g, const float *__restrict__ u, int n_tok, int64_t cols,
                                      __half *__restrict__ out, int64_t m_pad) {
    const int64_t kc_per_row = cols / 8;
    const int64_t total = m_pad * kc_per_row;
    const int64_t k_tiles = cols / GM_BK;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t r = i / kc_per_row;
        const int k0 = (int)(i % kc_per_row) * 8;
        float v[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        if (r < n_tok) {
            const float4 ga = *(const float4 *)(g + r * cols + k0), gb = *(const float4 *)(g + r * cols + k0 + 4);
            const float4 ua = *(const float4 *)(u + r * cols + k0), ub = *(const float4 *)(u + r * cols + k0 + 4);
            const float gg[8] = {ga.x, ga.y, ga.z, ga.w, gb.x, gb.y, gb.z, gb.w}, uu[8] = {ua.x, ua.y, ua.z, ua.w, ub.x, ub.y, ub.z, ub.w};
#pragma unroll
            for (int j = 0; j < 8; j++) v[j] = __fmul_rn(__fdiv_rn(gg[j], __fadd_rn(1.0f, expf(-gg[j]))), uu[j]);
        }
        const int64_t mt = r / GM_BM, kt = k0 / GM_BK;
        __half *dst = out + (mt * k_tiles + kt) * (int64_t)(GM_BM * GM_BK) + tile_elem_offset((int)(r % GM_BM), k0 % GM_BK);
        __half2 p4[4] = {gm_pack2(v[0], v[1]), gm_pack2(v[2], v[3]), gm_pack2(v[4], v[5]),
                                gm_pack2(v[6], v[7])};
        *(int4 *)dst = *(const int4 *)p4;
    }
}

// gelu(x + bias) (ggml_gelu's tanh form) fused with the fp16 tile conversion: the BERT feed-forward activation
__global__ void bias_gelu_x_tiles_kernel(const float *__restrict__ x, const float *__restrict__ bias, int n_tok, int64_t cols,
                                         __half *__restrict__ out, int64_t m_pad) {
    const int64_t kc_per_row = cols / 8;
    const int64_t total = m_pad * kc_per_row;
    const int64_t k_tiles = cols / GM_BK;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t r = i / kc_per_row;
        const int k0 = (int)(i % kc_per_row) * 8;
        float v[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        if (r < n_tok) {
            const float4 xa = *(const float4 *)(x + r * cols + k0), xb = *(const float4 *)(x + r * cols + k0 + 4);
            const float4 ba = *(const float4 *)(bias + k0), bb = *(const float4 *)(bias + k0 + 4);
            const float t[8] = {xa.x + ba.x, xa.y + ba.y, xa.z + ba.z, xa.w + ba.w, xb.x + bb.x, xb.y + bb.y, xb.z + bb.z, xb.w + bb.w};
#pragma unroll
            for (int j = 0; j < 8; j++) v[j] = 0.5f * t[j] * (1.f + tanhf(0.7978845608028654f * t[j] * (1.f + 0.044715f * t[j] * t[j])));
        }
        const int64_t mt = r / GM_BM, kt = k0 / GM_BK;
        __half *dst = out + (mt * k_tiles + kt) * (int64_t)(GM_BM * GM_BK) + tile_elem_offset((int)(r % GM_BM), k0 % GM_BK);
        __half2 p4[4] = {gm_pack2(v[0], v[1]), gm_pack2(v[2], v[3]), gm_pack2(v[4], v[5]),
                                gm_pack2(v[6], v[7])};
        *(int4 *)dst = *(const int4 *)p4;
    }
}

// ------------------------------------------------------------------------------------------------ PTX helpers
__device__ __forceinline__ uint32_t gm_s32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void gm_mbar_init(uint64_t *bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(gm_s32(bar)), "r"(count));
}
__device__ __forceinline__ void gm_mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(gm_s32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void gm_mbar_arrive(uint64_t *bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(gm_s32(bar)) : "memory");
}
__device__ __forceinline__ void gm_mbar_wait(uint64_t *bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "GM_WAIT:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra GM_DONE;\n"
        "bra GM_WAIT;\n"
        "GM_DONE:\n"
        "}\n" ::"r"(gm_s32(bar)),
        "r"(parity)
        : "memory");
}
__device__ __forceinline__ void gm_bulk_g2s(void *dst, const void *src, uint32_t bytes, uint64_t *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(gm_s32(dst)),
                 "l"(src), "r"(bytes), "r"(gm_s32(bar))
                 : "memory");
}
// K-major, no swizzle: LBO = 128 B between core matrices along K, SBO = 1024 B between 8-row groups
__device__ __forceinline__ uint64_t gm_desc(uint32_t saddr, uint32_t lbo, uint32_t sbo) {
    return (uint64_t)((saddr & 0x3FFFF) >> 4) | ((uint64_t)(lbo >> 4) << 16) | ((uint64_t)(sbo >> 4) << 32) | (1ull << 46);
}
__device__ __forceinline__ void gm_mma(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n"
        ".reg .pred p, e;\n"
        "elect.sync _|e, 0xffffffff;\n"
        "setp.ne.b32 p, %4, 0;\n"
        "@e tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n"
        "}\n" ::"r"(tmem_d),
        "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}
// gm_mma / gm_commit are called by a CONVERGED warp and issue from one elected lane: the operands of tcgen05.mma live in
// uniform registers, and inside a `lane == 0` branch the compiler cannot prove them warp-uniform -- it then wraps every
// MMA in an election loop (R2UR.BROADCAST + BRA.U.ANY), about 85 cycles of issue per instruction
__device__ __forceinline__ void gm_commit(uint64_t *bar) {
    asm volatile("{\n.reg .pred e;\nelect.sync _|e, 0xffffffff;\n@e tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n}\n" ::"r"(gm_s32(bar))
                 : "memory");
}
__device__ __forceinline__ bool gm_elect() {
    uint32_t e;
    asm volatile("{\n.reg .pred p;\nelect.sync _|p, 0xffffffff;\nselp.u32 %0, 1, 0, p;\n}\n" : "=r"(e));
    return e != 0;
}
__device__ __forceinline__ void gm_tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]),
          "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]),
          "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]),
          "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr)
        : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

struct GemmJobDev {             // one weight matrix of a launch: several matrices that share the activation
    const __half *b_tiles; // [n_tiles][k_tiles][256 x 64]   (wq|wk|wv, gate|up) run as ONE launch
    float *out;            // fp32 rows, or fp16 rows when GemmArgs::out_f16
    int64_t out_stride;
    const float *bias;     // added to the fp16 rows (may be null)
    int N, n_tiles, nt0;          // nt0: first n-tile of this matrix in the launch's concatenated n-tile space
    int col0;                     // first column of this matrix in the split-K partial rows
};
struct GemmArgs {
    const __half *a_tiles; // [m_tiles][k_tiles][128 x 64]
    GemmJobDev job[3];
    int n_jobs;
    const float *residual;        // single-matrix launches only
    int64_t res_stride;
    // epilogue that feeds the next GEMM: instead of fp32 rows, write fp16 operand tiles [m_tiles][K2 / 64][128 x 64] of
    // 1: silu(gate) * up over an interleaved gate|up matrix (columns c and c + 128 of a tile; K2 = N_total / 2),
    // 2: gelu(acc + bias[col]) (BERT feed-forward; K2 = N).  ksplit == 1 only.
    int epi_tiles;
    int out_f16;           // plain epilogue: rows of fp16 (+ the matrix's bias) instead of fp32 -- BERT's q | k | v
    __half *x2_tiles;
    const float *bias;
    int M, N_total, k_tiles, m_tiles, n_tiles; // n_tiles: over all matrices
    // Which index runs fastest over the CTAs: the tiles that run at the same time share the operand tile of the OTHER index
    // out of L2, so the smaller operand is the one streamed repeatedly.  0: m fastest (a weight tile is shared, the
    // activations are re-read per weight tile -- prefill, where they fit L2); 1: n fastest (long batches against a small
    // matrix, BERT: 32768 x 3072 activations were read three times for ffn_down's three weight tiles).
    int n_fastest;
    uint32_t lbo, sbo;
    // split-K for small M (batched decode): with one m-tile and few n-tiles most SMs would have no work item, and the
    // weight tiles are the bytes that have to stream.  A work item is (m-tile, n-tile, k-slice); slices write fp32
    // partial tiles to `partial` [ksplit][M][N_total], summed in slice order by gemm_splitk_reduce_kernel.
    int ksplit, k_per_split;
    float *partial;
};

__global__ void gemm_splitk_reduce_kernel(const float *__restrict__ partial, int ksplit, int M, int N, int N_total, int col0,
                                          float *__restrict__ out, int64_t out_stride, const float *__restrict__ residual,
                                          int64_t res_stride) {
    const int64_t total = (int64_t)M * N / 4;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int row = (int)(i * 4 / N), col = (int)(i * 4 % N);
        const int64_t e = (int64_t)row * N_total + col0 + col;
        float4 acc = *(const float4 *)(partial + e);
        for (int k = 1; k < ksplit; k++) {
            const float4 p = *(const float4 *)(partial + (int64_t)k * M * N_total + e);
            acc.x += p.x; acc.y += p.y; acc.z += p.z; acc.w += p.w;
        }
        if (residual) {
            const float4 r = *(const float4 *)(residual + (int64_t)row * res_stride + col);
            acc.x += r.x; acc.y += r.y; acc.z += r.z; acc.w += r.w;
        }
        *(float4 *)(out + (int64_t)row * out_stride + col) = acc;
    }
}

// ------------------------------------------------------------------------------------------------ the GEMM
__global__ void __launch_bounds__(GM_THREADS, 1)
gemm_tcgen05_kernel(const GemmArgs g) {
    extern __shared__ __align__(1024) uint8_t gsm[];
    uint8_t *stages = gsm;
    uint64_t *full = (uint64_t *)(gsm + (size_t)GM_STAGES * GM_STAGE_BYTES);
    uint64_t *empty = full + GM_STAGES;
    uint64_t *tfull = empty + GM_STAGES;   // [2] accumulator ready
    uint64_t *tempty = tfull + 2;          // [2] accumulator drained
    uint32_t *tmem_slot = (uint32_t *)(tempty + 2);
    const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31; // (the shuffle marks it warp-uniform)

    if (threadIdx.x == 0) {
        for (int i = 0; i < GM_STAGES; i++) { gm_mbar_init(full + i, 1); gm_mbar_init(empty + i, 1); }
        for (int i = 0; i < 2; i++) { gm_mbar_init(tfull + i, 1); gm_mbar_init(tempty + i, 8); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 2) { // TMEM: 512 columns = two 128 x 256 fp32 accumulators
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(gm_s32(tmem_slot)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_base = __shfl_sync(0xffffffffu, *tmem_slot, 0);
    const int n_tiles_total = g.m_tiles * g.n_tiles * g.ksplit; // work items
    // kind::f16, D = f32 (bit 4), A = B = f16 (format 0), both K-major, N = 256, M = 128
    const uint32_t idesc = (1u << 4) | ((uint32_t)(GM_BN >> 3) << 17) | ((uint32_t)(GM_BM >> 4) << 24);

    if (warp == 0) {
        { // ===== producer: one bulk copy per operand tile (the whole warp walks the schedule, an elected lane issues)
            uint32_t it = 0;
            for (int tile = blockIdx.x; tile < n_tiles_total; tile += gridDim.x) {
                const int ks = tile / (g.m_tiles * g.n_tiles), t2 = tile % (g.m_tiles * g.n_tiles);
                const int mt = g.n_fastest ? t2 / g.n_tiles : t2 % g.m_tiles, nt = g.n_fastest ? t2 % g.n_tiles : t2 / g.m_tiles;
                const int jb = (g.n_jobs > 2 && nt >= g.job[2].nt0) ? 2 : (g.n_jobs > 1 && nt >= g.job[1].nt0) ? 1 : 0;
                const __half *bt = g.job[jb].b_tiles + (int64_t)(nt - g.job[jb].nt0) * g.k_tiles * (GM_BN * GM_BK);
                const int kt0 = ks * g.k_per_split, kt1 = min(g.k_tiles, kt0 + g.k_per_split);
                for (int kt = kt0; kt < kt1; kt++, it++) {
                    const int s = it % GM_STAGES;
                    gm_mbar_wait(empty + s, ((it / GM_STAGES) & 1) ^ 1);
                    uint8_t *sa = stages + (size_t)s * GM_STAGE_BYTES;
                    if (gm_elect()) {
                        gm_mbar_expect_tx(full + s, GM_STAGE_BYTES);
                        gm_bulk_g2s(sa, g.a_tiles + ((int64_t)mt * g.k_tiles + kt) * (GM_BM * GM_BK), GM_A_BYTES, full + s);
                        gm_bulk_g2s(sa + GM_A_BYTES, bt + (int64_t)kt * (GM_BN * GM_BK), GM_B_BYTES, full + s);
                    }
                }
            }
        }
    } else if (warp == 1) {
        { // ===== MMA issuer (whole warp, elected lane)
            uint32_t it = 0, acc_it = 0;
            for (int tile = blockIdx.x; tile < n_tiles_total; tile += gridDim.x, acc_it++) {
                const int acc = acc_it & 1;
                gm_mbar_wait(tempty + acc, ((acc_it >> 1) & 1) ^ 1);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const uint32_t d_tmem = tmem_base + acc * GM_BN;
                const int ks = tile / (g.m_tiles * g.n_tiles);
                const int kt0 = ks * g.k_per_split, kt1 = min(g.k_tiles, kt0 + g.k_per_split);
                for (int kt = kt0; kt < kt1; kt++, it++) {
                    const int s = it % GM_STAGES;
                    gm_mbar_wait(full + s, (it / GM_STAGES) & 1);
                    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                    const uint32_t sa = gm_s32(stages + (size_t)s * GM_STAGE_BYTES), sb = sa + GM_A_BYTES;
#pragma unroll
                    for (int k4 = 0; k4 < GM_BK / 16; k4++) { // one MMA = K 16 = two core matrices = 256 B
                        gm_mma(d_tmem, gm_desc(sa + k4 * 256, g.lbo, g.sbo), gm_desc(sb + k4 * 256, g.lbo, g.sbo), idesc,
                               ((kt - kt0) | k4) != 0);
                    }
                    gm_commit(empty + s); // frees the smem stage when these MMAs have read it
                }
                gm_commit(tfull + acc);   // accumulator complete
            }
        }
    } else if (warp >= 4) { // ===== epilogue: TMEM -> registers -> global
        const int q = warp & 3; // TMEM lane quarter this warp may access
        const int chalf = (warp - 4) >> 2; // ... and the half of the tile's columns it drains (the other warp of the quarter takes the rest)
        float4 *stage4 = (float4 *)(gsm + (size_t)GM_STAGES * GM_STAGE_BYTES + 1024 + (size_t)(warp - 4) * 4096);
        uint32_t acc_it = 0;
        for (int tile = blockIdx.x; tile < n_tiles_total; tile += gridDim.x, acc_it++) {
            const int ks = tile / (g.m_tiles * g.n_tiles), t2 = tile % (g.m_tiles * g.n_tiles);
            const int mt = g.n_fastest ? t2 / g.n_tiles : t2 % g.m_tiles, nt = g.n_fastest ? t2 % g.n_tiles : t2 / g.m_tiles;
            const int acc = acc_it & 1;
            const GemmJobDev &J = g.job[(g.n_jobs > 2 && nt >= g.job[2].nt0) ? 2 : (g.n_jobs > 1 && nt >= g.job[1].nt0) ? 1 : 0];
            gm_mbar_wait(tfull + acc, (acc_it >> 1) & 1);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            if (g.epi_tiles) { // activation fused, output = the next GEMM's operand tiles (every row of the tile is written: padding rows are zeros in, zeros out)
                const int ncol = g.epi_tiles == 1 ? GM_BN / 2 : GM_BN;
                const int64_t k2_tiles = (g.epi_tiles == 1 ? g.N_total / 2 : g.N_total) / GM_BK;
                for (int c0 = chalf * (ncol / 2); c0 < (chalf + 1) * (ncol / 2); c0 += 32) {
                    uint32_t r[32];
                    float v[32];
                    gm_tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + acc * GM_BN + c0, r);
                    const int64_t k2 = (int64_t)nt * ncol + c0; // feature index of v[0] in the next GEMM's K
                    if (g.epi_tiles == 1) {
                        uint32_t u[32];
                        gm_tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + acc * GM_BN + 128 + c0, u);
#pragma unroll
                        for (int j = 0; j < 32; j++) { // ggml's silu(gate) * up
                            const float gt = __uint_as_float(r[j]);
                            v[j] = __fmul_rn(__fdiv_rn(gt, __fadd_rn(1.0f, expf(-gt))), __uint_as_float(u[j]));
                        }
                    } else {
#pragma unroll
                        for (int j = 0; j < 32; j++) { // ggml_gelu's tanh form on acc + bias
                            const float t = __uint_as_float(r[j]) + g.bias[k2 + j];
                            v[j] = 0.5f * t * (1.f + tanhf(0.7978845608028654f * t * (1.f + 0.044715f * t * t)));
                        }
                    }
                    if (k2 < (int64_t)k2_tiles * GM_BK) {
                        __half *dst = g.x2_tiles + ((int64_t)mt * k2_tiles + k2 / GM_BK) * (int64_t)(GM_BM * GM_BK) +
                                      tile_elem_offset(q * 32 + lane, (int)(k2 % GM_BK));
#pragma unroll
                        for (int j8 = 0; j8 < 4; j8++) { // 8 features = one 16-byte core-matrix row; consecutive groups are 64 halves apart
                            __half2 pk[4];
#pragma unroll
                            for (int j = 0; j < 4; j++) pk[j] = gm_pack2(v[8 * j8 + 2 * j], v[8 * j8 + 2 * j + 1]);
                            *(int4 *)(dst + j8 * 64) = *(const int4 *)pk;
                        }
                    }
                }
            } else
            for (int c0 = chalf * (GM_BN / 2); c0 < (chalf + 1) * (GM_BN / 2); c0 += 32) {
                // A lane holds 32 columns of ONE row: stored like that, a warp's store touches 32 rows.  The 32 x 32 block goes
                // through this warp's 4 KB of shared memory instead (16-byte chunks, chunk index xor row: no bank conflicts
                // either way) and leaves as whole 128-byte row segments, four rows per instruction -- and the bias / residual
                // are read the same way.
                uint32_t r[32];
                gm_tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + acc * GM_BN + c0, r);
                const int col = (nt - J.nt0) * GM_BN + c0; // column inside this matrix
#pragma unroll
                for (int j = 0; j < 8; j++)
                    stage4[lane * 8 + (j ^ (lane & 7))] = make_float4(__uint_as_float(r[4 * j]), __uint_as_float(r[4 * j + 1]),
                                                                      __uint_as_float(r[4 * j + 2]), __uint_as_float(r[4 * j + 3]));
                __syncwarp();
                const int row_base = mt * GM_BM + q * 32;
                if (g.out_f16) {
                    const int cc = (lane & 3) * 8; // 8 columns = 16 bytes of fp16 per lane, 8 rows per instruction
                    const bool vec = J.out_stride % 8 == 0 && col + cc + 8 <= J.N;
#pragma unroll
                    for (int i = 0; i < 4; i++) {
                        const int rr = i * 8 + (lane >> 2), grow = row_base + rr;
                        const float4 v0 = stage4[rr * 8 + ((2 * (lane & 3)) ^ (rr & 7))], v1 = stage4[rr * 8 + ((2 * (lane & 3) + 1) ^ (rr & 7))];
                        float v[8] = {v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w};
                        if (grow < g.M && col + cc < J.N) {
                            __half *oh = (__half *)J.out + (int64_t)grow * J.out_stride + col + cc;
                            if (J.bias) {
#pragma unroll
                                for (int e = 0; e < 8; e++)
                                    if (col + cc + e < J.N) v[e] += J.bias[col + cc + e];
                            }
                            if (vec) {
                                const __half2 pk[4] = {gm_pack2(v[0], v[1]), gm_pack2(v[2], v[3]), gm_pack2(v[4], v[5]), gm_pack2(v[6], v[7])};
                                *(int4 *)oh = *(const int4 *)pk;
                            } else {
                                for (int e = 0; e < 8 && col + cc + e < J.N; e++) oh[e] = __float2half_rn(fminf(fmaxf(v[e], -65504.f), 65504.f));
                            }
                        }
                    }
                } else {
                    const int cc = (lane & 7) * 4; // 4 columns = 16 bytes of fp32 per lane, 4 rows per instruction
                    const bool split = g.ksplit > 1;
                    const int64_t ostride = split ? g.N_total : J.out_stride;
                    float *obase = split ? g.partial + (int64_t)ks * g.M * g.N_total + J.col0 : J.out;
                    const float *rbase = (g.residual && !split) ? g.residual : nullptr;
                    const float *bj = (J.bias && !split) ? J.bias : nullptr;
                    const bool vec = ((ostride | (split ? J.col0 : 0) | (rbase ? g.res_stride : 0)) % 4) == 0 && col + cc + 4 <= J.N;
#pragma unroll
                    for (int i = 0; i < 8; i++) {
                        const int rr = i * 4 + (lane >> 3), grow = row_base + rr;
                        float4 v = stage4[rr * 8 + ((lane & 7) ^ (rr & 7))];
                        if (grow < g.M && col + cc < J.N) {
                            float *o = obase + (int64_t)grow * ostride + col + cc;
                            const float *res = rbase ? rbase + (int64_t)grow * g.res_stride + col + cc : nullptr;
                            if (vec) {
                                if (bj) {
                                    const float4 bb = *(const float4 *)(bj + col + cc);
                                    v.x += bb.x; v.y += bb.y; v.z += bb.z; v.w += bb.w;
                                }
                                if (res) {
                                    const float4 rv = *(const float4 *)res;
                                    v.x += rv.x; v.y += rv.y; v.z += rv.z; v.w += rv.w;
                                }
                                *(float4 *)o = v;
                            } else {
                                const float ve[4] = {v.x, v.y, v.z, v.w};
                                for (int e = 0; e < 4 && col + cc + e < J.N; e++) o[e] = ve[e] + (bj ? bj[col + cc + e] : 0.f) + (res ? res[e] : 0.f);
                            }
                        }
                    }
                }
                __syncwarp(); // the block is out before the next one overwrites the staging rows
            }
            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
            __syncwarp();
            if (lane == 0) gm_mbar_arrive(tempty + acc);
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 2) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem_base) : "memory");
}

// ------------------------------------------------------------------------------------------------ host side
size_t gemm_w_tiles_bytes(int64_t rows, int64_t cols) { return (size_t)align_up(rows, GM_BN) * cols * 2; }
size_t gemm_x_tiles_bytes(int n_tok, int64_t cols) { return (size_t)align_up(n_tok, GM_BM) * cols * 2; }

void launch_gemm_prepare_x(const float *x, int64_t x_stride, int n_tok, int64_t K, void *x_tiles, cudaStream_t s) {
    B2_CHECK(K % GM_BK == 0, B2_EINVAL, "gemm: K must be a multiple of 64");
    const int64_t m_pad = align_up(n_tok, GM_BM);
    const int64_t total = m_pad * (K / 8);
    const int blocks = (int)std::min<int64_t>((total + 255) / 256, 148 * 16);
    convert_x_tiles_kernel<<<blocks, 256, 0, s>>>(x, x_stride, n_tok, K, (__half *)x_tiles, m_pad);
    g_launch_count++;
    B2_CUDA(cudaGetLastError());
}

void launch_gemm_prepare_x_rmsnorm(const float *x, int64_t x_stride, const float *w, int n_tok, int64_t K, float eps, void *x_tiles,
                                   float *xn_out, cudaStream_t s) {
    B2_CHECK(K % GM_BK == 0, B2_EINVAL, "gemm: K must be a multiple of 64");
    const int m_pad = (int)align_up(n_tok, GM_BM);
    static const bool row_per_cta = getenv("B200LLM_RMSNORM_ROW_PER_CTA") && atoi(getenv("B200LLM_RMSNORM_ROW_PER_CTA")); // A/B
    const bool warp_rows = m_pad >= 1024 && x_stride % 4 == 0 && !row_per_cta;
    if (warp_rows && K == 4096) rmsnorm_x_tiles_warp_kernel<16><<<m_pad / 4, 128, 0, s>>>(x, x_stride, w, n_tok, eps, (__half *)x_tiles, xn_out, m_pad);
    else if (warp_rows && K == 2048) rmsnorm_x_tiles_warp_kernel<8><<<m_pad / 4, 128, 0, s>>>(x, x_stride, w, n_tok, eps, (__half *)x_tiles, xn_out, m_pad);
    else
    rmsnorm_x_tiles_kernel<<<m_pad, 256, 0, s>>>(x, x_stride, w, n_tok, K, eps, (__half *)x_tiles, xn_out);
    g_launch_count++;
    B2_CUDA(cudaGetLastError());
}

void launch_gemm_prepare_x_swiglu(const float *gate, const float *up, int n_tok, int64_t K, void *x_tiles, cudaStream_t s) {
    B2_CHECK(K % GM_BK == 0, B2_EINVAL, "gemm: K must be a multiple of 64");
    const int64_t m_pad = align_up(n_tok, GM_BM);
    const int64_t total = m_pad * (K / 8);
    const int blocks = (int)std::min<int64_t>((total + 255) / 256, 148 * 16);
    swiglu_x_tiles_kernel<<<blocks, 256, 0, s>>>(gate, up, n_tok, K, (__half *)x_tiles, m_pad);
    g_launch_count++;
    B2_CUDA(cudaGetLastError());
}

void launch_gemm_prepare_x_bias_residual_ln(float *x, const float *h, const float *bias, const float *w, const float *b, int n_tok,
                                            int64_t K, float eps, void *x_tiles, cudaStream_t s) {
    B2_CHECK(K % GM_BK == 0 && K <= 2048, B2_EINVAL, "gemm: LayerNorm rows must be a multiple of 64 and at most 2048 wide");
    const int m_pad = (int)align_up(n_tok, GM_BM);
    static const bool row_per_cta = getenv("B200LLM_BERT_LN_ROW_PER_CTA") && atoi(getenv("B200LLM_BERT_LN_ROW_PER_CTA")); // A/B
    // (measured on BERT-base: 16 x 512 tokens per call 3713 against 3655 seq/s; one 512-token text 675 against 696 -- too few CTAs)
    if (K <= 1024 && K % 8 == 0 && m_pad >= 2048 && !row_per_cta) bias_residual_ln_x_tiles_warp_kernel<<<m_pad / 8, 256, 0, s>>>(x, h, bias, w, b, n_tok, K, eps, (__half *)x_tiles, m_pad);
    else bias_residual_ln_x_tiles_kernel<<<m_pad, 256, 0, s>>>(x, h, bias, w, b, n_tok, K, eps, (__half *)x_tiles);
    g_launch_count++;
    B2_CUDA(cudaGetLastError());
}

void launch_gemm_prepare_x_bias_gelu(const float *x, const float *bias, int n_tok, int64_t K, void *x_tiles, cudaStream_t s) {
    B2_CHECK(K % GM_BK == 0, B2_EINVAL, "gemm: K must be a multiple of 64");
    const int64_t m_pad = align_up(n_tok, GM_BM);
    const int64_t total = m_pad * (K / 8);
    const int blocks = (int)std::min<int64_t>((total + 255) / 256, 148 * 16);
    bias_gelu_x_tiles_kernel<<<blocks, 256, 0, s>>>(x, bias, n_tok, K, (__half *)x_tiles, m_pad);
    g_launch_count++;
    B2_CUDA(cudaGetLastError());
}

void launch_gemm_dequant_w(const DTensor &W, void *w_tiles, cudaStream_t s) {
    B2_CHECK(W.cols % GM_BK == 0, B2_EINVAL, "gemm: K must be a multiple of 64");
    const int64_t rows_pad = align_up(W.rows, GM_BN);
    const int64_t total = rows_pad * (W.cols / 8);
    const int blocks = (int)std::min<int64_t>((total + 255) / 256, 148 * 32);
    dequant_tiles_kernel<<<blocks, 256, 0, s>>>(W.data, W.type, W.rows, W.cols, W.row_stride, (__half *)w_tiles, rows_pad, -1);
    g_launch_count++;
    B2_CUDA(cudaGetLastError());
}

void launch_gemm_tc_multi(const GemmJob *jobs, int n_jobs, const void *x_tiles, int n_tok, const float *residual, int64_t res_stride,
                          float *splitk_scratch, size_t splitk_bytes, cudaStream_t s, bool out_f16) {
    B2_CHECK(n_jobs >= 1 && n_jobs <= 3, B2_EINVAL, "gemm: 1 to 3 matrices per launch");
    B2_CHECK(!out_f16 || !residual, B2_EINVAL, "gemm: fp16 rows carry a bias, not a residual");
    B2_CHECK(!residual || n_jobs == 1, B2_EINVAL, "gemm: the residual add needs a single matrix");
    const int64_t cols = jobs[0].W->cols;
    B2_CHECK(cols % GM_BK == 0, B2_EINVAL, "gemm: K must be a multiple of 64");
    static PerDeviceOnce configured;
    configured.run([&] { B2_CUDA(cudaFuncSetAttribute(gemm_tcgen05_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)GM_SMEM)); });
    static const bool umma_swap = getenv("B200LLM_UMMA_SWAP") && atoi(getenv("B200LLM_UMMA_SWAP"));
    const uint32_t lbo = umma_swap ? 1024 : 128, sbo = umma_swap ? 128 : 1024;
    GemmArgs g;
    g.a_tiles = (const __half *)x_tiles;
    g.n_jobs = n_jobs;
    int nt = 0, col = 0;
    bool rows4 = true;
    for (int j = 0; j < n_jobs; j++) {
        const DTensor &W = *jobs[j].W;
        B2_CHECK(W.cols == cols, B2_EINVAL, "gemm: matrices of one launch share K");
        if (!jobs[j].w_tiles_ready) launch_gemm_dequant_w(W, jobs[j].w_tiles, s); // quantised W -> fp16 tiles (unless kept resident)
        g.job[j].b_tiles = (const __half *)jobs[j].w_tiles;
        g.job[j].out = jobs[j].out; g.job[j].out_stride = jobs[j].out_stride; g.job[j].bias = jobs[j].bias;
        B2_CHECK(!out_f16 || jobs[j].out_stride % 8 == 0, B2_EINVAL, "gemm: fp16 rows must be 16-byte aligned");
        g.job[j].N = (int)W.rows; g.job[j].n_tiles = (int)(align_up(W.rows, GM_BN) / GM_BN);
        g.job[j].nt0 = nt; g.job[j].col0 = col;
        nt += g.job[j].n_tiles; col += (int)W.rows;
        rows4 = rows4 && W.rows % 4 == 0;
    }
    g.residual = residual; g.res_stride = res_stride;
    g.epi_tiles = 0; g.out_f16 = out_f16 ? 1 : 0; g.x2_tiles = nullptr; g.bias = nullptr;
    g.M = n_tok; g.N_total = col; g.k_tiles = (int)(cols / GM_BK);
    g.m_tiles = (int)(align_up(n_tok, GM_BM) / GM_BM); g.n_tiles = nt;
    g.lbo = lbo; g.sbo = sbo;
    g.n_fastest = (int64_t)g.m_tiles * GM_BM > (int64_t)g.n_tiles * GM_BN ? 1 : 0;
    g.ksplit = 1; g.k_per_split = g.k_tiles; g.partial = nullptr;
    const int tiles = g.m_tiles * g.n_tiles;
    bool any_bias = false;
    for (int j = 0; j < n_jobs; j++) any_bias = any_bias || jobs[j].bias;
    if (!out_f16 && !any_bias && splitk_scratch && g.m_tiles == 1 && tiles <= 148 / 2 && rows4) { // at least 2 slices per tile or not worth a reduce pass
        int ks = std::min(148 / tiles, g.k_tiles / 4);
        while (ks > 1 && (size_t)ks * n_tok * g.N_total * sizeof(float) > splitk_bytes) ks--;
        if (ks > 1) {
            g.k_per_split = (g.k_tiles + ks - 1) / ks;
            g.ksplit = (g.k_tiles + g.k_per_split - 1) / g.k_per_split;
            g.partial = splitk_scratch;
        }
    }
    const int grid = std::min(tiles * g.ksplit, 148);
    gemm_tcgen05_kernel<<<grid, GM_THREADS, GM_SMEM, s>>>(g);
    g_launch_count++;
    B2_CUDA(cudaGetLastError());
    if (g.ksplit > 1) {
        for (int j = 0; j < n_jobs; j++) {
            const int64_t total = (int64_t)n_tok * g.job[j].N / 4;
            const int blocks = (int)std::min<int64_t>((total + 255) / 256, 148 * 8);
            gemm_splitk_reduce_kernel<<<blocks, 256, 0, s>>>(g.partial, g.ksplit, n_tok, g.job[j].N, g.N_total, g.job[j].col0, g.job[j].out,
                                                             g.job[j].out_stride, residual, res_stride);
            g_launch_count++;
            B2_CUDA(cudaGetLastError());
        }
    }
}

// ---- GEMMs whose epilogue applies the activation and writes the NEXT GEMM's operand tiles (no fp32 round trip through HBM)
size_t gemm_w_pair_tiles_bytes(int64_t rows, int64_t cols) { return (size_t)align_up(rows, 128) * 2 * cols * 2; }

void launch_gemm_dequant_w_pair(const DTensor &gate, const DTensor &up, void *w_tiles, cudaStream_t s) {
    B2_CHECK(gate.rows == up.rows && gate.cols == up.cols && gate.cols % GM_BK == 0, B2_EINVAL, "gemm: gate / up shapes differ");
    const int64_t rows_pad = align_up(gate.rows, 128);
    const int64_t total = rows_pad * (gate.cols / 8);
    const int blocks = (int)std::min<int64_t>((total + 255) / 256, 148 * 32);
    dequant_tiles_kernel<<<blocks, 256, 0, s>>>(gate.data, gate.type, gate.rows, gate.cols, gate.row_stride, (__half *)w_tiles, rows_pad, 0);
    dequant_tiles_kernel<<<blocks, 256, 0, s>>>(up.data, up.type, up.rows, up.cols, up.row_stride, (__half *)w_tiles, rows_pad, 1);
    g_launch_count += 2;
    B2_CUDA(cudaGetLastError());
}

static void launch_gemm_fused_epilogue(int epi, const void *w_tiles, int64_t N, int64_t cols, const void *x_tiles, int n_tok, const float *bias,
                                       void *x2_tiles, cudaStream_t s) {
    B2_CHECK(cols % GM_BK == 0 && N % GM_BN == 0, B2_EINVAL, "gemm: fused epilogue needs whole 256-column tiles");
    static PerDeviceOnce configured;
    configured.run([&] { B2_CUDA(cudaFuncSetAttribute(gemm_tcgen05_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)GM_SMEM)); });
    GemmArgs g;
    memset(&g, 0, sizeof(g));
    g.a_tiles = (const __half *)x_tiles;
    g.n_jobs = 1;
    g.job[0].b_tiles = (const __half *)w_tiles;
    g.job[0].N = (int)N; g.job[0].n_tiles = (int)(N / GM_BN); g.job[0].nt0 = 0; g.job[0].col0 = 0;
    g.M = n_tok; g.N_total = (int)N; g.k_tiles = (int)(cols / GM_BK);
    g.m_tiles = (int)(align_up(n_tok, GM_BM) / GM_BM); g.n_tiles = g.job[0].n_tiles;
    g.n_fastest = (int64_t)g.m_tiles * GM_BM > (int64_t)g.n_tiles * GM_BN ? 1 : 0;
    g.lbo = 128; g.sbo = 1024;
    g.ksplit = 1; g.k_per_split = g.k_tiles;
    g.epi_tiles = epi; g.x2_tiles = (__half *)x2_tiles; g.bias = bias;
    const int tiles = g.m_tiles * g.n_tiles;
    gemm_tcgen05_kernel<<<std::min(tiles, 148), GM_THREADS, GM_SMEM, s>>>(g);
    g_launch_count++;
    B2_CUDA(cudaGetLastError());
}

// x2 tiles [n_tok][ff] = silu(W_gate x) * (W_up x); pair_tiles from launch_gemm_dequant_w_pair (ff a multiple of 128)
void launch_gemm_tc_swiglu(int64_t ff, int64_t cols, const void *pair_tiles, const void *x_tiles, int n_tok, void *x2_tiles, cudaStream_t s) {
    B2_CHECK(ff % 128 == 0, B2_EINVAL, "gemm: fused SwiGLU needs ffn_gate rows in whole 128-row tiles");
    launch_gemm_fused_epilogue(1, pair_tiles, 2 * ff, cols, x_tiles, n_tok, nullptr, x2_tiles, s);
}
// x2 tiles [n_tok][N] = gelu(W x + bias)  (N a multiple of 256)
void launch_gemm_tc_bias_gelu(int64_t N, int64_t cols, const void *w_tiles, const void *x_tiles, int n_tok, const float *bias, void *x2_tiles,
                              cudaStream_t s) {
    launch_gemm_fused_epilogue(2, w_tiles, N, cols, x_tiles, n_tok, bias, x2_tiles, s);
}

void launch_gemm_tc(const DTensor &W, const void *x_tiles, int n_tok, float *out, int64_t out_stride, const float *residual,
                    int64_t res_stride, void *w_tiles, bool w_tiles_ready, float *splitk_scratch, size_t splitk_bytes, cudaStream_t s) {
    GemmJob j{&W, w_tiles, w_tiles_ready, out, out_stride};
    launch_gemm_tc_multi(&j, 1, x_tiles, n_tok, residual, res_stride, splitk_scratch, splitk_bytes, s);
}

} // namespace b2
