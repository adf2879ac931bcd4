// kernels_qgemm.cu -- batched decode mat-mul (2..64 tokens): the quantised weights are dequantised INSIDE the GEMM.
//
// out[tok][row] = W[row] . x[tok] (+ residual) for W in GGUF block formats (Q4_K, Q5_K, Q6_K, Q8_0), x in fp16.
// At 64 tokens the step is bound by streaming the weights once from HBM (ggml-cuda's mul_mat_q does the same job on
// mma.sync; the reference calls it through llama.cpp's ggml_mul_mat, ggml/src/ggml-cuda/mmq.cu).  The round-1 path read
// 3.3 x the bytes (fp16 shadow tiles); this kernel reads the 4.5-bit blocks and never writes a dequantised copy:
//
//   HBM --cp.async.bulk--> raw block ring (smem) --dequant warps--> fp16 A operand in TMEM --tcgen05.mma--> D in TMEM
//                                                 x tiles (fp16) --cp.async.bulk--> smem B operand ----^
//
// * "swap-AB": the 128 weight rows of a tile are the MMA's M, the tokens its N (16..64), so no tensor-core work is
//   spent on padding rows; the A operand comes from TMEM (tcgen05.mma [d], [a], b-desc), written by the dequant warps
//   with tcgen05.st -- one thread per weight row, which is also how the GGUF blocks are laid out.  The dequantised
//   values never touch shared memory (its bandwidth, not the tensor core, would bound an SS-mode design at N = 64).
// * weights are repacked once per model into tile-major chunks: chunk(tile, kb) = the 256-element super-blocks kb of
//   the tile's 128 rows, contiguous (18 KB for Q4_K), so a ring slot is ONE bulk copy.
// * work is split stream-K style: the (panel, kb) units, k fastest, are cut into gridDim.x equal contiguous ranges.  A
//   panel is T = 1 or 2 tiles that share the activation stage.  A range that covers a panel's whole K writes the
//   output directly; otherwise it writes an fp32 partial tile, and the LAST segment of that tile to arrive (one
//   atomic counter per tile and lane quarter) adds all partials in segment order (+ residual) -- deterministic, no
//   second launch, no spinning.
// * warps: 0 weight producer, 3 activation producer (bulk copies), 1 MMA issuer, 2 TMEM allocator, 4..7 epilogue (one TMEM
//   lane quarter each),
//   8..23 dequant (lane quarter = warp % 4; (warp - 8) / 4 = which 32 of the 128 k of an operand stage).
// fp16 arithmetic of the dequantisation: q exact, d*sc and dmin*m rounded to fp16, one fused multiply-add -- about 3
// fp16 roundings per weight (relative 2^-11 each), (the large-batch path, kernels_gemm.cu, rounds the dequantised fp32 value once: 2^-11).
// HBM-bound: algorithmic bytes = the quantised weights once; roofline = MEASURED_PEAKS.json hbm_gbs.
#include <cuda_fp16.h>

#include <vector>

#include "attention.cuh"
#include "kernels.h"
#include "quant.cuh"

namespace b2 {

constexpr int QG_ROWS = 128;     // weight rows per tile = MMA M = TMEM lanes
constexpr int QG_THREADS = 768;   // 24 warps: 0 producer, 1 MMA, 2 TMEM allocator, 4..7 epilogue, 8..23 dequant
constexpr int QG_DQ_WARPS = 16;
constexpr int QG_B_STAGES = 3;   // activation stages (256 k each)
constexpr int QG_A_STAGES = 4;   // TMEM operand stages (128 k = 64 columns each)
constexpr int QG_MAX_RAW = 6;    // raw ring slots at most
constexpr int QG_RAW_RING = 108 * 1024;
constexpr int QG_D_COL = 0;      // TMEM columns: accumulators [buf 2][t 2][64], then the A stages
constexpr int QG_A_COL = 256;

__host__ __device__ inline int qg_chunk_bytes(int type) {
    switch (type) {
        case B2_T_Q4_K: return QG_ROWS * 144;
        case B2_T_Q5_K: return QG_ROWS * 176;
        case B2_T_Q6_K: return QG_ROWS * 208 + QG_ROWS * 2; // [128][ql 128 | qh 64 | sc 16] then the 128 fp16 d
        case B2_T_Q8_0: return QG_ROWS * 272;               // [128][qs 256 | 8 x fp16 d]  (272 = 17 x 16: conflict-free 16-byte row loads)
    }
    return 0;
}
bool qgemm_type_supported(int type) { return qg_chunk_bytes(type) != 0; }
size_t qgemm_w_bytes(int type, int64_t rows, int64_t cols) {
    return (size_t)(align_up(rows, QG_ROWS) / QG_ROWS) * (size_t)(cols / 256) * qg_chunk_bytes(type);
}
size_t qgemm_x_bytes(int n_tok, int64_t cols) { return (size_t)align_up(n_tok, 16) * cols * 2; }

// ------------------------------------------------------------------------------------------------ weight repack
// device-layout rows (common.h: Q4_K / Q5_K raw GGUF blocks; Q6_K 208-byte blocks + d array; Q8_0 qs + d array) -> chunks
__global__ void qgemm_repack_kernel(const uint8_t *__restrict__ W, int type, int64_t rows, int64_t cols, int64_t row_stride,
                                    uint8_t *__restrict__ out) {
    const int nb = (int)(cols / 256);
    const int cb = qg_chunk_bytes(type);
    const int64_t n_chunks = (rows + QG_ROWS - 1) / QG_ROWS * nb;
    for (int64_t c = blockIdx.x; c < n_chunks; c += gridDim.x) {
        const int64_t tile = c / nb;
        const int kb = (int)(c % nb);
        uint8_t *dst = out + c * cb;
        const int bb = type == B2_T_Q4_K ? 144 : type == B2_T_Q5_K ? 176 : type == B2_T_Q6_K ? 208 : 272;
        const int per_row = bb / 16;
        for (int i = threadIdx.x; i < QG_ROWS * per_row; i += blockDim.x) {
            const int r = i / per_row, p = i % per_row;
            const int64_t R = tile * QG_ROWS + r;
            int4 v = make_int4(0, 0, 0, 0);
            if (R < rows) {
                const uint8_t *row = W + R * row_stride;
                if (type == B2_T_Q8_0) {
                    if (p < 16) v = *(const int4 *)(row + (size_t)kb * 256 + p * 16);
                    else v = *(const int4 *)(row + cols + (size_t)kb * 16); // the 8 block scales of this super-block
                } else {
                    v = *(const int4 *)(row + (size_t)kb * bb + p * 16);
                }
            }
            *(int4 *)(dst + (size_t)r * bb + p * 16) = v;
        }
        if (type == B2_T_Q6_K) {
            for (int r = threadIdx.x; r < QG_ROWS; r += blockDim.x) {
                const int64_t R = tile * QG_ROWS + r;
                uint16_t d = 0;
                if (R < rows) d = *(const uint16_t *)(W + R * row_stride + (size_t)nb * Q6_BLK + 2 * kb);
                *(uint16_t *)(dst + QG_ROWS * 208 + 2 * r) = d;
            }
        }
    }
}

void launch_qgemm_repack(const DTensor &W, void *q_tiles, cudaStream_t s) {
    B2_CHECK(qgemm_type_supported(W.type) && W.cols % 256 == 0, B2_EINVAL, "qgemm: unsupported weight type or K not a multiple of 256");
    const int64_t n_chunks = align_up(W.rows, QG_ROWS) / QG_ROWS * (W.cols / 256);
    qgemm_repack_kernel<<<(int)std::min<int64_t>(n_chunks, 148 * 16), 256, 0, s>>>(W.data, W.type, W.rows, W.cols, W.row_stride,
                                                                                 (uint8_t *)q_tiles);
    g_launch_count++;
    B2_CUDA(cudaGetLastError());
}

// ------------------------------------------------------------------------------------------------ activation tiles
// x tiles: [kb = K/256][tok/8][k/8 = 32][8 tok][8 k] fp16 -- one 256-k stage of all (padded) tokens is contiguous, and a
// K = 16 MMA step is two adjacent 128-byte core matrices (LBO 128 B along K, SBO 4096 B between 8-token groups).
// (qg_x_offset: attention.cuh -- the fused decode-attention kernel writes these tiles too)
__device__ __forceinline__ void qg_store8(__half *out, int r, int64_t k0, int n_pad, const float (&v)[8]) {
    __half2 p[4];
#pragma unroll
    for (int j = 0; j < 4; j++) { // operand order (0, 2, 1, 3) inside each group of four k, as the weights; saturate: an fp16 inf would poison the accumulator
        const int a = (j >> 1) * 4 + (j & 1), b = a + 2;
        p[j] = __floats2half2_rn(fminf(fmaxf(v[a], -65504.f), 65504.f), fminf(fmaxf(v[b], -65504.f), 65504.f));
    }
    *(int4 *)(out + qg_x_offset(r, k0, n_pad)) = *(const int4 *)p;
}

__global__ void qgemm_convert_x_kernel(const float *__restrict__ x, int64_t x_stride, int n_tok, int64_t cols, __half *__restrict__ out,
                                       int n_pad) {
    B2_PDL_TRIGGER();
    B2_PDL_WAIT();
    const int64_t kc = cols / 8, total = (int64_t)n_pad * kc;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int r = (int)(i / kc);
        const int64_t k0 = (i % kc) * 8;
        float v[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        if (r < n_tok) {
            const float4 a = *(const float4 *)(x + r * x_stride + k0), b = *(const float4 *)(x + r * x_stride + k0 + 4);
            v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
        }
        qg_store8(out, r, k0, n_pad, v);
    }
}

// RMSNorm (ggml arithmetic, as rmsnorm_x_tiles_kernel) -> fp16 tiles; one CTA per padded token row
__global__ void __launch_bounds__(256)
qgemm_rmsnorm_x_kernel(const float *__restrict__ x, int64_t x_stride, const float *__restrict__ w, int n_tok, int64_t cols, float eps,
                       __half *__restrict__ out, int n_pad) {
    B2_PDL_TRIGGER();
    B2_PDL_WAIT();
    const int r = blockIdx.x;
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
        }
        qg_store8(out, r, k0, n_pad, v);
    }
}

__global__ void qgemm_swiglu_x_kernel(const float *__restrict__ g, const float *__restrict__ u, int n_tok, int64_t cols,
                                      __half *__restrict__ out, int n_pad) {
    B2_PDL_TRIGGER();
    B2_PDL_WAIT();
    const int64_t kc = cols / 8, total = (int64_t)n_pad * kc;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int r = (int)(i / kc);
        const int64_t k0 = (i % kc) * 8;
        float v[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        if (r < n_tok) {
            const float4 ga = *(const float4 *)(g + r * cols + k0), gb = *(const float4 *)(g + r * cols + k0 + 4);
            const float4 ua = *(const float4 *)(u + r * cols + k0), ub = *(const float4 *)(u + r * cols + k0 + 4);
            const float gg[8] = {ga.x, ga.y, ga.z, ga.w, gb.x, gb.y, gb.z, gb.w}, uu[8] = {ua.x, ua.y, ua.z, ua.w, ub.x, ub.y, ub.z, ub.w};
#pragma unroll
            for (int j = 0; j < 8; j++) v[j] = __fmul_rn(__fdiv_rn(gg[j], __fadd_rn(1.0f, expf(-gg[j]))), uu[j]);
        }
        qg_store8(out, r, k0, n_pad, v);
    }
}

void launch_qgemm_prepare_x(const float *x, int64_t x_stride, int n_tok, int64_t K, void *x_tiles, cudaStream_t s) {
    B2_CHECK(K % 256 == 0 && n_tok >= 1 && n_tok <= QGEMM_MAX_TOKENS, B2_EINVAL, "qgemm: K must be a multiple of 256, 1..64 tokens");
    const int n_pad = (int)align_up(n_tok, 16);
    const int64_t total = (int64_t)n_pad * (K / 8);
    launch_pdl(qgemm_convert_x_kernel, dim3((unsigned)std::min<int64_t>((total + 255) / 256, 148 * 8)), dim3(256), 0, s, x, x_stride, n_tok, K, (__half *)x_tiles, n_pad);
    g_launch_count++;
    B2_CUDA(cudaGetLastError());
}
void launch_qgemm_prepare_x_rmsnorm(const float *x, int64_t x_stride, const float *w, int n_tok, int64_t K, float eps, void *x_tiles,
                                    cudaStream_t s) {
    B2_CHECK(K % 256 == 0 && n_tok >= 1 && n_tok <= QGEMM_MAX_TOKENS, B2_EINVAL, "qgemm: K must be a multiple of 256, 1..64 tokens");
    const int n_pad = (int)align_up(n_tok, 16);
    launch_pdl(qgemm_rmsnorm_x_kernel, dim3(n_pad), dim3(256), 0, s, x, x_stride, w, n_tok, K, eps, (__half *)x_tiles, n_pad);
    g_launch_count++;
    B2_CUDA(cudaGetLastError());
}
void launch_qgemm_prepare_x_swiglu(const float *gate, const float *up, int n_tok, int64_t K, void *x_tiles, cudaStream_t s) {
    B2_CHECK(K % 256 == 0 && n_tok >= 1 && n_tok <= QGEMM_MAX_TOKENS, B2_EINVAL, "qgemm: K must be a multiple of 256, 1..64 tokens");
    const int n_pad = (int)align_up(n_tok, 16);
    const int64_t total = (int64_t)n_pad * (K / 8);
    launch_pdl(qgemm_swiglu_x_kernel, dim3((unsigned)std::min<int64_t>((total + 255) / 256, 148 * 8)), dim3(256), 0, s, gate, up, n_tok, K, (__half *)x_tiles, n_pad);
    g_launch_count++;
    B2_CUDA(cudaGetLastError());
}

// ------------------------------------------------------------------------------------------------ PTX helpers
__device__ __forceinline__ uint32_t qg_s32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void qg_mbar_init(uint64_t *bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(qg_s32(bar)), "r"(count));
}
__device__ __forceinline__ void qg_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(qg_s32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void qg_arrive(uint64_t *bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(qg_s32(bar)) : "memory");
}
__device__ __forceinline__ void qg_wait(uint64_t *bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "QG_WAIT:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra QG_DONE;\n"
        "bra QG_WAIT;\n"
        "QG_DONE:\n"
        "}\n" ::"r"(qg_s32(bar)),
        "r"(parity)
        : "memory");
}
__device__ __forceinline__ void qg_bulk_g2s(void *dst, const void *src, uint32_t bytes, uint64_t *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(qg_s32(dst)), "l"(src),
                 "r"(bytes), "r"(qg_s32(bar))
                 : "memory");
}
// one lane of a converged warp (the operands of tcgen05.mma / cp.async.bulk live in uniform registers: issued from a
// `lane == 0` branch the compiler cannot prove them warp-uniform and wraps every instruction in an election loop)
__device__ __forceinline__ bool qg_elect() {
    uint32_t e;
    asm volatile("{\n.reg .pred p;\nelect.sync _|p, 0xffffffff;\nselp.u32 %0, 1, 0, p;\n}\n" : "=r"(e));
    return e != 0;
}
__device__ __forceinline__ uint64_t qg_desc(uint32_t saddr, uint32_t lbo, uint32_t sbo) { // K-major, no swizzle
    return (uint64_t)((saddr & 0x3FFFF) >> 4) | ((uint64_t)(lbo >> 4) << 16) | ((uint64_t)(sbo >> 4) << 32) | (1ull << 46);
}
// D[tmem] (+)= A[tmem] . B[smem]
__device__ __forceinline__ void qg_mma_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n"
        ".reg .pred p, e;\n"
        "elect.sync _|e, 0xffffffff;\n"
        "setp.ne.b32 p, %4, 0;\n"
        "@e tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n"
        "}\n" ::"r"(tmem_d),
        "r"(tmem_a), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}
__device__ __forceinline__ void qg_commit(uint64_t *bar) { // called by a converged warp
    asm volatile("{\n.reg .pred e;\nelect.sync _|e, 0xffffffff;\n@e tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n}\n" ::"r"(qg_s32(bar))
                 : "memory");
}
__device__ __forceinline__ void qg_tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {
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
__device__ __forceinline__ void qg_tmem_st16(uint32_t taddr, const uint32_t (&r)[16]) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};" ::"r"(taddr), "r"(r[0]),
                 "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]), "r"(r[10]), "r"(r[11]),
                 "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15])
                 : "memory");
    asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void qg_tmem_st32(uint32_t taddr, const uint32_t (&r)[32]) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
        "{%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31,%32};" ::"r"(taddr),
        "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]), "r"(r[10]),
        "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]), "r"(r[16]), "r"(r[17]), "r"(r[18]), "r"(r[19]), "r"(r[20]),
        "r"(r[21]), "r"(r[22]), "r"(r[23]), "r"(r[24]), "r"(r[25]), "r"(r[26]), "r"(r[27]), "r"(r[28]), "r"(r[29]), "r"(r[30]),
        "r"(r[31])
        : "memory");
    asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
}

// ------------------------------------------------------------------------------------------------ dequantisation
__device__ __forceinline__ uint32_t qg_h2u(__half2 h) { return *(const uint32_t *)&h; }
__device__ __forceinline__ __half2 qg_u2h(uint32_t u) { return *(const __half2 *)&u; }
// Operand order: inside every group of four consecutive k the operand positions hold elements (0, 2, 1, 3) -- what the
// mask-and-or conversion of a 32-bit word of nibbles produces without a byte permute ((w & 0x000F000F) | 0x64006400 is
// the half2 (1024 + nibble of byte 0, 1024 + nibble of byte 2)).  A dot product does not care about the order of k as
// long as both operands agree: qg_store8 writes the activations in the same order.
// bytes (b0, b2) / (b1, b3) of v -> the half2 (1024 + b, 1024 + b'): 0x6400 | b is the fp16 1024 + b for b < 1024
__device__ __forceinline__ __half2 qg_pair02(uint32_t v) { return qg_u2h(__byte_perm(v, 0x64646464u, 0x4240)); }
__device__ __forceinline__ __half2 qg_pair13(uint32_t v) { return qg_u2h(__byte_perm(v, 0x64646464u, 0x4341)); }

// scale / min of sub-block sb from the 12 packed bytes (ggml get_scale_min_k4), the bytes in three words
__device__ __forceinline__ void qg_scale_min(uint32_t s0, uint32_t s1, uint32_t s2, int sb, int &sc, int &m) {
    if (sb < 4) {
        sc = (s0 >> (8 * sb)) & 63;
        m = (s1 >> (8 * sb)) & 63;
    } else {
        const int b = sb - 4;
        sc = ((s2 >> (8 * b)) & 0xF) | (((s0 >> (8 * b + 6)) & 3) << 4);
        m = ((s2 >> (8 * b + 4)) & 0xF) | (((s1 >> (8 * b + 6)) & 3) << 4);
    }
}

// one sub-block (32 consecutive k) of one Q4_K / Q5_K row -> 16 half2 registers.  q is exact in fp16; d * sc and
// dmin * m are rounded to fp16 once; value = fma(q, d*sc, -dmin*m) rounds once.
template <bool Q5>
__device__ __forceinline__ void qg_dq_k45(const uint8_t *blk, const uint4 &hdr, int sb, uint32_t (&r)[16]) {
    const __half2 dd = qg_u2h(hdr.x);
    int sc, m;
    qg_scale_min(hdr.y, hdr.z, hdr.w, sb, sc, m);
    const __half2 d1 = __float2half2_rn(__low2float(dd) * (float)sc), nm = __float2half2_rn(-(__high2float(dd) * (float)m));
    const uint8_t *qs = blk + (Q5 ? 48 : 16) + 32 * (sb >> 1);
    const uint4 q0 = *(const uint4 *)qs, q1 = *(const uint4 *)(qs + 16);
    const uint32_t w[8] = {q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, q1.z, q1.w};
    uint32_t hb[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    if (Q5) { // bit sb of qh[l] is the fifth bit of element l of sub-block sb
        const uint4 h0 = *(const uint4 *)(blk + 16), h1 = *(const uint4 *)(blk + 32);
        hb[0] = h0.x; hb[1] = h0.y; hb[2] = h0.z; hb[3] = h0.w; hb[4] = h1.x; hb[5] = h1.y; hb[6] = h1.z; hb[7] = h1.w;
    }
    if (!(sb & 1)) { // low nibbles: 0x6400 | q = 1024 + q
        const __half2 k1024 = __float2half2_rn(1024.f);
#pragma unroll
        for (int i = 0; i < 8; i++) {
            uint32_t x02 = (w[i] & 0x000F000Fu) | 0x64006400u, x13 = ((w[i] >> 8) & 0x000F000Fu) | 0x64006400u;
            if (Q5) {
                const uint32_t t = (hb[i] >> sb) & 0x01010101u;
                x02 |= (t << 4) & 0x00100010u;
                x13 |= (t >> 4) & 0x00100010u;
            }
            r[2 * i] = qg_h2u(__hfma2(__hsub2(qg_u2h(x02), k1024), d1, nm));
            r[2 * i + 1] = qg_h2u(__hfma2(__hsub2(qg_u2h(x13), k1024), d1, nm));
        }
    } else { // high nibbles, unshifted: 0x6400 | (q << 4) = 1024 + 16 q, and fma(., 1/16, -64) = q exactly
        const __half2 k16th = __float2half2_rn(0.0625f), km64 = __float2half2_rn(-64.f);
#pragma unroll
        for (int i = 0; i < 8; i++) {
            uint32_t x02 = (w[i] & 0x00F000F0u) | 0x64006400u, x13 = ((w[i] >> 8) & 0x00F000F0u) | 0x64006400u;
            if (Q5) { // + 16 * 16 for the fifth bit
                const uint32_t t = (hb[i] >> sb) & 0x01010101u;
                x02 |= (t << 8) & 0x01000100u;
                x13 |= t & 0x01000100u;
            }
            r[2 * i] = qg_h2u(__hfma2(__hfma2(qg_u2h(x02), k16th, km64), d1, nm));
            r[2 * i + 1] = qg_h2u(__hfma2(__hfma2(qg_u2h(x13), k16th, km64), d1, nm));
        }
    }
}

// Q6_K: 32-element group c (0..3) of half n of the super-block: ql[64 n + 32 (c & 1) + l] nibble (c >> 1), qh[32 n + l] bits 2c, 2c+1
__device__ __forceinline__ void qg_dq_q6k(const uint8_t *blk, float d, int n, int c, uint32_t (&r)[16]) {
    const uint4 *qlp = (const uint4 *)(blk + 64 * n + 32 * (c & 1));
    const uint4 *qhp = (const uint4 *)(blk + 128 + 32 * n);
    const uint4 l0 = qlp[0], l1 = qlp[1], h0 = qhp[0], h1 = qhp[1];
    const uint32_t ql[8] = {l0.x, l0.y, l0.z, l0.w, l1.x, l1.y, l1.z, l1.w};
    const uint32_t qh[8] = {h0.x, h0.y, h0.z, h0.w, h1.x, h1.y, h1.z, h1.w};
    const uint32_t scw = *(const uint16_t *)(blk + 192 + 8 * n + 2 * c); // the two int8 scales of these 32 elements (16 each)
    const __half2 ds0 = __float2half2_rn(d * (float)(int8_t)(scw & 0xFF)), ds1 = __float2half2_rn(d * (float)(int8_t)(scw >> 8));
    const __half2 k1056 = __float2half2_rn(1056.f); // 1024 (the bit trick) + 32 (Q6_K's offset)
#pragma unroll
    for (int i = 0; i < 8; i++) {
        const uint32_t lo4 = ((c >> 1) ? (ql[i] >> 4) : ql[i]) & 0x0F0F0F0Fu;
        const uint32_t hsh = c == 0 ? (qh[i] << 4) : c == 1 ? (qh[i] << 2) : c == 2 ? qh[i] : (qh[i] >> 2);
        const uint32_t v4 = lo4 | (hsh & 0x30303030u);
        const __half2 ds = i < 4 ? ds0 : ds1;
        r[2 * i] = qg_h2u(__hmul2(__hsub2(qg_pair02(v4), k1056), ds));
        r[2 * i + 1] = qg_h2u(__hmul2(__hsub2(qg_pair13(v4), k1056), ds));
    }
}

// Q8_0: block b (0..7) of the row's 256-k group: [qs 256][d x 8]
__device__ __forceinline__ void qg_dq_q80(const uint8_t *blk, int b, uint32_t (&r)[16]) {
    const __half2 k1152 = __float2half2_rn(1152.f); // 1024 + 128: the bytes are biased by 0x80 to make them unsigned
    const __half2 d2 = __half2half2(*(const __half *)(blk + 256 + 2 * b));
    const uint4 a0 = *(const uint4 *)(blk + 32 * b), a1 = *(const uint4 *)(blk + 32 * b + 16);
    const uint32_t w[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
#pragma unroll
    for (int i = 0; i < 8; i++) {
        const uint32_t u = w[i] ^ 0x80808080u;
        r[2 * i] = qg_h2u(__hmul2(__hsub2(qg_pair02(u), k1152), d2));
        r[2 * i + 1] = qg_h2u(__hmul2(__hsub2(qg_pair13(u), k1152), d2));
    }
}

// ------------------------------------------------------------------------------------------------ the kernel
#ifdef QG_DBG // profiling build: cycles each role spends waiting / working, summed over a launch (scratch/qgemm_dbg.py)
static __device__ long long g_qg_dbg[148][16];
#define QG_T0() long long qg_t0 = clock64()
#define QG_ACC(slot) do { const long long qg_t1 = clock64(); qg_acc[slot] += qg_t1 - qg_t0; qg_t0 = qg_t1; } while (0)
#else
#define QG_T0()
#define QG_ACC(slot)
#endif
int qgemm_debug_read(long long *host, int n) {
#ifdef QG_DBG
    if (host && n >= 148 * 16) cudaMemcpyFromSymbol(host, g_qg_dbg, sizeof(long long) * 148 * 16);
    return 148 * 16;
#else
    (void)host; (void)n;
    return 0;
#endif
}

struct QGemmJobDev {
    const uint8_t *q_tiles; // [tiles][K/256][chunk]
    float *out;
    int64_t out_stride;
    int N, nt0, type, chunk_bytes; // nt0: first tile of this matrix among the launch's tiles
};
struct QGemmArgs {
    const __half *x_tiles;
    QGemmJobDev job[3];
    int n_jobs;
    const float *residual;
    int64_t res_stride;
    int n_tok, n_pad, nkb, T, units;  // nkb = K / 256; units = panels * nkb
    int raw_slots, raw_slot_bytes;
    float *partial;                   // [2 * CTA + (segment starts at the panel's first k-block)][T][n_pad][128]: a CTA's range has
                                      // at most two partial segments, its first (starts inside a panel) and its last (ends inside one)
    unsigned *counters;               // [tile] warp arrivals (4 per segment), zero between launches (the last arriver re-zeroes its counter)
};

__device__ __forceinline__ const QGemmJobDev &qg_job_of(const QGemmArgs &g, int tile) {
    return g.job[(g.n_jobs > 2 && tile >= g.job[2].nt0) ? 2 : (g.n_jobs > 1 && tile >= g.job[1].nt0) ? 1 : 0];
}
// the CTA whose unit range [c * U / G, (c + 1) * U / G) holds unit x
__host__ __device__ __forceinline__ int qg_cta_of(int64_t x, int64_t U, int64_t G) { return (int)(((x + 1) * G + U - 1) / U - 1); }

__global__ void __launch_bounds__(QG_THREADS, 1)
qgemm_kernel(const __grid_constant__ QGemmArgs g) {
    extern __shared__ __align__(1024) uint8_t qsm[];
    B2_PDL_TRIGGER();
#ifdef QG_DBG
    const long long qg_kstart = clock64();
#endif
    const int b_bytes = g.n_pad * 512; // one activation stage: n_pad tokens x 256 k x 2 B
    uint8_t *raw = qsm;
    uint8_t *bst = qsm + QG_RAW_RING;
    uint64_t *bars = (uint64_t *)(bst + QG_B_STAGES * 64 * 512);
    uint64_t *raw_full = bars, *raw_empty = raw_full + QG_MAX_RAW;
    uint64_t *b_full = raw_empty + QG_MAX_RAW, *b_empty = b_full + QG_B_STAGES;
    uint64_t *a_full = b_empty + QG_B_STAGES, *a_empty = a_full + QG_A_STAGES;
    uint64_t *d_full = a_empty + QG_A_STAGES, *d_empty = d_full + 2;
    uint32_t *tmem_slot = (uint32_t *)(d_empty + 2);
    int *n_fix = (int *)(tmem_slot + 1);   // tiles this CTA has to sum up after its main loop (it held their last segment)
    int4 *fix = (int4 *)(tmem_slot + 4);   // [<= 4 per CTA range: two partial segments of T tiles]  {panel, t, first CTA, segments}
    const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31; // (the shuffle tells the compiler it is warp-uniform)

    if (threadIdx.x == 0) {
        *n_fix = 0;
        for (int i = 0; i < QG_MAX_RAW; i++) { qg_mbar_init(raw_full + i, 1); qg_mbar_init(raw_empty + i, QG_DQ_WARPS); }
        for (int i = 0; i < QG_B_STAGES; i++) { qg_mbar_init(b_full + i, 1); qg_mbar_init(b_empty + i, 1); }
        for (int i = 0; i < QG_A_STAGES; i++) { qg_mbar_init(a_full + i, QG_DQ_WARPS); qg_mbar_init(a_empty + i, 1); }
        for (int i = 0; i < 2; i++) { qg_mbar_init(d_full + i, 1); qg_mbar_init(d_empty + i, 4); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 2) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(qg_s32(tmem_slot)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = __shfl_sync(0xffffffffu, *tmem_slot, 0);
    // (programmatic dependent launch: the roles that touch what earlier kernels of the stream wrote or still read -- the
    // activation producer, the epilogue warps, the fix-up -- wait for the previous kernel themselves; the weight producer,
    // the dequant and MMA warps start at once)
#ifdef QG_DBG
    if (threadIdx.x == 0) g_qg_dbg[blockIdx.x][13] = clock64() - qg_kstart; // prologue
#endif
    const int64_t U = g.units, G = gridDim.x;
    const int u0 = (int)((int64_t)blockIdx.x * U / G), u1 = (int)(((int64_t)blockIdx.x + 1) * U / G);
    const int nkb = g.nkb, T = g.T;

    if (warp == 0) {
        { // ===== weight producer (the whole warp walks the schedule; one elected lane issues the copies).  The weights are
          // immutable, so this warp does not wait for the previous kernel of the stream: the ring fills while that one drains.
            uint32_t rit = 0;
#ifdef QG_DBG
            long long qg_acc[4] = {0, 0, 0, 0};
#endif
            QG_T0();
            for (int u = u0; u < u1;) {
                const int p = u / nkb, kb0 = u - p * nkb, kb1 = min(nkb, kb0 + (u1 - u));
                for (int kb = kb0; kb < kb1; kb++) {
                    for (int t = 0; t < T; t++, rit++) {
                        const int tile = p * T + t;
                        const QGemmJobDev &J = qg_job_of(g, tile);
                        const int rs = rit % g.raw_slots;
                        QG_ACC(2);
                        qg_wait(raw_empty + rs, ((rit / g.raw_slots) & 1) ^ 1);
                        QG_ACC(1);
                        if (qg_elect()) {
                            qg_expect_tx(raw_full + rs, J.chunk_bytes);
                            qg_bulk_g2s(raw + (size_t)rs * g.raw_slot_bytes, J.q_tiles + ((int64_t)(tile - J.nt0) * nkb + kb) * J.chunk_bytes,
                                        J.chunk_bytes, raw_full + rs);
                        }
                    }
                }
                u += kb1 - kb0;
            }
#ifdef QG_DBG
            for (int i = 0; i < 3; i++) g_qg_dbg[blockIdx.x][10 + i] = qg_acc[i];
#endif
        }
    } else if (warp == 3) {
        { // ===== activation producer: one 256-k stage of all tokens per unit (written by the previous kernel: wait for it)
            B2_PDL_WAIT();
            uint32_t bit = 0;
            for (int u = u0; u < u1;) {
                const int p = u / nkb, kb0 = u - p * nkb, kb1 = min(nkb, kb0 + (u1 - u));
                for (int kb = kb0; kb < kb1; kb++, bit++) {
                    const int bs = bit % QG_B_STAGES;
                    qg_wait(b_empty + bs, ((bit / QG_B_STAGES) & 1) ^ 1);
                    if (qg_elect()) {
                        qg_expect_tx(b_full + bs, b_bytes);
                        qg_bulk_g2s(bst + (size_t)bs * b_bytes, g.x_tiles + (int64_t)kb * g.n_pad * 256, b_bytes, b_full + bs);
                    }
                }
                u += kb1 - kb0;
            }
        }
    } else if (warp == 1) {
        { // ===== MMA issuer (whole warp, elected lane)
            // kind::f16: D f32 (bit 4), A = B = f16 (format 0), both K-major; N = n_pad, M = 128
            const uint32_t idesc = (1u << 4) | ((uint32_t)(g.n_pad >> 3) << 17) | ((uint32_t)(QG_ROWS >> 4) << 24);
            uint32_t ait = 0, bit = 0, sit = 0;
#ifdef QG_DBG
            long long qg_acc[4] = {0, 0, 0, 0};
#endif
            QG_T0();
            for (int u = u0; u < u1; sit++) {
                const int p = u / nkb, kb0 = u - p * nkb, kb1 = min(nkb, kb0 + (u1 - u));
                const int buf = sit & 1;
                qg_wait(d_empty + buf, ((sit >> 1) & 1) ^ 1);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                QG_ACC(0);
                for (int kb = kb0; kb < kb1; kb++, bit++) {
                    const int bs = bit % QG_B_STAGES;
                    qg_wait(b_full + bs, (bit / QG_B_STAGES) & 1);
                    QG_ACC(1);
                    const uint32_t sb = qg_s32(bst + (size_t)bs * b_bytes);
                    for (int t = 0; t < T; t++) {
                        const uint32_t d_tmem = tmem + QG_D_COL + (buf * 2 + t) * 64;
                        for (int hh = 0; hh < 2; hh++, ait++) {
                            const int as = ait % QG_A_STAGES;
                            qg_wait(a_full + as, (ait / QG_A_STAGES) & 1);
                            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                            QG_ACC(2);
#pragma unroll
                            for (int i = 0; i < 8; i++) // one MMA = K 16 = 8 TMEM columns of A, two core matrices (256 B) of B
                                qg_mma_ts(d_tmem, tmem + QG_A_COL + as * 64 + i * 8, qg_desc(sb + (hh * 8 + i) * 256, 128, 4096), idesc,
                                          (uint32_t)((kb - kb0) | hh | i) != 0);
                            qg_commit(a_empty + as);
                            __syncwarp();
                            QG_ACC(3);
                        }
                    }
                    qg_commit(b_empty + bs);
                }
                qg_commit(d_full + buf);
                u += kb1 - kb0;
            }
#ifdef QG_DBG
            for (int i = 0; i < 4; i++) g_qg_dbg[blockIdx.x][i] = qg_acc[i];
#endif
        }
    } else if (warp >= 8) { // ===== dequant: raw blocks (smem) -> fp16 A operand (TMEM)
        const int dw = warp - 8, q = dw & 3, ksel = dw >> 2; // lane quarter; which 32 of the 128 k of an operand stage
        const int row = q * 32 + lane;
        uint32_t rit = 0, ait = 0;
#ifdef QG_DBG
        long long qg_acc[4] = {0, 0, 0, 0};
        const long long qg_start = clock64();
#endif
        QG_T0();
        for (int u = u0; u < u1;) {
            const int p = u / nkb, kb0 = u - p * nkb, kb1 = min(nkb, kb0 + (u1 - u));
            for (int kb = kb0; kb < kb1; kb++) {
                for (int t = 0; t < T; t++, rit++) {
                    const QGemmJobDev &J = qg_job_of(g, p * T + t);
                    const int rs = rit % g.raw_slots;
                    qg_wait(raw_full + rs, (rit / g.raw_slots) & 1);
                    QG_ACC(0);
                    const uint8_t *chunk = raw + (size_t)rs * g.raw_slot_bytes;
                    const int type = J.type;
                    uint4 hdr = make_uint4(0, 0, 0, 0);
                    float d6 = 0.f;
                    const uint8_t *blk;
                    if (type == B2_T_Q4_K) { blk = chunk + row * 144; hdr = *(const uint4 *)blk; }
                    else if (type == B2_T_Q5_K) { blk = chunk + row * 176; hdr = *(const uint4 *)blk; }
                    else if (type == B2_T_Q6_K) { blk = chunk + row * 208; d6 = __half2float(*(const __half *)(chunk + QG_ROWS * 208 + 2 * row)); }
                    else blk = chunk + row * 272;
                    for (int hh = 0; hh < 2; hh++, ait++) {
                        uint32_t r[16];
                        if (type == B2_T_Q4_K) qg_dq_k45<false>(blk, hdr, 4 * hh + ksel, r);
                        else if (type == B2_T_Q5_K) qg_dq_k45<true>(blk, hdr, 4 * hh + ksel, r);
                        else if (type == B2_T_Q6_K) qg_dq_q6k(blk, d6, hh, ksel, r);
                        else qg_dq_q80(blk, 4 * hh + ksel, r);
                        const int as = ait % QG_A_STAGES;
                        QG_ACC(1);
                        qg_wait(a_empty + as, ((ait / QG_A_STAGES) & 1) ^ 1);
                        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                        QG_ACC(2);
                        qg_tmem_st16(tmem + ((uint32_t)(q * 32) << 16) + QG_A_COL + as * 64 + ksel * 16, r);
                        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
                        __syncwarp();
                        if (lane == 0) qg_arrive(a_full + as);
                        QG_ACC(3);
                    }
                    __syncwarp();
                    if (lane == 0) qg_arrive(raw_empty + rs); // every lane's loads of this slot have completed (their values were used)
                }
            }
            u += kb1 - kb0;
        }
#ifdef QG_DBG
        if (dw == 0 && lane == 0) {
            for (int i = 0; i < 4; i++) g_qg_dbg[blockIdx.x][4 + i] = qg_acc[i];
            g_qg_dbg[blockIdx.x][8] = clock64() - qg_start;
            g_qg_dbg[blockIdx.x][9] = rit;
        }
#endif
    } else if (warp >= 4) { // ===== epilogue: TMEM -> registers -> global (direct, or partial + last-arriver reduce)
        const int q = warp & 3;
        uint32_t sit = 0;
        B2_PDL_WAIT();
        for (int u = u0; u < u1; sit++) {
            const int p = u / nkb, kb0 = u - p * nkb, kb1 = min(nkb, kb0 + (u1 - u));
            const int buf = sit & 1;
            const bool full = kb0 == 0 && kb1 == nkb;
            int c_first = 0, nseg = 1;
            if (!full) {
                c_first = qg_cta_of((int64_t)p * nkb, U, G);
                nseg = qg_cta_of((int64_t)(p + 1) * nkb - 1, U, G) - c_first + 1;
            }
            const int64_t slot_floats = (int64_t)T * g.n_pad * QG_ROWS;
            qg_wait(d_full + buf, (sit >> 1) & 1);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            for (int t = 0; t < T; t++) {
                const int tile = p * T + t;
                const QGemmJobDev &J = qg_job_of(g, tile);
                const int row = (tile - J.nt0) * QG_ROWS + q * 32 + lane;
                const bool valid = row < J.N;
                const int64_t in_slot = (int64_t)t * g.n_pad * QG_ROWS + q * 32 + lane;
                float *part = g.partial + (2 * (int64_t)blockIdx.x + (kb0 == 0)) * slot_floats + in_slot;
                for (int c0 = 0; c0 < g.n_pad; c0 += 32) {
                    uint32_t r[32];
                    qg_tmem_ld32(tmem + ((uint32_t)(q * 32) << 16) + QG_D_COL + (buf * 2 + t) * 64 + c0, r);
                    if (t == T - 1 && c0 + 32 >= g.n_pad) { // the accumulators are in registers: hand the buffer back to the MMA warp
                        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
                        __syncwarp();
                        if (lane == 0) qg_arrive(d_empty + buf);
                    }
                    if (full) {
                        if (valid) {
#pragma unroll
                            for (int j = 0; j < 32; j++) {
                                const int tok = c0 + j;
                                if (tok < g.n_tok) {
                                    float v = __uint_as_float(r[j]);
                                    if (g.residual) v += g.residual[(int64_t)tok * g.res_stride + row];
                                    J.out[(int64_t)tok * J.out_stride + row] = v;
                                }
                            }
                        }
                    } else {
#pragma unroll
                        for (int j = 0; j < 32; j++)
                            if (c0 + j < g.n_tok) __stcg(part + (int64_t)(c0 + j) * QG_ROWS, __uint_as_float(r[j]));
                    }
                }
            }
            if (!full) { // one fence for the segment's tiles, then their counters
                __threadfence();
                __syncwarp();
                if (lane < T) {
                    const int tile = p * T + lane;
                    const unsigned old = atomicAdd(g.counters + tile, 1u);
                    if (old == 4u * (unsigned)nseg - 1) { // the last of the tile's 4 x nseg warp arrivals: this CTA adds the partials
                        __threadfence();
                        g.counters[tile] = 0;
                        const int e = atomicAdd(n_fix, 1);
                        fix[e] = make_int4(p, lane, c_first, nseg);
                    }
                }
            }
            u += kb1 - kb0;
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 2) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem) : "memory");
#ifdef QG_DBG
    if (threadIdx.x == 0) g_qg_dbg[blockIdx.x][14] = clock64() - qg_kstart; // main loops done (all roles)
#endif
    // ===== fix-up: the tiles whose last partial segment arrived in this CTA.  All 24 warps: a warp takes whole tokens (lane =
    // four consecutive rows, 16-byte loads), three tokens each, four segments' loads in flight; partials are added in
    // segment order, so the sum does not depend on which CTA happens to be last.
    const int nf = *n_fix;
    B2_PDL_WAIT();
    for (int e = 0; e < nf; e++) {
        const int4 f = fix[e];
        const int p = f.x, t = f.y, c_first = f.z, nseg = f.w;
        const int tile = p * T + t;
        const QGemmJobDev &J = qg_job_of(g, tile);
        const int r = lane * 4, row = (tile - J.nt0) * QG_ROWS + r;
        const int64_t slot_floats = (int64_t)T * g.n_pad * QG_ROWS;
        const float *p0 = g.partial + (2 * (int64_t)c_first + 1) * slot_floats + (int64_t)t * g.n_pad * QG_ROWS + r; // segment 0: starts at the panel's first k-block
        const float *p1 = g.partial + (2 * (int64_t)c_first) * slot_floats + (int64_t)t * g.n_pad * QG_ROWS + r;     // segment s: + 2 s slots
        constexpr int NW = QG_THREADS / 32, NJ = (QGEMM_MAX_TOKENS + NW - 1) / NW; // warps; tokens per warp
        const int w = threadIdx.x >> 5;
        float4 acc[NJ];
#pragma unroll
        for (int j = 0; j < NJ; j++) {
            const int tok = w + NW * j;
            acc[j] = tok < g.n_pad ? __ldcg((const float4 *)(p0 + (int64_t)tok * QG_ROWS)) : make_float4(0.f, 0.f, 0.f, 0.f);
        }
        for (int s0 = 1; s0 < nseg; s0 += 4) {
            float4 v[4][NJ];
#pragma unroll
            for (int ss = 0; ss < 4; ss++)
#pragma unroll
                for (int j = 0; j < NJ; j++) {
                    const int tok = w + NW * j;
                    v[ss][j] = (s0 + ss < nseg && tok < g.n_pad) ? __ldcg((const float4 *)(p1 + 2 * (s0 + ss) * slot_floats + (int64_t)tok * QG_ROWS))
                                                                 : make_float4(0.f, 0.f, 0.f, 0.f);
                }
#pragma unroll
            for (int ss = 0; ss < 4; ss++)
#pragma unroll
                for (int j = 0; j < NJ; j++) { // (+ 0 for a segment past the last: the same value in every run)
                    acc[j].x += v[ss][j].x; acc[j].y += v[ss][j].y; acc[j].z += v[ss][j].z; acc[j].w += v[ss][j].w;
                }
        }
        const bool vec = row + 3 < J.N && (J.out_stride & 3) == 0 && (((uintptr_t)J.out) & 15) == 0 &&
                         (!g.residual || ((g.res_stride & 3) == 0 && (((uintptr_t)g.residual) & 15) == 0));
#pragma unroll
        for (int j = 0; j < NJ; j++) {
            const int tok = w + NW * j;
            if (tok >= g.n_tok) continue;
            float *o = J.out + (int64_t)tok * J.out_stride + row;
            const float *res = g.residual ? g.residual + (int64_t)tok * g.res_stride + row : nullptr;
            if (vec) {
                float4 v = acc[j];
                if (res) {
                    const float4 rr = *(const float4 *)res;
                    v.x += rr.x; v.y += rr.y; v.z += rr.z; v.w += rr.w;
                }
                *(float4 *)o = v;
            } else {
                const float a4[4] = {acc[j].x, acc[j].y, acc[j].z, acc[j].w};
                for (int i = 0; i < 4; i++)
                    if (row + i < J.N) o[i] = a4[i] + (res ? res[i] : 0.f);
            }
        }
    }
#ifdef QG_DBG
    __syncthreads();
    if (threadIdx.x == 0) g_qg_dbg[blockIdx.x][15] = clock64() - qg_kstart; // with the fix-up
#endif
}

// ------------------------------------------------------------------------------------------------ host side
constexpr size_t QG_SMEM = (size_t)QG_RAW_RING + (size_t)QG_B_STAGES * 64 * 512 + 512;

// the plan of one launch (host arithmetic only; tests/test_capi_cpu.py replays it through b2_qgemm_plan_check)
void qgemm_plan(int n_tiles, int nkb, int n_sm, bool allow_pairs, QGemmPlan *pl) {
    int T = 1;
    // pairs of tiles share the activation stage; only when a CTA's range then still spans whole panels (few split tiles)
    if (allow_pairs && n_tiles % 2 == 0 && (int64_t)(n_tiles / 2) * nkb >= (int64_t)nkb * n_sm) T = 2;
    if (const char *e = getenv("B200LLM_QG_T")) { // A/B measurements: 1 / 2 force the panel width where it divides
        const int f = atoi(e);
        if (f == 1 || (f == 2 && n_tiles % 2 == 0 && n_tiles / 2 * (int64_t)nkb >= n_sm)) T = f;
    }
    pl->T = T;
    pl->panels = n_tiles / T;
    pl->units = (int64_t)pl->panels * nkb;
    pl->grid = (int)std::min<int64_t>(n_sm, pl->units);
}

// host-only replay of the kernel's work split (the CPU unit test of the planner): every (panel, kb) unit belongs to exactly
// one CTA range; the segment arithmetic the epilogue uses (first CTA of a panel, number of segments) matches the ranges; a
// CTA never needs more than its two partial slots.  Throws b2::Error on an inconsistency; returns the number of split tiles.
int64_t qgemm_plan_check(int n_tiles, int nkb, int n_sm) {
    QGemmPlan pl;
    qgemm_plan(n_tiles, nkb, n_sm, true, &pl);
    B2_CHECK(pl.T >= 1 && pl.T <= 2 && pl.panels * pl.T == n_tiles && pl.units == (int64_t)pl.panels * nkb, B2_EINVAL, "qgemm plan: panels");
    B2_CHECK(pl.grid >= 1 && pl.grid <= n_sm && pl.grid <= pl.units, B2_EINVAL, "qgemm plan: grid");
    const int64_t U = pl.units, G = pl.grid;
    std::vector<int> seen_first(pl.panels, -1), seen_n(pl.panels, 0);
    int64_t covered = 0, split = 0;
    for (int c = 0; c < G; c++) {
        const int64_t u0 = c * U / G, u1 = (c + 1) * U / G;
        B2_CHECK(u1 > u0, B2_EINVAL, "qgemm plan: empty CTA range");
        int slots_used[2] = {0, 0};
        for (int64_t u = u0; u < u1;) {
            const int64_t p = u / nkb, kb0 = u - p * nkb, kb1 = std::min<int64_t>(nkb, kb0 + (u1 - u));
            const bool full = kb0 == 0 && kb1 == nkb;
            if (!full) {
                const int c_first = qg_cta_of(p * nkb, U, G), c_last = qg_cta_of((p + 1) * nkb - 1, U, G);
                B2_CHECK(c >= c_first && c <= c_last, B2_EINVAL, "qgemm plan: CTA outside its panel's segment range");
                // segment index c - c_first: 0 must be the one that starts at the panel's first k-block
                B2_CHECK((kb0 == 0) == (c == c_first), B2_EINVAL, "qgemm plan: segment 0 is not the panel's first");
                B2_CHECK(++slots_used[kb0 == 0] == 1, B2_EINVAL, "qgemm plan: a CTA needs a partial slot twice");
                if (seen_first[p] < 0) seen_first[p] = c_first;
                seen_n[p]++;
                if (c == c_last) B2_CHECK(seen_n[p] == c_last - c_first + 1 && kb1 == nkb, B2_EINVAL, "qgemm plan: segment count");
            }
            covered += kb1 - kb0;
            u += kb1 - kb0;
        }
    }
    B2_CHECK(covered == U, B2_EINVAL, "qgemm plan: units not covered exactly once");
    for (int p = 0; p < pl.panels; p++) split += seen_n[p] > 0 ? pl.T : 0;
    return split;
}

void launch_qgemm_multi(const QGemmJob *jobs, int n_jobs, const void *x_tiles, int n_tok, const float *residual, int64_t res_stride,
                        float *partial, size_t partial_bytes, unsigned *counters, size_t n_counters, int n_sm, cudaStream_t s) {
    B2_CHECK(n_jobs >= 1 && n_jobs <= 3, B2_EINVAL, "qgemm: 1 to 3 matrices per launch");
    B2_CHECK(!residual || n_jobs == 1, B2_EINVAL, "qgemm: the residual add needs a single matrix");
    B2_CHECK(n_tok >= 1 && n_tok <= QGEMM_MAX_TOKENS, B2_EINVAL, "qgemm: 1..64 tokens");
    const int64_t cols = jobs[0].W->cols;
    B2_CHECK(cols % 256 == 0, B2_EINVAL, "qgemm: K must be a multiple of 256");
    static PerDeviceOnce configured;
    configured.run([&] { B2_CUDA(cudaFuncSetAttribute(qgemm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)QG_SMEM)); });
    QGemmArgs g;
    memset(&g, 0, sizeof(g));
    g.x_tiles = (const __half *)x_tiles;
    g.n_jobs = n_jobs;
    int nt = 0, max_chunk = 0;
    for (int j = 0; j < n_jobs; j++) {
        const DTensor &W = *jobs[j].W;
        B2_CHECK(W.cols == cols && qgemm_type_supported(W.type), B2_EINVAL, "qgemm: matrices of one launch share K; Q4_K/Q5_K/Q6_K/Q8_0 only");
        g.job[j].q_tiles = (const uint8_t *)jobs[j].q_tiles;
        g.job[j].out = jobs[j].out; g.job[j].out_stride = jobs[j].out_stride;
        g.job[j].N = (int)W.rows; g.job[j].nt0 = nt; g.job[j].type = W.type; g.job[j].chunk_bytes = qg_chunk_bytes(W.type);
        max_chunk = std::max(max_chunk, g.job[j].chunk_bytes);
        nt += (int)(align_up(W.rows, QG_ROWS) / QG_ROWS);
    }
    g.residual = residual; g.res_stride = res_stride;
    g.n_tok = n_tok; g.n_pad = (int)align_up(n_tok, 16); g.nkb = (int)(cols / 256);
    QGemmPlan pl;
    qgemm_plan(nt, g.nkb, n_sm, true, &pl);
    g.T = pl.T; g.units = (int)pl.units;
    g.raw_slot_bytes = (int)align_up(max_chunk, 128);
    g.raw_slots = std::min(QG_MAX_RAW, QG_RAW_RING / g.raw_slot_bytes);
    g.partial = partial; g.counters = counters;
    B2_CHECK((size_t)2 * pl.grid * pl.T * g.n_pad * QG_ROWS * sizeof(float) <= partial_bytes, B2_EINVAL, "qgemm: partial-tile scratch too small");
    B2_CHECK((size_t)nt <= n_counters, B2_EINVAL, "qgemm: too few tile counters");
    launch_pdl(qgemm_kernel, dim3(pl.grid), dim3(QG_THREADS), QG_SMEM, s, g);
    g_launch_count++;
    B2_CUDA(cudaGetLastError());
}

} // namespace b2
