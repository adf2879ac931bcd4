// kernels_gemv.cu -- streaming dequant-GEMV for decode (and small token counts).
//
// out[t][r] = dot(W[r,:], xq[t,:]) where W is a quantised weight matrix streamed ONCE from HBM with 128-bit
// loads, and xq is the activation row already quantised the way llama.cpp's CPU path does it (Q8_K for
// K-quants, Q8_0 for Q8_0, f16 for F16).  The dot is integer (dp4a) inside each block and is scaled in fp32,
// i.e. the same arithmetic as ggml_vec_dot_q4_K_q8_K & co -- only the fp32 summation ORDER differs.
//
// Work split: one warp per output row; a Q4_K/Q5_K/Q6_K super-block (256 weights) is spread over 8 lanes so
// that the 8 lanes x 16 B they load are contiguous (4 blocks = 4 x 128 contiguous bytes per warp load).
// The activation rows live in shared memory.  HBM-bound: algorithmic bytes = rows x row_stride.
#include "kernels.h"
#include "gemv_rows.cuh"

namespace b2 {

constexpr int GEMV_WARPS = 8;
constexpr int GEMV_MAX_NT = 4;

template <int TYPE, int NT>
__global__ void __launch_bounds__(GEMV_WARPS * 32)
gemv_kernel(const uint8_t *__restrict__ W, int64_t rows, int64_t K, int64_t row_stride, const uint8_t *__restrict__ xq,
            int64_t xrb, float *__restrict__ out, int64_t out_stride, const float *__restrict__ residual, int64_t res_stride) {
    extern __shared__ __align__(16) uint8_t xs[];
    {   // stage the NT activation rows
        const int n16 = (int)(xrb * NT / 16);
        const int4 *src = (const int4 *)xq;
        int4 *dst = (int4 *)xs;
        for (int i = threadIdx.x; i < n16; i += blockDim.x) dst[i] = src[i];
    }
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int nb = (int)(K / 256);
    for (int64_t r = (int64_t)blockIdx.x * GEMV_WARPS + warp; r < rows; r += (int64_t)gridDim.x * GEMV_WARPS) {
        const uint8_t *row = W + r * row_stride;
        float acc[NT];
#pragma unroll
        for (int t = 0; t < NT; t++) acc[t] = 0.f;
        if (TYPE == B2_T_Q4_K) row_dot_q4_K<NT>(row, nb, K, xs, xrb, lane, acc);
        else if (TYPE == B2_T_Q5_K) row_dot_q5_K<NT>(row, nb, K, xs, xrb, lane, acc);
        else if (TYPE == B2_T_Q6_K) row_dot_q6_K<NT>(row, nb, K, xs, xrb, lane, acc);
        else if (TYPE == B2_T_Q8_0) row_dot_q8_0<NT>(row, K, xs, xrb, lane, acc);
        else if (TYPE == B2_T_F16) row_dot_f16<NT>(row, K, xs, xrb, lane, acc);
        else row_dot_f32<NT>(row, K, xs, xrb, lane, acc);
#pragma unroll
        for (int t = 0; t < NT; t++) {
            const float v = warp_sum(acc[t]);
            if (lane == 0) {
                float o = v;
                if (residual) o = v + residual[t * res_stride + r];
                out[t * out_stride + r] = o;
            }
        }
    }
}

template <int TYPE>
static void launch_gemv_typed(const DTensor &W, const uint8_t *xq, int64_t xrb, int n_tok, float *out, int64_t out_stride,
                              const float *residual, int64_t res_stride, cudaStream_t s) {
    // rows per CTA pass = 8; aim at ~4 CTAs per SM worth of work, at least 2 rows per warp
    int64_t ctas = (W.rows + GEMV_WARPS - 1) / GEMV_WARPS;
    const int64_t cap = 148 * 8;
    if (ctas > cap) ctas = cap;
    for (int t0 = 0; t0 < n_tok; t0 += GEMV_MAX_NT) {
        const int nt = std::min(GEMV_MAX_NT, n_tok - t0);
        const size_t smem = (size_t)xrb * nt;
        const uint8_t *xp = xq + (size_t)t0 * xrb;
        float *op = out + (size_t)t0 * out_stride;
        const float *rp = residual ? residual + (size_t)t0 * res_stride : nullptr;
#define B2_GEMV_CASE(N)                                                                                             \
    case N: {                                                                                                       \
        auto kfn = gemv_kernel<TYPE, N>;                                                                            \
        static PerDeviceOnce cfg; /* once per device and instantiation, not per launch (the call takes a driver lock) */   \
        if (smem > 48 * 1024) cfg.run([&] { B2_CUDA(cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024)); }); \
        kfn<<<(int)ctas, GEMV_WARPS * 32, smem, s>>>(W.data, W.rows, W.cols, W.row_stride, xp, xrb, op, out_stride, rp, \
                                                     res_stride);                                                  \
    } break;
        switch (nt) {
            B2_GEMV_CASE(1)
            B2_GEMV_CASE(2)
            B2_GEMV_CASE(3)
            B2_GEMV_CASE(4)
        }
#undef B2_GEMV_CASE
        g_launch_count++;
        B2_CUDA(cudaGetLastError());
    }
}

void launch_gemv(const DTensor &W, const uint8_t *xq, ActFmt f, int n_tok, float *out, int64_t out_stride,
                 const float *residual, int64_t res_stride, cudaStream_t s) {
    B2_CHECK(f == act_format_for(W.type), B2_EINVAL, "gemv: activation format does not match weight type");
    const int64_t xrb = act_row_bytes(f, W.cols);
    B2_CHECK(xrb * std::min(n_tok, GEMV_MAX_NT) <= 200 * 1024, B2_EUNSUPPORTED, "gemv: activation rows exceed shared memory");
    switch (W.type) {
        case B2_T_Q4_K: launch_gemv_typed<B2_T_Q4_K>(W, xq, xrb, n_tok, out, out_stride, residual, res_stride, s); break;
        case B2_T_Q5_K: launch_gemv_typed<B2_T_Q5_K>(W, xq, xrb, n_tok, out, out_stride, residual, res_stride, s); break;
        case B2_T_Q6_K: launch_gemv_typed<B2_T_Q6_K>(W, xq, xrb, n_tok, out, out_stride, residual, res_stride, s); break;
        case B2_T_Q8_0: launch_gemv_typed<B2_T_Q8_0>(W, xq, xrb, n_tok, out, out_stride, residual, res_stride, s); break;
        case B2_T_F16: launch_gemv_typed<B2_T_F16>(W, xq, xrb, n_tok, out, out_stride, residual, res_stride, s); break;
        case B2_T_F32: launch_gemv_typed<B2_T_F32>(W, xq, xrb, n_tok, out, out_stride, residual, res_stride, s); break;
        default: throw Error(B2_EUNSUPPORTED, "gemv: unsupported weight type");
    }
}

} // namespace b2
