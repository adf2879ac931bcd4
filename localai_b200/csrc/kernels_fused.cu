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
#include "gemv_rows2.cuh"
#include "kernels.h"

namespace b2 {

constexpr int FW = 8;         // warps per CTA
constexpr int FUSED_MAX_NT = 4;

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
    if (a.xq_pre) { // rows quantised by the stand-alone kernels: copy, swizzling the int8 section (act_swz)
        const int per_row = (int)(a.xrb / 16);
        const int n_q8 = (a.fmt == ACT_Q8_K || a.fmt == ACT_Q8_0) ? (int)(K / 16) : 0;
        const int4 *src = (const int4 *)a.xq_pre;
        int4 *dst = (int4 *)xs;
        for (int i = threadIdx.x; i < per_row * NT; i += blockDim.x) {
            const int t = i / per_row, c = i - t * per_row;
            dst[t * per_row + (c < n_q8 ? act_swz(c) : c)] = src[i];
        }
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
                quant_block<true>(a.fmt, v, lane, row, K, b);
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
        rows2_dot<LdGlobal, 2, NT>(a.job[jb].type, r0, r1, K, 0, 1, xs, a.xrb, lane, acc);
#pragma unroll
        for (int t = 0; t < NT; t++) {
            float v0 = warp_sum(acc[0][t]), v1 = warp_sum(acc[1][t]);
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
                if (const float *bj = a.qkv_bias[jb]) { // ggml_add after the mat-mul, before the rotation
                    v0 = __fadd_rn(v0, bj[r_local]);
                    v1 = __fadd_rn(v1, bj[r_local + 1]);
                }
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
        static PerDeviceOnce cfg; /* once per device and instantiation, not per launch (the call takes a driver lock) */   \
        if (smem > 48 * 1024) cfg.run([&] { B2_CUDA(cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024)); }); \
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
