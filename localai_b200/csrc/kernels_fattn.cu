// kernels_fattn.cu -- prefill attention: flash-style tiled kernel over the paged F16 KV cache.
//
// One CTA = 64 consecutive query tokens of one sequence x one head; 4 warps x 16 queries.  K/V tiles of 64 positions
// are gathered from the KV pages into shared memory with cp.async (double buffered); S = Q.K^T and O += P.V run on
// the tensor cores through mma.sync.m16n8k16 (f16 inputs, fp32 accumulate) with the online soft-max between them in
// registers (S accumulator fragment == P operand fragment, so P never leaves the register file).  q is rounded to
// f16 and P to f16 exactly where ggml's CPU path rounds them (mul_mat on an F16 cache); the only difference to the
// per-token kernel is that P is rounded before the final 1/sum normalisation.
// Attention is ~4 % of the prefill flops (2 x n^2/2 x head_dim x 2 per head); the dense mat-muls, which are the rest,
// run on tcgen05 (kernels_gemm.cu).  L2-bound on K/V re-reads across the query tiles of a head.
#include "kernels.h"
#include "quant.cuh"

namespace b2 {

#ifndef FA_MINB
#define FA_MINB 2 // CTAs per SM the 64-query variant is compiled for (171 registers; 3 = 168 registers measured the same, 4 = 128 registers spills: 40.6 against 33.7 ms per 2048-token prefill)
#endif
constexpr int FA_BK = 64; // keys per tile; a CTA has BQ queries (64 or 128), 16 per warp

__device__ __forceinline__ void fa_cp16(void *smem, const void *gmem) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(smem)), "l"(gmem));
}
__device__ __forceinline__ void fa_commit() { asm volatile("cp.async.commit_group;"); }
template <int N> __device__ __forceinline__ void fa_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N)); }
__device__ __forceinline__ void fa_mma(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ uint32_t fa_pack(float lo, float hi) {
    const __half2 h = __floats2half2_rn(lo, hi);
    return *(const uint32_t *)&h;
}

// q: [n_tok][n_head*HD] fp32 (rotated); tokens [t0, t0 + n_q) belong to `slot`, token t0 + i sits at position pos0 + i
template <int HD, int FA_BQ>
__global__ void __launch_bounds__(FA_BQ * 2, FA_BQ == 64 ? FA_MINB : 1)
fattn_prefill_kernel(const float *__restrict__ q, int t0, int n_q, int pos0, int slot, int n_head, int n_head_kv, KVLayout kv,
                     float *__restrict__ out, float scale, __half *__restrict__ x_tiles) {
    constexpr int LD = HD + 8; // padded row (halves): conflict-free fragment loads
    constexpr int FA_THREADS = FA_BQ * 2;
    extern __shared__ __align__(16) __half fsm[];
    __half *sK = fsm;                       // [2][FA_BK][LD]
    __half *sV = fsm + 2 * FA_BK * LD;      // [2][FA_BK][LD]
    const int h = blockIdx.y, qt = blockIdx.x;
    const int hk = h / (n_head / n_head_kv);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, tq = lane & 3;
    const int q_lo = qt * FA_BQ;                       // first query of this CTA (relative to t0)
    const int n_kv = pos0 + min(q_lo + FA_BQ, n_q);    // keys this CTA may need: positions [0, n_kv)
    const int32_t *pt = kv.page_table + (int64_t)slot * kv.max_pages;

    auto load_tile = [&](int buf, int j0) {
        // 64 keys x HD halves = 64 x HD/8 16-byte chunks for K and the same for V
        constexpr int CH = HD / 8;
        for (int i = threadIdx.x; i < FA_BK * CH; i += FA_THREADS) {
            const int r = i / CH, c = i % CH;
            const int j = min(j0 + r, n_kv - 1); // clamp: rows past the end are masked anyway
            const int64_t cell = ((int64_t)pt[j / kv.page_size] * kv.page_size + j % kv.page_size) * kv.kvd + hk * HD + c * 8;
            fa_cp16(sK + ((size_t)buf * FA_BK + r) * LD + c * 8, kv.k_pool + cell);
            fa_cp16(sV + ((size_t)buf * FA_BK + r) * LD + c * 8, kv.v_pool + cell);
        }
        fa_commit();
    };

    // Q fragments for this warp's 16 queries: rows (g, g+8), k-chunks of 16 dims
    const int r0 = q_lo + warp * 16 + g, r1 = r0 + 8;
    uint32_t qa[HD / 16][4];
    {
        const float *q0 = q + ((int64_t)(t0 + min(r0, n_q - 1)) * n_head + h) * HD;
        const float *q1 = q + ((int64_t)(t0 + min(r1, n_q - 1)) * n_head + h) * HD;
#pragma unroll
        for (int kc = 0; kc < HD / 16; kc++) {
            const int c = kc * 16 + tq * 2;
            qa[kc][0] = fa_pack(q0[c], q0[c + 1]);
            qa[kc][1] = fa_pack(q1[c], q1[c + 1]);
            qa[kc][2] = fa_pack(q0[c + 8], q0[c + 9]);
            qa[kc][3] = fa_pack(q1[c + 8], q1[c + 9]);
        }
    }
    float o[HD / 8][4];
#pragma unroll
    for (int i = 0; i < HD / 8; i++) o[i][0] = o[i][1] = o[i][2] = o[i][3] = 0.f;
    float m0 = -INFINITY, m1 = -INFINITY, l0 = 0.f, l1 = 0.f;
    const int p_r0 = pos0 + r0, p_r1 = pos0 + r1; // absolute positions of this thread's two query rows

    const int n_tiles = (n_kv + FA_BK - 1) / FA_BK;
    const float sl = scale * 1.4426950408889634f; // scores are kept in units of log2: p = exp2(s sl - max)
    load_tile(0, 0);
    for (int it = 0; it < n_tiles; it++) {
        const int buf = it & 1, j0 = it * FA_BK;
        fa_wait<0>();
        __syncthreads(); // tile `it` is visible to everyone, and everyone is done with tile it - 1: its buffer may be refilled
        if (it + 1 < n_tiles) load_tile(buf ^ 1, j0 + FA_BK);
        const __half *K = sK + (size_t)buf * FA_BK * LD, *V = sV + (size_t)buf * FA_BK * LD;
        if (j0 <= pos0 + q_lo + warp * 16 + 15) { // (a tile wholly after this warp's last query is masked out: nothing to add)
        // ---- S = Q K^T : 8 n-tiles of 8 keys
        float s[FA_BK / 8][4];
#pragma unroll
        for (int nt = 0; nt < FA_BK / 8; nt++) {
            s[nt][0] = s[nt][1] = s[nt][2] = s[nt][3] = 0.f;
            // B fragments of two k-steps per ldmatrix.x4: matrices (dims +0, +8) of k-step 2 kp and of 2 kp + 1, 8 keys each
            const uint32_t kaddr = (uint32_t)__cvta_generic_to_shared(K + (size_t)(nt * 8 + (lane & 7)) * LD + (lane >> 3) * 8);
#pragma unroll
            for (int kp = 0; kp < HD / 32; kp++) {
                uint32_t b[4];
                asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
                             : "=r"(b[0]), "=r"(b[1]), "=r"(b[2]), "=r"(b[3])
                             : "r"(kaddr + kp * 64));
                fa_mma(s[nt], qa[2 * kp], b[0], b[1]);
                fa_mma(s[nt], qa[2 * kp + 1], b[2], b[3]);
            }
        }
        // ---- scale, causal mask, online soft-max (rows r0: regs 0,1 ; r1: regs 2,3)
        float mx0 = m0, mx1 = m1;
#pragma unroll
        for (int nt = 0; nt < FA_BK / 8; nt++) {
            const int j = j0 + nt * 8 + tq * 2;
            s[nt][0] = (j <= p_r0) ? s[nt][0] * sl : -INFINITY;
            s[nt][1] = (j + 1 <= p_r0) ? s[nt][1] * sl : -INFINITY;
            s[nt][2] = (j <= p_r1) ? s[nt][2] * sl : -INFINITY;
            s[nt][3] = (j + 1 <= p_r1) ? s[nt][3] * sl : -INFINITY;
            mx0 = fmaxf(mx0, fmaxf(s[nt][0], s[nt][1]));
            mx1 = fmaxf(mx1, fmaxf(s[nt][2], s[nt][3]));
        }
        mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 1));
        mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 2));
        mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 1));
        mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 2));
        // (position 0 is visible to every query, so after the first tile every maximum is finite: exp2(-inf - finite) = 0)
        const float c0 = exp2f(m0 - mx0), c1 = exp2f(m1 - mx1);
        m0 = mx0;
        m1 = mx1;
        l0 *= c0;
        l1 *= c1;
#pragma unroll
        for (int i = 0; i < HD / 8; i++) { o[i][0] *= c0; o[i][1] *= c0; o[i][2] *= c1; o[i][3] *= c1; }
        uint32_t pa[FA_BK / 16][4];
#pragma unroll
        for (int nt = 0; nt < FA_BK / 8; nt++) {
            const float e0 = exp2f(s[nt][0] - mx0), e1 = exp2f(s[nt][1] - mx0), e2 = exp2f(s[nt][2] - mx1), e3 = exp2f(s[nt][3] - mx1);
            const __half2 h01 = __floats2half2_rn(e0, e1), h23 = __floats2half2_rn(e2, e3);
            // the row sums use the f16-rounded probabilities, i.e. exactly what multiplies V
            const float2 f01 = __half22float2(h01), f23 = __half22float2(h23);
            l0 += f01.x + f01.y;
            l1 += f23.x + f23.y;
            pa[nt >> 1][(nt & 1) * 2 + 0] = *(const uint32_t *)&h01;
            pa[nt >> 1][(nt & 1) * 2 + 1] = *(const uint32_t *)&h23;
        }
        // ---- O += P V : B fragments straight from the row-major V tile with ldmatrix.trans
#pragma unroll
        for (int kc = 0; kc < FA_BK / 16; kc++) {
            // V[16 keys][8 dims] transposed, two groups of 8 dims per ldmatrix.x4 (lanes 16-31 name the rows of the second)
            const uint32_t vaddr = (uint32_t)__cvta_generic_to_shared(V + (size_t)(kc * 16 + (lane & 15)) * LD + (lane >> 4) * 8);
#pragma unroll
            for (int dp = 0; dp < HD / 16; dp++) {
                uint32_t b[4];
                asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
                             : "=r"(b[0]), "=r"(b[1]), "=r"(b[2]), "=r"(b[3])
                             : "r"(vaddr + dp * 32));
                fa_mma(o[2 * dp], pa[kc], b[0], b[1]);
                fa_mma(o[2 * dp + 1], pa[kc], b[2], b[3]);
            }
        }
        }
    }
    // ---- normalise and write: rows r0 / r1, dims dt*8 + tq*2 (+1)
    l0 += __shfl_xor_sync(0xffffffffu, l0, 1);
    l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
    l1 += __shfl_xor_sync(0xffffffffu, l1, 1);
    l1 += __shfl_xor_sync(0xffffffffu, l1, 2);
    const float i0 = 1.f / l0, i1 = 1.f / l1;
    if (x_tiles) { // straight into attn_output's fp16 operand tiles ([token tile][k tile][128 x 64], core-matrix order; kernels_gemm.cu)
        const int64_t k_tiles = (int64_t)n_head * HD / 64;
#pragma unroll
        for (int hh = 0; hh < 2; hh++) {
            const int r = hh ? r1 : r0;
            if (r >= n_q) continue;
            const int64_t row = t0 + r;
            const float inv = hh ? i1 : i0;
            __half *tile_row = x_tiles + (row >> 7) * k_tiles * 8192 + (int64_t)((row & 127) >> 3) * 512 + (row & 7) * 8 + tq * 2;
#pragma unroll
            for (int dt = 0; dt < HD / 8; dt++) {
                const int kc = h * (HD / 8) + dt; // 8-feature chunk index along K
                *(__half2 *)(tile_row + (int64_t)(kc >> 3) * 8192 + (kc & 7) * 64) = __floats2half2_rn(o[dt][2 * hh] * inv, o[dt][2 * hh + 1] * inv);
            }
        }
        return;
    }
    if (r0 < n_q) {
        float *orow = out + ((int64_t)(t0 + r0) * n_head + h) * HD + tq * 2;
#pragma unroll
        for (int dt = 0; dt < HD / 8; dt++) *(float2 *)(orow + dt * 8) = make_float2(o[dt][0] * i0, o[dt][1] * i0);
    }
    if (r1 < n_q) {
        float *orow = out + ((int64_t)(t0 + r1) * n_head + h) * HD + tq * 2;
#pragma unroll
        for (int dt = 0; dt < HD / 8; dt++) *(float2 *)(orow + dt * 8) = make_float2(o[dt][2] * i1, o[dt][3] * i1);
    }
}

bool fattn_supported(int hd) { return hd == 64 || hd == 128; }

void launch_fattn_prefill(const float *q, int t0, int n_q, int pos0, int slot, int n_head, int n_head_kv, int hd,
                          const KVLayout &kv, float *out, cudaStream_t s, void *x_tiles) {
    __half *xt = (__half *)x_tiles;
    if (n_q <= 0) return;
    const float scale = 1.0f / sqrtf((float)hd);
    // 64 queries per CTA; 128 (8 warps: half the K / V tile traffic and barriers per query) measured slower on a 2048-token
    // prefill of Llama-3-8B (34.4 against 33.7 ms), kept behind B200LLM_FATTN_BQ=128
    static const int bq_env = getenv("B200LLM_FATTN_BQ") ? atoi(getenv("B200LLM_FATTN_BQ")) : 0;
    const int bq = bq_env == 128 ? 128 : 64;
    const dim3 grid((n_q + bq - 1) / bq, n_head);
    const size_t smem = (size_t)4 * FA_BK * (hd + 8) * sizeof(__half);
    if (hd == 128) {
        static PerDeviceOnce cfg;
        cfg.run([&] {
            B2_CUDA(cudaFuncSetAttribute(fattn_prefill_kernel<128, 64>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            B2_CUDA(cudaFuncSetAttribute(fattn_prefill_kernel<128, 128>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        });
        if (bq == 128) fattn_prefill_kernel<128, 128><<<grid, 256, smem, s>>>(q, t0, n_q, pos0, slot, n_head, n_head_kv, kv, out, scale, xt);
        else fattn_prefill_kernel<128, 64><<<grid, 128, smem, s>>>(q, t0, n_q, pos0, slot, n_head, n_head_kv, kv, out, scale, xt);
    } else if (hd == 64) {
        if (bq == 128) fattn_prefill_kernel<64, 128><<<grid, 256, smem, s>>>(q, t0, n_q, pos0, slot, n_head, n_head_kv, kv, out, scale, xt);
        else fattn_prefill_kernel<64, 64><<<grid, 128, smem, s>>>(q, t0, n_q, pos0, slot, n_head, n_head_kv, kv, out, scale, xt);
    } else {
        throw Error(B2_EUNSUPPORTED, "flash prefill attention: head_dim must be 64 or 128");
    }
    g_launch_count++;
    B2_CUDA(cudaGetLastError());
}

} // namespace b2
