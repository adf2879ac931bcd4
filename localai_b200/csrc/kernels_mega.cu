// kernels_mega.cu -- the whole batch-1 decode step as ONE persistent cooperative kernel (one CTA per SM).
//
// Why: a decode step streams every weight byte exactly once (4.6 GB for Llama-3-8B Q4_K_M), so the only way to sit
// on the HBM roofline is to never let the memory pipeline drain.  With one kernel per operator (~165 launches) each
// launch pays ramp-up, an activation-quantisation prologue with no loads in flight, and a tail.  Here:
//   * the 126 MB L2 is the staging buffer.  A PREFETCH warp walks the static schedule of this CTA's weight rows for
//     every phase of every layer and issues bulk L2 prefetches (cp.async.bulk.prefetch.L2, the TMA engine) a bounded
//     distance ahead of what the consumers have finished.  Weights do not depend on activations, so it runs ahead
//     across phase boundaries: while the consumers sit in a grid barrier or re-quantise the activation row, HBM keeps
//     streaming into L2.
//   * 31 CONSUMER warps (64 registers each: latency is hidden by thread-level parallelism, and the loads hit L2) take
//     (row, K-segment) tasks and do the integer block dots against the activation row held in shared memory (same
//     arithmetic as the stand-alone kernels).  Segment partials are combined in a fixed order (deterministic) by the
//     phase epilogue (RoPE + KV store | residual | SwiGLU).
//   * phases are separated by a sense-reversing grid barrier; activations written by other SMs are read with ld.cg.
// Phases per layer: QKV -> attention -> wo -> gate/up -> down; then the LM head.  HBM-bound by construction;
// algorithmic bytes per launch = decode_stream_bytes + KV bytes.
#include "act_quant.cuh"
#include "attention.cuh"
#include "gemv_rows2.cuh"
#include "kernels.h"

namespace b2 {

constexpr int MEGA_CW = 31;                       // consumer warps
constexpr int MEGA_CT = MEGA_CW * 32;             // consumer threads
constexpr int MEGA_THREADS = MEGA_CT + 32;        // + prefetch warp = 1024
constexpr int MEGA_XS_BYTES = 36 * 1024;          // quantised activation row, or attention scratch
constexpr int MEGA_MAX_PART = 4096;               // per-phase partial sums (rows of this CTA x K-segments)
constexpr int MEGA_PB = 2;                        // prologue: 256-element blocks held in registers per warp
constexpr long long MEGA_PF_AHEAD = 320 * 1024;   // bytes per CTA the prefetcher may run ahead (x148 = 46 MB of L2)
constexpr int MEGA_PF_CHUNK = 16 * 1024;

constexpr size_t MEGA_OFF_PART = MEGA_XS_BYTES;                                  // float [MAX_PART]
constexpr size_t MEGA_OFF_RED = MEGA_OFF_PART + MEGA_MAX_PART * sizeof(float);   // double [32]
constexpr size_t MEGA_OFF_CNT = MEGA_OFF_RED + 32 * sizeof(double);              // unsigned long long consumed; int stop
constexpr size_t MEGA_SMEM = MEGA_OFF_CNT + 64;

// weights that the prefetcher has already pulled into L2: plain 128-bit loads, no L1 allocation, no register
// double buffer (L2 latency is hidden by the 31 warps)
struct LdL2 {
    static constexpr bool kPrefetch = false;
    static __device__ __forceinline__ int4 v4(const void *p) { return ld_stream(p); }
    static __device__ __forceinline__ int2 v2(const void *p) { return *(const int2 *)p; }
    static __device__ __forceinline__ uint32_t u16(const void *p) { return *(const uint16_t *)p; }
};

__device__ __forceinline__ int64_t act_row_bytes_dev(int fmt, int64_t K) {
    const int64_t b = fmt == ACT_Q8_K ? K + K / 256 * 4 + K / 16 * 2 : fmt == ACT_Q8_0 ? K + K / 32 * 4 : fmt == ACT_F16 ? K * 2 : K * 4;
    return (b + 15) / 16 * 16;
}
__device__ __forceinline__ void prefetch_l2(const void *p, uint32_t bytes) {
    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(p), "r"(bytes) : "memory");
}
__device__ __forceinline__ void consumer_sync() { asm volatile("bar.sync 2, %0;" ::"n"(MEGA_CT) : "memory"); }
__device__ __forceinline__ unsigned ld_acquire_u32(const unsigned *p) {
    unsigned v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

// sense-reversing grid barrier; called by ONE thread per CTA after a CTA-level barrier.  `gen` is the caller's copy
// of the generation counter (read once at kernel start).  arrive = acq_rel RMW on the count, the last arriver resets
// the count and release-increments the generation, everyone else acquire-polls the generation.
__device__ __forceinline__ void grid_barrier(unsigned *bar, unsigned n_cta, unsigned &gen) {
    unsigned prev;
    asm volatile("atom.add.acq_rel.gpu.global.u32 %0, [%1], 1;" : "=r"(prev) : "l"(bar) : "memory");
    if (prev == n_cta - 1) {
        asm volatile("st.relaxed.gpu.global.u32 [%0], 0;" ::"l"(bar) : "memory");
        asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(bar + 1) : "memory");
    } else {
        while (ld_acquire_u32(bar + 1) == gen) { }
    }
    gen++;
}

// This CTA's share of a phase: a contiguous range of rows (pair-aligned for PLAIN/QKV) in the concatenated row space.
__device__ __forceinline__ void cta_range(const MegaPhase &ph, int cta, int n_cta, int &u, int &u_end) {
    if (ph.kind == PH_ATT) { u = u_end = 0; return; }
    const int64_t a = (int64_t)ph.units * cta / n_cta, b = (int64_t)ph.units * (cta + 1) / n_cta;
    const int mul = ph.kind == PH_SWIGLU ? 1 : 2;
    u = (int)a * mul;
    u_end = (int)b * mul;
}

// row `u` of the concatenated row space -> (job, local row)
__device__ __forceinline__ void locate_row(const MegaPhase &ph, int u, int &jb, int &row) {
    jb = 0;
    row = u;
    while (jb + 1 < ph.n_jobs && row >= (int)ph.job[jb].rows) { row -= (int)ph.job[jb].rows; jb++; }
}

// ------------------------------------------------------------------------------------------------ the kernel
__global__ void __launch_bounds__(MEGA_THREADS, 1)
decode_mega_kernel(const MegaArgs a) {
    extern __shared__ __align__(128) uint8_t smem[];
    uint8_t *xs = smem;
    float *partial = (float *)(smem + MEGA_OFF_PART);
    double *s_part = (double *)(smem + MEGA_OFF_RED);
    volatile unsigned long long *consumed = (volatile unsigned long long *)(smem + MEGA_OFF_CNT);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int cta = blockIdx.x, n_cta = gridDim.x;

    if (tid == 0) *consumed = 0ull;
    __syncthreads();

    if (warp == MEGA_CW) {
        // ================= PREFETCHER: one lane pulls this CTA's weights into L2, a bounded distance ahead =================
        if (lane != 0) return;
        unsigned long long issued = 0;
        auto pf = [&](const uint8_t *p, long long bytes) {
            while (bytes > 0) {
                const uint32_t n = (uint32_t)min((long long)MEGA_PF_CHUNK, bytes);
                while ((long long)(issued - *consumed) > MEGA_PF_AHEAD) __nanosleep(100);
                prefetch_l2(p, n);
                p += n;
                bytes -= n;
                issued += n;
            }
        };
        for (int p = 0; p < a.n_phases; p++) {
            const MegaPhase &ph = a.phases[p];
            int u, u_end;
            cta_range(ph, cta, n_cta, u, u_end);
            if (ph.kind == PH_SWIGLU) {
                const int64_t s0 = ph.job[0].row_stride, s1 = ph.job[1].row_stride;
                const int step = max(1, (int)(MEGA_PF_CHUNK / s0));
                for (int r = u; r < u_end; r += step) {
                    const int n = min(step, u_end - r);
                    pf(ph.job[0].W + (int64_t)r * s0, (long long)n * s0);
                    pf(ph.job[1].W + (int64_t)r * s1, (long long)n * s1);
                }
            } else {
                while (u < u_end) {
                    int jb, row;
                    locate_row(ph, u, jb, row);
                    const int n = min(u_end - u, (int)ph.job[jb].rows - row);
                    pf(ph.job[jb].W + (int64_t)row * ph.job[jb].row_stride, (long long)n * ph.job[jb].row_stride);
                    u += n;
                }
            }
        }
        return;
    }

    // ================= CONSUMERS =================
    const int pos = a.pos[0], sl = a.slot[0];
    unsigned bar_gen = tid == 0 ? ld_acquire_u32(a.bar + 1) : 0;
    for (int p = 0; p < a.n_phases; p++) {
        const MegaPhase &ph = a.phases[p];
        long long t0 = 0, t1 = 0, t2 = 0;
        if (a.dbg && tid == 0) t0 = clock64();
        if (ph.kind == PH_ATT) {
            KVLayout kv{ph.k_pool, ph.v_pool, a.page_table, a.page_size, a.max_pages, a.kvd};
            for (int h = cta; h < a.n_head; h += n_cta)
                attention_head(ph.x + (int64_t)h * a.hd, pos + 1, a.page_table + (int64_t)sl * a.max_pages, kv,
                               h / (a.n_head / a.n_head_kv), a.hd, a.att_scale, ph.out + (int64_t)h * a.hd, (float *)xs, tid, MEGA_CT);
        } else {
            const int64_t K = ph.K;
            const int nb = (int)(K / 256);
            const int64_t xrb = act_row_bytes_dev(ph.fmt, K);
            int u, u_end;
            cta_range(ph, cta, n_cta, u, u_end);
            if (u < u_end) {
                // ---- prologue: (RMSNorm ->) quantise the activation row into shared memory (redundantly per CTA).
                // One L2 round trip: every warp pulls its (<= MEGA_PB) 256-element blocks into registers with all
                // loads in flight, the sum of squares is reduced from those registers, then they are quantised.
                const float *xr = ph.x;
                float v[MEGA_PB][8];
#pragma unroll
                for (int k = 0; k < MEGA_PB; k++) {
                    const int b = warp + k * MEGA_CW;
                    if (b < nb) {
                        const float4 *xb = (const float4 *)(xr + b * 256 + lane * 8);
                        const float4 a0 = __ldcg(xb), a1 = __ldcg(xb + 1);
                        v[k][0] = a0.x; v[k][1] = a0.y; v[k][2] = a0.z; v[k][3] = a0.w;
                        v[k][4] = a1.x; v[k][5] = a1.y; v[k][6] = a1.z; v[k][7] = a1.w;
                    }
                }
                float scale = 1.f;
                if (ph.norm_w) {
                    double sum = 0.0;
#pragma unroll
                    for (int k = 0; k < MEGA_PB; k++)
                        if (warp + k * MEGA_CW < nb) {
#pragma unroll
                            for (int i = 0; i < 8; i++) sum += (double)__fmul_rn(v[k][i], v[k][i]);
                        }
                    sum = warp_sum_d(sum);
                    if (lane == 0) s_part[warp] = sum;
                    consumer_sync();
                    double tot = 0.0;
                    for (int i = 0; i < MEGA_CW; i++) tot += s_part[i]; // every thread: same order, same result
                    const float mean = (float)(tot / (double)K);
                    scale = __fdiv_rn(1.0f, __fsqrt_rn(__fadd_rn(mean, a.eps)));
                }
#pragma unroll
                for (int k = 0; k < MEGA_PB; k++) {
                    const int b = warp + k * MEGA_CW;
                    if (b < nb) {
                        if (ph.norm_w) {
                            const float4 *wb = (const float4 *)(ph.norm_w + b * 256 + lane * 8);
                            const float4 w0 = wb[0], w1 = wb[1];
                            const float wv[8] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w};
#pragma unroll
                            for (int i = 0; i < 8; i++) v[k][i] = __fmul_rn(__fmul_rn(v[k][i], scale), wv[i]);
                            if (ph.xn_out && cta == 0) {
                                float *o = ph.xn_out + b * 256 + lane * 8;
                                *(float4 *)o = make_float4(v[k][0], v[k][1], v[k][2], v[k][3]);
                                *(float4 *)(o + 4) = make_float4(v[k][4], v[k][5], v[k][6], v[k][7]);
                            }
                        }
                        quant_block<true>(ph.fmt, v[k], lane, xs, K, b);
                    }
                }
                consumer_sync();
                if (a.dbg && tid == 0) t1 = clock64();
                // ---- tasks: (dot row, K-segment); dot rows = this CTA's rows (SwiGLU: its gate rows, then its up rows)
                const int n_r = u_end - u;
                const int n_dot = ph.kind == PH_SWIGLU ? 2 * n_r : n_r;
                int sh = 0;
                while ((n_dot << sh) < 2 * MEGA_CW && (nb >> (sh + 1)) >= 8 && (n_dot << (sh + 1)) <= MEGA_MAX_PART) sh++;
                const int n_tasks = n_dot << sh;
                for (int task = warp; task < n_tasks; task += MEGA_CW) {
                    const int dr = task >> sh, seg = task & ((1 << sh) - 1);
                    int jb, row;
                    if (ph.kind == PH_SWIGLU) {
                        jb = dr >= n_r;
                        row = u + (jb ? dr - n_r : dr);
                    } else {
                        locate_row(ph, u + dr, jb, row);
                    }
                    const FusedJob &j = ph.job[jb];
                    const uint8_t *r0 = j.W + (int64_t)row * j.row_stride;
                    float acc[1][1] = {{0.f}};
                    rows2_dot<LdL2, 1, 1>(j.type, r0, r0, K, seg, 1 << sh, xs, xrb, lane, acc);
                    const float s0 = warp_sum(acc[0][0]);
                    if (lane == 0) {
                        partial[task] = s0;
                        const unsigned long long by = (unsigned long long)((((int64_t)(seg + 1) * j.row_stride) >> sh) -
                                                                           (((int64_t)seg * j.row_stride) >> sh));
                        atomicAdd((unsigned long long *)consumed, by);
                    }
                }
                consumer_sync();
                // ---- epilogue: combine the segment partials in a fixed order (deterministic), then the phase's tail op
                const int n_seg = 1 << sh;
                auto dot_sum = [&](int dr) {
                    float s = 0.f;
                    for (int sg = 0; sg < n_seg; sg++) s += partial[(dr << sh) + sg];
                    return s;
                };
                if (ph.kind == PH_SWIGLU) {
                    for (int r = tid; r < n_r; r += MEGA_CT) {
                        const float g = dot_sum(r), up = dot_sum(n_r + r);
                        ph.out[u + r] = __fmul_rn(__fdiv_rn(g, __fadd_rn(1.0f, expf(-g))), up);
                    }
                } else if (ph.kind == PH_PLAIN) {
                    for (int r = tid; r < n_r; r += MEGA_CT) {
                        float s = dot_sum(r);
                        if (ph.residual) s += __ldcg(ph.residual + u + r);
                        ph.out[u + r] = s;
                    }
                } else { // PH_QKV: job 0 = q, 1 = k, 2 = v; rows (r, r+1) form a rotary pair inside one head
                    const int page = a.page_table[(int64_t)sl * a.max_pages + pos / a.page_size];
                    const int64_t cell = ((int64_t)page * a.page_size + pos % a.page_size) * a.kvd;
                    for (int pr = tid; pr < n_r / 2; pr += MEGA_CT) {
                        int jb, r;
                        locate_row(ph, u + 2 * pr, jb, r);
                        const float v0 = dot_sum(2 * pr), v1 = dot_sum(2 * pr + 1);
                        if (jb == 2) {
                            *(__half2 *)(ph.v_pool + cell + r) = __halves2half2(__float2half_rn(v0), __float2half_rn(v1));
                        } else {
                            const float *cs = a.rope_table + (int64_t)pos * a.hd + (r % a.hd);
                            const float c = cs[0], sn = cs[1];
                            const float o0 = __fsub_rn(__fmul_rn(v0, c), __fmul_rn(v1, sn));
                            const float o1 = __fadd_rn(__fmul_rn(v0, sn), __fmul_rn(v1, c));
                            if (jb == 0) *(float2 *)(ph.out + r) = make_float2(o0, o1);
                            else *(__half2 *)(ph.k_pool + cell + r) = __halves2half2(__float2half_rn(o0), __float2half_rn(o1));
                        }
                    }
                }
            }
        }
        // ---- phase boundary
        if (a.dbg && tid == 0) t2 = clock64();
        if (p + 1 < a.n_phases) {
            consumer_sync();
            if (tid == 0) grid_barrier(a.bar, n_cta, bar_gen);
            consumer_sync();
        }
        if (a.dbg && tid == 0) {
            long long *dd = a.dbg + ((int64_t)cta * a.n_phases + p) * 8;
            dd[0] = t0; dd[1] = t1 ? t1 : t0; dd[2] = t2; dd[3] = clock64(); dd[4] = 0; dd[5] = 0; dd[6] = 0; dd[7] = 0;
        }
    }
}

void launch_decode_mega(const MegaArgs &a, int n_sm, cudaStream_t s) {
    static thread_local bool configured = false;
    if (!configured) {
        B2_CUDA(cudaFuncSetAttribute(decode_mega_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)MEGA_SMEM));
        configured = true;
    }
    void *args[] = {(void *)&a};
    B2_CUDA(cudaLaunchCooperativeKernel((const void *)decode_mega_kernel, dim3(n_sm), dim3(MEGA_THREADS), args, MEGA_SMEM, s));
    g_launch_count++;
}

bool mega_phase_supported(const MegaPhase &ph) {
    if (ph.kind == PH_ATT) return true;
    if (ph.fmt != ACT_Q8_K && ph.fmt != ACT_Q8_0) return false;
    if (act_row_bytes((ActFmt)ph.fmt, ph.K) > MEGA_XS_BYTES || ph.K / 256 > MEGA_PB * MEGA_CW) return false;
    int64_t rows = 0;
    for (int j = 0; j < ph.n_jobs; j++) {
        if (ph.job[j].rows % 2 != 0) return false;
        rows += ph.job[j].rows;
    }
    // the per-CTA share of the rows (x2 K-segments at most when it is small) must fit the partial buffer
    return rows / 132 + 4 <= MEGA_MAX_PART;
}

size_t mega_attention_ctx_limit(int hd) {
    // scores + page table + q + per-warp partials must fit the activation region
    const size_t fixed = sizeof(float) * ((size_t)hd + 32 + (size_t)MEGA_CW * hd + 16);
    return fixed >= (size_t)MEGA_XS_BYTES ? 0 : (size_t)((MEGA_XS_BYTES - fixed) / 4.6) - 8;
}

} // namespace b2
