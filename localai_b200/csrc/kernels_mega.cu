// kernels_mega.cu -- the whole batch-1 decode step as ONE persistent cooperative kernel (one CTA per SM).
//
// A decode step streams every weight byte exactly once (4.6 GB for Llama-3-8B Q4_K_M) and does ~2 integer MACs per
// byte, so it lives or dies by how close HBM stays to 100 % busy.  Four things keep it there:
//
//   1. THE WEIGHTS ARE LAID OUT FOR THIS KERNEL.  At LoadModel the rows of every matrix are scattered (MegaStreams)
//      into ONE CONTIGUOUS BYTE STREAM PER (CTA, WARP), in exactly the order that warp consumes them across all phases
//      of all layers: tile = [16-byte header][n row-splits].  The header says what the tile is (type, rows, pitch,
//      where the partial sums go), how many bytes the tile two ahead has, and whether it is the warp's last tile of
//      its phase.  The kernel therefore holds no schedule at all: "copy the next tile of my stream" is one
//      cp.async.bulk of a size read from shared memory.  (Same bytes as the GGUF rows + 16 B per ~9 KB tile.)
//   2. WEIGHTS STREAM THROUGH SHARED MEMORY, AHEAD OF THE MATH, ACROSS PHASE BOUNDARIES.  Every warp owns a private
//      two-stage ring (2 x 9 KB) that it fills itself (cp.async.bulk -> mbarrier complete_tx, the TMA engine).
//      Weights do not depend on activations, so while the CTA sits in a grid barrier or re-quantises the activation
//      row its 16 stages (144 KB per SM, 21 MB per GPU) are already filling with the next phase's rows.  No warp ever
//      waits for another warp's data; the only CTA-wide syncs are the phase boundaries.
//   3. ACTIVATIONS ARE REGISTER-STATIONARY.  A lane always handles the same K-slice of a row (K-quants: half a
//      256-weight super-block; Q8_0: every 32nd 16-byte piece), so after the phase prologue has normalised and
//      quantised the activation row (bit-exact with quantize_row_q8_K / q8_0) each lane keeps its 128 int8
//      activations, their scale and block sums in registers for the whole phase.  Rows longer than one warp pass are
//      split over 2/4/8 warps (K-splits); split partials are combined in a fixed order by the phase epilogue.
//   4. ONE LAUNCH.  Phases (QKV+RoPE+KV store -> attention -> wo+residual -> gate/up+SwiGLU -> down+residual, then
//      the LM head) are separated by a grid barrier on one monotonically growing counter; activations written by
//      other SMs are read with ld.cg.
//
// Arithmetic = ggml_vec_dot_{q4_K,q5_K,q6_K}_q8_K / q8_0_q8_0 (integer block dots, fp32 scaling); only the fp32
// summation order differs from the scalar reference.  Algorithmic bytes per launch = decode_stream_bytes + KV bytes read.
#include <algorithm>
#include <vector>

#include "act_quant.cuh"
#include "kernels.h"
#include "quant.cuh"

namespace b2 {

constexpr int MG_WARPS = 8;
constexpr int MG_THREADS = MG_WARPS * 32;
constexpr int MG_STAGES = 2;                 // per warp
constexpr int MG_TILE_HDR = 16;              // tile header, in-band
constexpr int MG_MAX_PART = 2048;            // per-phase partial sums: rows of this CTA x K-splits
constexpr int MG_PB = 8;                     // prologue: 256-element blocks a warp holds in registers (K <= 16384)
constexpr int MG_MAX_PASS = 8;               // warp passes per tile
constexpr int MG_MAX_PAGES = 512;            // page-table entries of the sequence kept in shared memory
constexpr int MG_PH_WIN = 3;                 // phase descriptors kept in shared memory: phases p - 1, p, p + 1
constexpr int MG_PH_SLOT = (int)((sizeof(MegaPhase) + 15) / 16 * 16);
constexpr int MG_SMEM_MAX = 227 * 1024;
static_assert(sizeof(MegaPhase) % 4 == 0, "MegaPhase is copied word by word");
// Shared memory: [rings: 8 warps x 2 stages][activation row / attention scratch][partial sums][reduction scratch][mbarriers]
// [phase descriptors][page table].  The first two are sized per model (MegaGeo).
struct MgSmem {
    uint32_t xs, part, red, bar, ph, pt, total;
};
__host__ __device__ inline MgSmem mg_smem_layout(const MegaGeo &g) {
    MgSmem l;
    l.xs = (uint32_t)(MG_WARPS * MG_STAGES * g.stage_bytes);
    l.part = l.xs + (uint32_t)g.xs_bytes;
    l.red = l.part + MG_MAX_PART * (uint32_t)sizeof(float);
    l.bar = l.red + 32 * (uint32_t)sizeof(double);
    l.ph = l.bar + MG_WARPS * MG_STAGES * (uint32_t)sizeof(uint64_t);
    l.pt = l.ph + MG_PH_WIN * MG_PH_SLOT;
    l.total = l.pt + MG_MAX_PAGES * (uint32_t)sizeof(int);
    return l;
}
constexpr long long MG_WATCHDOG = 40000000000ll; // cycles (~20 s) after which a wait gives up and traps instead of hanging the GPU

// ------------------------------------------------------------------------------------------------ geometry
// How a phase's rows map onto lanes (computed on the host, mega_phase_prepare).  K-quants: item = half a super-block
// (128 weights); a (row, K-split) takes lpr = 2*nb/S lanes; when that is a power of two below 32, rpp = 32/lpr rows
// share a warp pass.  Q8_0: lane l takes the 16-byte pieces l, l+32, ... (np of them) of its K-split.
inline bool mg_split(int fmt, int64_t K, bool pre_quantised, int &S, int &lpr, int &rpp, int &np) {
    np = 0;
    if (fmt == ACT_Q8_K) {
        if (K % 256) return false;
        const int nb = (int)(K / 256);
        if (nb > (pre_quantised ? 16 * MG_WARPS : MG_PB * MG_WARPS)) return false; // the register prologue holds MG_PB blocks per warp
        for (S = 1; S <= MG_WARPS; S *= 2)
            if (nb % S == 0 && 2 * nb / S <= 32) {
                lpr = 2 * nb / S;
                rpp = (lpr & (lpr - 1)) == 0 ? 32 / lpr : 1;
                return true;
            }
        return false;
    }
    if (fmt == ACT_Q8_0) {
        if (K / 256 > MG_PB * MG_WARPS) return false;
        for (S = 1; S <= MG_WARPS; S *= 2)
            if (K % (S * 512) == 0 && K / S <= 8192) {
                lpr = 32;
                rpp = 1;
                np = (int)(K / S / 512);
                return true;
            }
        return false;
    }
    return false;
}

// Tile geometry of one matrix: rb = bytes of one row's K-split inside a tile (= the row pitch), R = rows per tile.
//   Q4_K / Q5_K: the K-split's blocks.  Q6_K: its 208-byte blocks, then its fp16 super-block scales (padded to 16).
//   Q8_0: its int8 range, then its fp16 block scales.
inline bool mg_job_geo(int type, int64_t K, int S, int rpp, int tile_data, int &rb, int &R) {
    const int64_t nbs = K / 256 / S;
    if (type == B2_T_Q4_K) rb = (int)(144 * nbs);
    else if (type == B2_T_Q5_K) rb = (int)(176 * nbs);
    else if (type == B2_T_Q6_K) rb = (int)(Q6_BLK * nbs + (2 * nbs + 15) / 16 * 16);
    else if (type == B2_T_Q8_0) rb = (int)(K / S + K / S / 16);
    else return false;
    if (rb <= 0 || rb % 16 || rb > tile_data) return false;
    R = tile_data / rb;
    if (R > MG_MAX_PASS * rpp) R = MG_MAX_PASS * rpp;
    R = R / rpp * rpp;
    if (rpp == 1) { // the row loops work on groups of 2 rows with a merged reduction
        if (R >= 2) R = R / 2 * 2;
    }
    return R >= rpp && R > 0 && R < 256;
}

// ------------------------------------------------------------------------------------------------ primitives
__device__ __forceinline__ uint32_t mg_s32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mg_mbar_init(uint64_t *bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(mg_s32(bar)), "r"(count));
}
__device__ __forceinline__ void mg_prefetch_l2(const void *p, uint32_t bytes) {
    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(p), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mg_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mg_try_wait(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n"
        : "=r"(ok)
        : "r"(bar), "r"(parity)
        : "memory");
    return ok != 0;
}
// Wait for a stage to fill.  A wait that lasts MG_WATCHDOG cycles (a bug, never a slow copy) traps: the launch fails
// with an error instead of occupying the GPU for ever.
__device__ __forceinline__ void mg_wait(uint32_t bar, uint32_t parity) {
    if (mg_try_wait(bar, parity)) return;
    const long long t0 = clock64();
    while (!mg_try_wait(bar, parity))
        if (clock64() - t0 > MG_WATCHDOG) __trap();
}
__device__ __forceinline__ void mg_bulk(uint32_t dst, const void *src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst), "l"(src),
                 "r"(bytes), "r"(bar)
                 : "memory");
}
__device__ __forceinline__ long long mg_gtime() { // nanoseconds, one clock for all SMs (profiling build only)
    long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}
__device__ __forceinline__ int4 mg_lds128(uint32_t a) {
    int4 r;
    asm volatile("ld.shared.v4.b32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "r"(a));
    return r;
}
__device__ __forceinline__ int2 mg_lds64(uint32_t a) {
    int2 r;
    asm volatile("ld.shared.v2.b32 {%0,%1}, [%2];" : "=r"(r.x), "=r"(r.y) : "r"(a));
    return r;
}
__device__ __forceinline__ uint32_t mg_lds_u16(uint32_t a) {
    uint32_t r;
    asm volatile("ld.shared.u16 %0, [%1];" : "=r"(r) : "r"(a));
    return r;
}
__device__ __forceinline__ uint32_t mg_ldg_u16(const void *p) {
    uint32_t r;
    asm volatile("ld.global.nc.u16 %0, [%1];" : "=r"(r) : "l"(p));
    return r;
}
__device__ __forceinline__ int mg_dp4a_us(uint32_t a, int b, int c) { // 4 x (u8 * s8) + c
    int d;
    asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}
// ---- LL words: {fp32 value, 32-bit tag} moved as ONE 64-bit access, so value and "ready" flag travel together ----
__device__ __forceinline__ void ll_st(uint2 *p, float v, uint32_t tag) {
    const unsigned long long w = ((unsigned long long)tag << 32) | (unsigned long long)__float_as_uint(v);
    asm volatile("st.relaxed.gpu.global.b64 [%0], %1;" ::"l"(p), "l"(w) : "memory");
}
__device__ __forceinline__ void ll_st2(uint2 *p, float v0, float v1, uint32_t tag) { // p 16-byte aligned
    const unsigned long long w0 = ((unsigned long long)tag << 32) | (unsigned long long)__float_as_uint(v0);
    const unsigned long long w1 = ((unsigned long long)tag << 32) | (unsigned long long)__float_as_uint(v1);
    asm volatile("st.relaxed.gpu.global.v2.b64 [%0], {%1, %2};" ::"l"(p), "l"(w0), "l"(w1) : "memory");
}
__device__ __forceinline__ unsigned long long ll_ld(const uint2 *p) {
    unsigned long long w;
    asm volatile("ld.relaxed.gpu.global.b64 %0, [%1];" : "=l"(w) : "l"(p) : "memory");
    return w;
}
__device__ __forceinline__ void ll_ld2(const uint2 *p, unsigned long long &w0, unsigned long long &w1) { // p 16-byte aligned
    asm volatile("ld.relaxed.gpu.global.v2.b64 {%0, %1}, [%2];" : "=l"(w0), "=l"(w1) : "l"(p) : "memory");
}
// the same across GPUs (peer memory over NVLink): system scope
__device__ __forceinline__ void ll_st_sys(uint2 *p, float v, uint32_t tag) {
    const unsigned long long w = ((unsigned long long)tag << 32) | (unsigned long long)__float_as_uint(v);
    asm volatile("st.relaxed.sys.global.b64 [%0], %1;" ::"l"(p), "l"(w) : "memory");
}
__device__ __forceinline__ unsigned long long ll_ld_sys(const uint2 *p) {
    unsigned long long w;
    asm volatile("ld.relaxed.sys.global.b64 %0, [%1];" : "=l"(w) : "l"(p) : "memory");
    return w;
}
__device__ __forceinline__ uint32_t ll_tag(unsigned long long w) { return (uint32_t)(w >> 32); }
__device__ __forceinline__ float ll_val(unsigned long long w) { return __uint_as_float((uint32_t)w); }
// Spin budget of every LL wait: a wait that long is a bug (a producer that never ran), and traps instead of hanging.
constexpr unsigned MG_SPIN_LIMIT = 1u << 27;
__constant__ unsigned g_ll_poll_ns = 32; // sleep between two polls of an LL word (set per launch from MegaArgs.ll_poll_ns when it changes)
__device__ __forceinline__ void ll_backoff(unsigned &spins) {
    const unsigned ns = g_ll_poll_ns;
    if (ns) __nanosleep(ns);
    if (++spins > MG_SPIN_LIMIT) __trap();
}
// The phase counter: a HINT that lets a CTA wait quietly.  Every CTA adds 1 after storing a phase's outputs -- a relaxed
// reduction, not ordered against those stores -- and a consumer's thread 0 polls the one word until all CTAs of the
// producing phase have added theirs; only then do its warps load the row.  The LL tags remain the proof that a word is
// there: a word that lags behind the counter is simply polled again.
__device__ __forceinline__ void mg_hint_arrive(unsigned *cnt) {
    asm volatile("red.relaxed.gpu.global.add.u32 [%0], 1;" ::"l"(cnt) : "memory");
}
__device__ __forceinline__ void mg_hint_wait(const unsigned *cnt, unsigned target) {
    unsigned v, spins = 0;
    while (true) {
        asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(cnt) : "memory");
        if ((int)(v - target) >= 0) return;
        ll_backoff(spins);
    }
}
// one word, polled by one thread until it carries `tag`
__device__ __forceinline__ float ll_wait(const uint2 *p, uint32_t tag) {
    unsigned long long w = ll_ld(p);
    unsigned spins = 0;
    while (ll_tag(w) != tag) {
        ll_backoff(spins);
        w = ll_ld(p);
    }
    return ll_val(w);
}
// 8 consecutive words (64 bytes, 16-byte aligned) -> v; false when some word does not carry `tag` yet
__device__ __forceinline__ bool ll_ld8(const uint2 *p, uint32_t tag, float (&v)[8]) {
    unsigned long long w[8];
    ll_ld2(p, w[0], w[1]);
    ll_ld2(p + 2, w[2], w[3]);
    ll_ld2(p + 4, w[4], w[5]);
    ll_ld2(p + 6, w[6], w[7]);
    bool ok = true;
#pragma unroll
    for (int i = 0; i < 8; i++) {
        ok = ok && ll_tag(w[i]) == tag;
        v[i] = ll_val(w[i]);
    }
    return ok;
}

inline int64_t mg_act_row_bytes_host(int fmt, int64_t K) {
    const int64_t b = fmt == ACT_Q8_K ? K + K / 256 * 4 + K / 16 * 2 : K + K / 32 * 4;
    return (b + 15) / 16 * 16;
}

// This CTA's share of a phase: a contiguous range of rows (pair-aligned for PLAIN/QKV) in the concatenated row space.
__host__ __device__ __forceinline__ void mg_cta_range(const MegaPhase &ph, int cta, int n_cta, int &u, int &u_end) {
    if (ph.kind == PH_ATT) { u = u_end = 0; return; }
    // units_q = units / n_cta and units_r = units % n_cta come from the host: the first units_r CTAs take one more
    const int a = cta * ph.units_q + (cta < ph.units_r ? cta : ph.units_r), b = a + ph.units_q + (cta < ph.units_r ? 1 : 0);
    const int mul = ph.kind == PH_SWIGLU ? 1 : 2;
    u = a * mul;
    u_end = b * mul;
}
// row `u` of the concatenated row space -> (job, local row)
__device__ __forceinline__ void mg_locate_row(const MegaPhase &ph, int u, int &jb, int &row) {
    jb = 0;
    row = u;
    while (jb + 1 < ph.n_jobs && row >= (int)ph.job[jb].rows) { row -= (int)ph.job[jb].rows; jb++; }
}


// ------------------------------------------------------------------------------------------------ block dots
// A lane's place in a phase (fixed for the whole phase, so its activations stay in registers).
struct MgLane {
    int sr;       // which of the rpp rows of a warp pass
    int kb, h;    // K-quants: super-block inside the K-split, half of it
    int span;     // lanes that hold one row's partial sums (32, or lpr when rows share a pass)
    bool on, head; // lane takes part; lane writes the row's sum
};

// Q4_K, one half super-block (sub-blocks 4h..4h+3 = 64 bytes of nibbles) against the lane's 128 activations.
//   block: [d f16][dmin f16][12 B packed 6-bit scales/mins][128 B nibbles]; group j (32 B): low nibbles = sub-block
//   2j, high nibbles = sub-block 2j+1.  xr[8s..8s+7] = int8 activations of the lane's sub-block s, bs32[s] their sum.
__device__ __forceinline__ float mg_dot_q4k(uint32_t blk, int h, const int (&xr)[32], float dx, const int (&bs32)[4]) {
    const int4 hdr = mg_lds128(blk);
    const uint32_t qa = blk + 16 + 64 * h;
    const int4 q0 = mg_lds128(qa), q1 = mg_lds128(qa + 16), q2 = mg_lds128(qa + 32), q3 = mg_lds128(qa + 48);
    // the 4 scales / 4 mins of this half, one per byte (ggml's kmask unpacking of the 12 packed bytes)
    const uint32_t u0 = hdr.y, u1 = hdr.z, u2 = hdr.w;
    const uint32_t sc_hi = (u2 & 0x0F0F0F0Fu) | ((u0 >> 2) & 0x30303030u);
    const uint32_t mn_hi = ((u2 >> 4) & 0x0F0F0F0Fu) | ((u1 >> 2) & 0x30303030u);
    const uint32_t sc4 = h ? sc_hi : (u0 & 0x3F3F3F3Fu), mn4 = h ? mn_hi : (u1 & 0x3F3F3F3Fu);
    const uint32_t w[16] = {(uint32_t)q0.x, (uint32_t)q0.y, (uint32_t)q0.z, (uint32_t)q0.w, (uint32_t)q1.x, (uint32_t)q1.y,
                            (uint32_t)q1.z, (uint32_t)q1.w, (uint32_t)q2.x, (uint32_t)q2.y, (uint32_t)q2.z, (uint32_t)q2.w,
                            (uint32_t)q3.x, (uint32_t)q3.y, (uint32_t)q3.z, (uint32_t)q3.w};
    int s0 = 0, s1 = 0, s2 = 0, s3 = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) {
        s0 = mg_dp4a_us(w[i] & 0x0F0F0F0Fu, xr[i], s0);
        s1 = mg_dp4a_us(w[i] & 0xF0F0F0F0u, xr[8 + i], s1); // 16 * (high nibbles . x)
        s2 = mg_dp4a_us(w[8 + i] & 0x0F0F0F0Fu, xr[16 + i], s2);
        s3 = mg_dp4a_us(w[8 + i] & 0xF0F0F0F0u, xr[24 + i], s3);
    }
    const int isum = (int)(sc4 & 0xFF) * s0 + (int)((sc4 >> 8) & 0xFF) * (s1 >> 4) + (int)((sc4 >> 16) & 0xFF) * s2 +
                     (int)(sc4 >> 24) * (s3 >> 4);
    const int imin = (int)(mn4 & 0xFF) * bs32[0] + (int)((mn4 >> 8) & 0xFF) * bs32[1] + (int)((mn4 >> 16) & 0xFF) * bs32[2] +
                     (int)(mn4 >> 24) * bs32[3];
    const float2 dd = __half22float2(*(const __half2 *)&hdr.x);
    return (dd.x * dx) * (float)isum - (dd.y * dx) * (float)imin;
}

// Q5_K, one half super-block of a 176-byte block [d f16][dmin f16][12 B scales/mins][qh 32 B][qs 128 B]: as Q4_K, and bit s
// of qh[l] is the fifth bit of weight l of sub-block s (dequantize_row_q5_K).  Only the 70B-class attn_v rows use it.
__device__ __forceinline__ float mg_dot_q5k(uint32_t blk, int h, const int (&xr)[32], float dx, const int (&bs32)[4]) {
    const int4 hdr = mg_lds128(blk);
    const int4 h0 = mg_lds128(blk + 16), h1 = mg_lds128(blk + 32);
    const uint32_t qa = blk + 48 + 64 * h;
    const int4 q0 = mg_lds128(qa), q1 = mg_lds128(qa + 16), q2 = mg_lds128(qa + 32), q3 = mg_lds128(qa + 48);
    const uint32_t u0 = hdr.y, u1 = hdr.z, u2 = hdr.w;
    const uint32_t sc_hi = (u2 & 0x0F0F0F0Fu) | ((u0 >> 2) & 0x30303030u);
    const uint32_t mn_hi = ((u2 >> 4) & 0x0F0F0F0Fu) | ((u1 >> 2) & 0x30303030u);
    const uint32_t sc4 = h ? sc_hi : (u0 & 0x3F3F3F3Fu), mn4 = h ? mn_hi : (u1 & 0x3F3F3F3Fu);
    const uint32_t w[16] = {(uint32_t)q0.x, (uint32_t)q0.y, (uint32_t)q0.z, (uint32_t)q0.w, (uint32_t)q1.x, (uint32_t)q1.y,
                            (uint32_t)q1.z, (uint32_t)q1.w, (uint32_t)q2.x, (uint32_t)q2.y, (uint32_t)q2.z, (uint32_t)q2.w,
                            (uint32_t)q3.x, (uint32_t)q3.y, (uint32_t)q3.z, (uint32_t)q3.w};
    const uint32_t H[8] = {(uint32_t)h0.x, (uint32_t)h0.y, (uint32_t)h0.z, (uint32_t)h0.w, (uint32_t)h1.x, (uint32_t)h1.y, (uint32_t)h1.z, (uint32_t)h1.w};
    int s0 = 0, s1 = 0, s2 = 0, s3 = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) {
        const uint32_t hb = H[i] >> (4 * h); // bits 0..3 of every byte: the fifth bits of this half's four sub-blocks
        s0 = mg_dp4a_us((w[i] & 0x0F0F0F0Fu) | ((hb << 4) & 0x10101010u), xr[i], s0);
        s1 = mg_dp4a_us(((w[i] >> 4) & 0x0F0F0F0Fu) | ((hb << 3) & 0x10101010u), xr[8 + i], s1);
        s2 = mg_dp4a_us((w[8 + i] & 0x0F0F0F0Fu) | ((hb << 2) & 0x10101010u), xr[16 + i], s2);
        s3 = mg_dp4a_us(((w[8 + i] >> 4) & 0x0F0F0F0Fu) | ((hb << 1) & 0x10101010u), xr[24 + i], s3);
    }
    const int isum = (int)(sc4 & 0xFF) * s0 + (int)((sc4 >> 8) & 0xFF) * s1 + (int)((sc4 >> 16) & 0xFF) * s2 + (int)(sc4 >> 24) * s3;
    const int imin = (int)(mn4 & 0xFF) * bs32[0] + (int)((mn4 >> 8) & 0xFF) * bs32[1] + (int)((mn4 >> 16) & 0xFF) * bs32[2] +
                     (int)(mn4 >> 24) * bs32[3];
    const float2 dd = __half22float2(*(const __half2 *)&hdr.x);
    return (dd.x * dx) * (float)isum - (dd.y * dx) * (float)imin;
}

// Q6_K, one half super-block n (weights 128n..128n+127) of a 208-byte device block [ql 128][qh 64][scales 16]:
//   ql[64n + l] low/high nibble -> weights l / 64+l, ql[64n + 32 + l] -> 32+l / 96+l, qh[32n + l] holds their four
//   2-bit tops; scale of the 16-weight sub-block (column c = 0..3, half hf = l/16): sc[8n + 2c + hf].
//   xr[8c + i] = activations of column c, word i; bs16[2c + hf] = their 16-element sums.  d is the fp16 row-tail scale.
__device__ __forceinline__ float mg_dot_q6k(uint32_t blk, int n, const int (&xr)[32], float dx, const int (&bs16)[8], uint32_t d16) {
    const uint32_t la = blk + 64 * n;
    const int4 a0 = mg_lds128(la), a1 = mg_lds128(la + 16), b0 = mg_lds128(la + 32), b1 = mg_lds128(la + 48);
    const int4 h0 = mg_lds128(blk + 128 + 32 * n), h1 = mg_lds128(blk + 144 + 32 * n);
    const int2 scw = mg_lds64(blk + 192 + 8 * n);
    const uint32_t A[8] = {(uint32_t)a0.x, (uint32_t)a0.y, (uint32_t)a0.z, (uint32_t)a0.w, (uint32_t)a1.x, (uint32_t)a1.y, (uint32_t)a1.z, (uint32_t)a1.w};
    const uint32_t B[8] = {(uint32_t)b0.x, (uint32_t)b0.y, (uint32_t)b0.z, (uint32_t)b0.w, (uint32_t)b1.x, (uint32_t)b1.y, (uint32_t)b1.z, (uint32_t)b1.w};
    const uint32_t H[8] = {(uint32_t)h0.x, (uint32_t)h0.y, (uint32_t)h0.z, (uint32_t)h0.w, (uint32_t)h1.x, (uint32_t)h1.y, (uint32_t)h1.z, (uint32_t)h1.w};
    int s[4][2] = {{0, 0}, {0, 0}, {0, 0}, {0, 0}};
#pragma unroll
    for (int i = 0; i < 8; i++) {
        const int hf = i >> 2;
        const uint32_t hh = H[i];
        s[0][hf] = mg_dp4a_us((A[i] & 0x0F0F0F0Fu) | ((hh << 4) & 0x30303030u), xr[i], s[0][hf]);
        s[1][hf] = mg_dp4a_us((B[i] & 0x0F0F0F0Fu) | ((hh << 2) & 0x30303030u), xr[8 + i], s[1][hf]);
        s[2][hf] = mg_dp4a_us(((A[i] >> 4) & 0x0F0F0F0Fu) | (hh & 0x30303030u), xr[16 + i], s[2][hf]);
        s[3][hf] = mg_dp4a_us(((B[i] >> 4) & 0x0F0F0F0Fu) | ((hh >> 2) & 0x30303030u), xr[24 + i], s[3][hf]);
    }
    // stored quants are q+32: subtract 32 * sum(x) of each 16-weight sub-block (exact, integer)
    int isum = 0;
#pragma unroll
    for (int c = 0; c < 4; c++)
#pragma unroll
        for (int hf = 0; hf < 2; hf++) {
            const int k = 2 * c + hf;
            const uint32_t word = k < 4 ? (uint32_t)scw.x : (uint32_t)scw.y;
            const int scale = (int)(int8_t)((word >> (8 * (k & 3))) & 0xFF);
            isum += scale * (s[c][hf] - 32 * bs16[k]);
        }
    return (h2f((uint16_t)d16) * dx) * (float)isum;
}

// warp-reduce over the `span` lanes that hold one row (span = 32, or a power of two when rows share a pass)
__device__ __forceinline__ float mg_row_sum(float v, int span) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1)
        if (o < span) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// Merged butterfly reduction: the sums of 2 per-lane values over the whole warp in 5 shuffles instead of 10.
// Result: lanes 0 / 16 hold rows 0 / 1.
__device__ __forceinline__ float mg_reduce2(float v0, float v1, int lane) {
    const bool b4 = lane & 16;
    float t = (b4 ? v1 : v0) + __shfl_xor_sync(0xffffffffu, b4 ? v0 : v1, 16);
    t += __shfl_xor_sync(0xffffffffu, t, 8);
    t += __shfl_xor_sync(0xffffffffu, t, 4);
    t += __shfl_xor_sync(0xffffffffu, t, 2);
    t += __shfl_xor_sync(0xffffffffu, t, 1);
    return t;
}


// ------------------------------------------------------------------------------------------------ attention
// row of position j in a layer's K or V pool; 32-bit element index (the host checks that a layer pool has < 2^31 halves)
__device__ __forceinline__ const __half *mg_kv_row(const __half *pool, const int *pbase, int j, int page_shift, int kvd) {
    const uint32_t cell = ((uint32_t)pbase[j >> page_shift] << page_shift) + (uint32_t)(j & ((1 << page_shift) - 1));
    return pool + cell * (uint32_t)kvd;
}

// Attention of ONE (head, query token), shared by `parts` CTAs (contiguous position ranges; parts = 1: one CTA), written
// for LATENCY: at short context the phase is a chain of dependent round trips, so everything whose address is known up
// front is requested up front, and the CTAs of a head talk through LL words -- no barrier, no atomic:
//   0. the K rows (4 threads per position, 64 contiguous bytes per 4-lane group and load) and V rows of my first
//      positions are requested from the cache BEFORE q is waited for (the QKV phase prefetched them into L2);
//   1. q arrives (LL, from the QKV phase); the row of the CURRENT position is not in the cache yet as far as this SM can
//      know, so its K and V values come as LL words too (kcur / vcur) and replace the cache row in registers;
//   2. scores of my positions -> the head's score row (LL);
//   3. every part reads ALL the scores of the head and does the soft-max redundantly (a few hundred elements), so the
//      probabilities -- including their rounding to f16 -- are exactly those of the unsplit computation;
//   4. p.V over my positions -> a partial output row (LL); part 0 adds the partial rows in part order and publishes
//      the head's output (LL) for the wo phase.
// Numerics as attention_head (attention.cuh): q and the normalised probabilities rounded to f16, fp32 accumulation.
// scratch: q[hd] | red[32] | part[8][hd] | scores[n_kv].  hd = 64 or 128.
__device__ __noinline__ void mg_attention_ll(const uint2 *__restrict__ q_ll, const uint2 *__restrict__ kcur_ll, const uint2 *__restrict__ vcur_ll,
                                             uint32_t tag_in, int n_kv, const int *pbase, const __half *k_pool, const __half *v_pool,
                                             int page_shift, int kvd, int hk, int hd, float scale, uint2 *__restrict__ out_ll, float *scratch,
                                             int tid, int part, int parts, uint2 *g_scores, uint2 *g_part, uint32_t tag, bool combine_here) {
    const int lane = tid & 31, warp = tid >> 5;
    float *qs = scratch, *red = qs + hd, *pw = red + 32, *sc = pw + MG_WARPS * hd;
    const int pos = n_kv - 1;                // the current token's position
    const int chunk = (n_kv + parts - 1) / parts;
    const int j_lo = min(n_kv, part * chunk), j_hi = min(n_kv, j_lo + chunk);
    const int qd = tid & 3;                  // pass 1: 4 threads per position; load i covers bytes [64 i + 16 qd, +16) of the row
    const int nld = hd / 32;                 // 16-byte loads per thread (hd 128 -> 4, hd 64 -> 2)
    const int dpl = hd / 32, nw = dpl / 2;   // pass 2: halves / words per lane
    constexpr int KP = MG_THREADS / 4;       // 64 positions per round
    constexpr int VP = 8;                    // V: local position = j0 + warp + 8u
    int4 kq[4];
    uint32_t vq[VP][2];
    // NOTE: the loads are UNCONDITIONAL on clamped (always valid) positions.  A load under `if (j < j_hi)` into a
    // register that keeps its old value otherwise is compiled as load + select, and the select waits for the load:
    // eight such loads in a row cost eight L2 round trips instead of one.
    const int j_last = max(j_hi - 1, 0);
    auto load_k = [&](int j0) {
        const int j = min(j0 + (tid >> 2), j_last);
        const __half *kr = mg_kv_row(k_pool, pbase, j, page_shift, kvd) + hk * hd + qd * 8;
#pragma unroll
        for (int i = 0; i < 4; i++) kq[i] = *(const int4 *)(kr + (i < nld ? i : 0) * 32); // unconditional, see the note above
    };
    auto load_v = [&](int j0) {
#pragma unroll
        for (int u = 0; u < VP; u++) {
            const int jv = min(j0 + warp + MG_WARPS * u, j_last);
            const __half *vr = mg_kv_row(v_pool, pbase, jv, page_shift, kvd) + hk * hd + lane * dpl;
            if (nw == 2) { const int2 t = *(const int2 *)vr; vq[u][0] = t.x; vq[u][1] = t.y; }
            else { vq[u][0] = *(const uint32_t *)vr; vq[u][1] = 0u; }
        }
    };
    load_k(j_lo);
    load_v(j_lo);
    for (int i = tid; i < hd; i += MG_THREADS) qs[i] = __half2float(__float2half_rn(ll_wait(q_ll + i, tag_in)));
    __syncthreads();

    // ---- 1./2. scores of my positions -> the head's score row
    for (int j0 = j_lo; j0 < j_hi; j0 += KP) {
        if (j0 > j_lo) load_k(j0);
        const int j = j0 + (tid >> 2);
        if (j == pos && j < j_hi) { // the current token: its K row comes as LL words (values already rounded to f16)
            float kc[4][8];
            unsigned spins = 0;
            while (true) { // all loads of a round are in flight together; a round trip per missing word would serialise
                bool ok = true;
#pragma unroll
                for (int i = 0; i < 4; i++)
                    if (i < nld) ok = ll_ld8(kcur_ll + hk * hd + i * 32 + qd * 8, tag_in, kc[i]) && ok;
                if (ok) break;
                ll_backoff(spins);
            }
#pragma unroll
            for (int i = 0; i < 4; i++)
                if (i < nld) {
                    __half2 h2[4];
#pragma unroll
                    for (int e = 0; e < 4; e++) h2[e] = __halves2half2(__float2half_rn(kc[i][2 * e]), __float2half_rn(kc[i][2 * e + 1]));
                    kq[i] = *(const int4 *)h2;
                }
        }
        float acc = 0.f;
        if (j < j_hi) {
            float a0 = 0.f, a1 = 0.f;
#pragma unroll
            for (int i = 0; i < 4; i++)
                if (i < nld) {
                    const __half2 *kh = (const __half2 *)&kq[i];
                    const float *qh = qs + i * 32 + qd * 8;
                    const float4 qa = *(const float4 *)qh, qb = *(const float4 *)(qh + 4);
                    const float2 k0 = __half22float2(kh[0]), k1 = __half22float2(kh[1]), k2 = __half22float2(kh[2]), k3 = __half22float2(kh[3]);
                    a0 += k0.x * qa.x; a1 += k0.y * qa.y; a0 += k1.x * qa.z; a1 += k1.y * qa.w;
                    a0 += k2.x * qb.x; a1 += k2.y * qb.y; a0 += k3.x * qb.z; a1 += k3.y * qb.w;
                }
            acc = a0 + a1;
        }
        acc += __shfl_xor_sync(0xffffffffu, acc, 1);
        acc += __shfl_xor_sync(0xffffffffu, acc, 2);
        if (j < j_hi && qd == 0) ll_st(g_scores + j, acc * scale, tag);
    }
    // ---- 3. soft-max over ALL positions of the head (redundantly in every part); the scores of the other parts arrive as LL words
    {
        unsigned spins = 0;
        while (true) { // every round requests all of this thread's words at once
            bool ok = true;
#pragma unroll 4
            for (int j = tid; j < n_kv; j += MG_THREADS) {
                const unsigned long long w = ll_ld(g_scores + j);
                if (ll_tag(w) == tag) sc[j] = ll_val(w);
                else ok = false;
            }
            if (ok) break;
            ll_backoff(spins);
        }
    }
    float lmax = -INFINITY;
    for (int j = tid; j < n_kv; j += MG_THREADS) lmax = fmaxf(lmax, sc[j]);
    lmax = warp_max(lmax);
    if (lane == 0) red[warp] = lmax;
    __syncthreads();
    float mx = red[0];
#pragma unroll
    for (int i = 1; i < MG_WARPS; i++) mx = fmaxf(mx, red[i]);
    float lsum = 0.f;
    for (int j = tid; j < n_kv; j += MG_THREADS) {
        const float e = expf(sc[j] - mx);
        sc[j] = e;
        lsum += e;
    }
    lsum = warp_sum(lsum);
    if (lane == 0) red[16 + warp] = lsum;
    __syncthreads();
    double tot = 0.0;
#pragma unroll
    for (int i = 0; i < MG_WARPS; i++) tot += (double)red[16 + i];
    const float inv = (float)(1.0 / tot);
    for (int j = j_lo + tid; j < j_hi; j += MG_THREADS) sc[j] = __half2float(__float2half_rn(__fmul_rn(sc[j], inv)));
    __syncthreads();
    // ---- 4. partial output over my positions
    float acc[4] = {0.f, 0.f, 0.f, 0.f};
    for (int j0 = j_lo; j0 < j_hi; j0 += VP * MG_WARPS) {
        if (j0 > j_lo) load_v(j0);
#pragma unroll
        for (int u = 0; u < VP; u++) {
            const int jv = j0 + warp + MG_WARPS * u;
            if (jv == pos && jv < j_hi) { // the current token's V row: LL words
                const uint2 *vp = vcur_ll + hk * hd + lane * dpl;
                unsigned long long w0, w1, w2 = 0, w3 = 0;
                unsigned spins = 0;
                while (true) {
                    ll_ld2(vp, w0, w1);
                    if (nw == 2) ll_ld2(vp + 2, w2, w3);
                    if (ll_tag(w0) == tag_in && ll_tag(w1) == tag_in && (nw != 2 || (ll_tag(w2) == tag_in && ll_tag(w3) == tag_in))) break;
                    ll_backoff(spins);
                }
                const __half2 h0 = __halves2half2(__float2half_rn(ll_val(w0)), __float2half_rn(ll_val(w1)));
                vq[u][0] = *(const uint32_t *)&h0;
                if (nw == 2) {
                    const __half2 h1 = __halves2half2(__float2half_rn(ll_val(w2)), __float2half_rn(ll_val(w3)));
                    vq[u][1] = *(const uint32_t *)&h1;
                }
            }
            if (jv < j_hi) {
                const float pj = sc[jv];
#pragma unroll
                for (int k = 0; k < 2; k++)
                    if (k < nw) {
                        const float2 vv = __half22float2(*(const __half2 *)&vq[u][k]);
                        acc[2 * k] += vv.x * pj;
                        acc[2 * k + 1] += vv.y * pj;
                    }
            }
        }
    }
#pragma unroll
    for (int i = 0; i < 4; i++)
        if (i < dpl) pw[warp * hd + lane * dpl + i] = acc[i];
    __syncthreads();
    for (int i = tid; i < hd; i += MG_THREADS) {
        float o = 0.f;
#pragma unroll
        for (int w = 0; w < MG_WARPS; w++) o += pw[w * hd + i];
        if (part != 0 || !combine_here) {
            ll_st(g_part + part * hd + i, o, tag); // (parts == 1 never gets here without combine_here)
        } else { // part 0 adds the partial rows in part order (fixed order: deterministic) and publishes the head's output
            float pv[3] = {0.f, 0.f, 0.f}; // parts <= 4
            unsigned spins = 0;
            while (true) {
                bool ok = true;
#pragma unroll
                for (int pp = 1; pp < 4; pp++)
                    if (pp < parts) {
                        const unsigned long long w = ll_ld(g_part + pp * hd + i);
                        ok = ok && ll_tag(w) == tag;
                        pv[pp - 1] = ll_val(w);
                    }
                if (ok) break;
                ll_backoff(spins);
            }
#pragma unroll
            for (int pp = 1; pp < 4; pp++)
                if (pp < parts) o += pv[pp - 1];
            ll_st(out_ll + i, o, tag);
        }
    }
    __syncthreads(); // the scratch is reused by the next phase's prologue
}

// The partial output rows of a head's parts -> the head's output, added in part order.  Run by the CTAs that have no share
// of the attention itself (148 CTAs, 128 shares for 32 heads x 4 parts): the CTAs that do the attention then all carry the
// same load, and none of them starts the next phase late because it also had to collect its peers' rows.
__device__ __noinline__ void mg_attention_combine(const uint2 *g_part, int parts, int hd, uint2 *out_ll, uint32_t tag, int tid) {
    for (int i = tid; i < hd; i += MG_THREADS) {
        float pv[4] = {0.f, 0.f, 0.f, 0.f};
        unsigned spins = 0;
        while (true) {
            bool ok = true;
#pragma unroll
            for (int pp = 0; pp < 4; pp++)
                if (pp < parts) {
                    const unsigned long long w = ll_ld(g_part + pp * hd + i);
                    ok = ok && ll_tag(w) == tag;
                    pv[pp] = ll_val(w);
                }
            if (ok) break;
            ll_backoff(spins);
        }
        float o = pv[0];
#pragma unroll
        for (int pp = 1; pp < 4; pp++)
            if (pp < parts) o += pv[pp];
        ll_st(out_ll + i, o, tag);
    }
}

// Pull the K/V rows (positions < n_kv) of one KV head into L2 ahead of the attention phase.
__device__ __forceinline__ void mg_prefetch_kv(int n_kv, const int *pbase, const __half *k_pool, const __half *v_pool, int page_shift,
                                               int kvd, int hk, int hd, int tid) {
    for (int j = tid; j < n_kv; j += MG_THREADS) {
        const __half *kr = mg_kv_row(k_pool, pbase, j, page_shift, kvd) + hk * hd;
        const __half *vr = mg_kv_row(v_pool, pbase, j, page_shift, kvd) + hk * hd;
        for (int b = 0; b < hd * 2; b += 128) {
            asm volatile("prefetch.global.L2 [%0];" ::"l"((const char *)kr + b));
            asm volatile("prefetch.global.L2 [%0];" ::"l"((const char *)vr + b));
        }
    }
}

// ------------------------------------------------------------------------------------------------ phase prologue
// 16-byte chunk c of the int8 activation section is stored at chunk c ^ ((c >> 3) & 7): the lanes of a quarter warp
// then read their register slices (stride 128 B = one bank group) from 8 different bank groups.
__device__ __forceinline__ int mg_swz(int c) { return c ^ ((c >> 3) & 7); }

// One warp quantises ONE 256-element block of a row to Q8_K and publishes it as LL words that hold the 32-bit words of
// the shared-memory activation row image (chunk-swizzled int8 | d | bsums): word w of the image -> out[w].  Used for
// activation rows so long (ffn_down: 56 blocks) that re-quantising them redundantly in every CTA costs more than one
// more hop: block b is quantised by CTA b as soon as its 256 inputs have arrived, then every CTA copies the 16 KB
// image.  Arithmetic = mg_prologue's.
__device__ __forceinline__ void mg_quant_block_ll(const uint2 *x_ll, uint32_t tag_in, int b, int64_t K, int nb, int lane, uint2 *out,
                                                   uint32_t tag_out) {
    float v[8];
    unsigned spins = 0;
    while (!__all_sync(0xffffffffu, ll_ld8(x_ll + b * 256 + lane * 8, tag_in, v))) ll_backoff(spins);
    float m = 0.f;
#pragma unroll
    for (int i = 0; i < 8; i++) m = fmaxf(m, fabsf(v[i]));
    const uint32_t amax = __reduce_max_sync(0xffffffffu, __float_as_uint(m));
    uint32_t li = 0xffffffffu;
#pragma unroll
    for (int i = 7; i >= 0; i--)
        if (__float_as_uint(fabsf(v[i])) == amax) li = ((uint32_t)(lane * 8 + i) << 1) | (__float_as_uint(v[i]) >> 31);
    const uint32_t key = __reduce_min_sync(0xffffffffu, li);
    const float mx = __uint_as_float(amax | (key << 31));
    const bool zero = amax == 0u;
    const float iscale = zero ? 0.f : __fdiv_rn(-127.f, mx);
    int sum = 0, qv[8];
#pragma unroll
    for (int i = 0; i < 8; i++) {
        int vi = __float2int_rn(__fmul_rn(iscale, v[i]));
        vi = min(127, vi);
        qv[i] = vi;
        sum += vi;
    }
    sum += __shfl_xor_sync(0xffffffffu, sum, 1); // 16-element sums, in both lanes of a pair
    const uint32_t p0 = (qv[0] & 0xFF) | ((qv[1] & 0xFF) << 8) | ((qv[2] & 0xFF) << 16) | ((uint32_t)(qv[3] & 0xFF) << 24);
    const uint32_t p1 = (qv[4] & 0xFF) | ((qv[5] & 0xFF) << 8) | ((qv[6] & 0xFF) << 16) | ((uint32_t)(qv[7] & 0xFF) << 24);
    const int w0 = (mg_swz(b * 16 + (lane >> 1)) * 16 + (lane & 1) * 8) >> 2; // image word of this lane's 8 bytes
    ll_st2(out + w0, __uint_as_float(p0), __uint_as_float(p1), tag_out);
    // bsums: int16 index b*16 + lane/2; two of them per image word: lanes 4k (low half) and 4k+2 (high half)
    const int hi = __shfl_down_sync(0xffffffffu, sum, 2);
    if ((lane & 3) == 0) {
        const uint32_t wv = ((uint32_t)sum & 0xFFFFu) | ((uint32_t)hi << 16);
        ll_st(out + ((K + nb * 4) >> 2) + b * 8 + (lane >> 2), __uint_as_float(wv), tag_out);
    }
    if (lane == 0) ll_st(out + (K >> 2) + b, zero ? 0.f : __fdiv_rn(1.f, iscale), tag_out);
}

// Copy the image of a quantised activation row (mg_quant_block_ll: LL words, one per 32-bit word of the image) into
// shared memory.  A light probe first -- the scale word of every block -- then 8 paired loads per thread in flight.
__device__ __noinline__ void mg_copy_xq(const uint2 *xq_ll, uint32_t tag, int64_t K, int nb, uint8_t *xs, int tid) {
    const int n_w = (int)((K + nb * 4 + K / 8) >> 2);
    unsigned spins = 0;
    while (true) {
        bool ok = true;
        for (int b = tid; b < nb; b += MG_THREADS) ok = ok && ll_tag(ll_ld(xq_ll + (K >> 2) + b)) == tag;
        if (__syncthreads_and(ok)) break;
        ll_backoff(spins);
    }
    for (int base = 0; base < n_w; base += 16 * MG_THREADS) {
        unsigned long long w[8][2];
        bool pending[8];
#pragma unroll
        for (int k = 0; k < 8; k++) {
            const int i0 = base + 2 * (tid + k * MG_THREADS);
            pending[k] = i0 < n_w;
            if (pending[k]) ll_ld2(xq_ll + i0, w[k][0], w[k][1]);
        }
#pragma unroll
        for (int k = 0; k < 8; k++) {
            const int i0 = base + 2 * (tid + k * MG_THREADS);
            if (!pending[k]) continue;
            while (ll_tag(w[k][0]) != tag || (i0 + 1 < n_w && ll_tag(w[k][1]) != tag)) {
                ll_backoff(spins);
                ll_ld2(xq_ll + i0, w[k][0], w[k][1]);
            }
            ((uint32_t *)xs)[i0] = (uint32_t)w[k][0];
            if (i0 + 1 < n_w) ((uint32_t *)xs)[i0 + 1] = (uint32_t)w[k][1];
        }
    }
    __syncthreads();
}

// (RMSNorm ->) quantise the activation row into shared memory, redundantly per CTA.  Written for LATENCY: warp w owns
// the 256-element blocks w, w+8, ... (PBN of them, all in registers), every global load is issued before the first
// use, invalid blocks are carried as zeros instead of branches so that the PBN dependency chains interleave, and the
// warp-wide maxima are hardware reductions (redux.sync) instead of shuffle trees.  Results are bit-exact with
// quantize_row_q8_K_ref / quantize_row_q8_0_ref (act_quant.cuh holds the same arithmetic for the other kernels).
template <int PBN>
__device__ __noinline__ void mg_prologue(const MegaPhase &ph, uint32_t tag_in, float eps, uint8_t *xs, double *s_part, int nb, int k_first,
                                         int warp, int lane, float *xn_out) {
    const bool write_xn = xn_out != nullptr;
    const int64_t K = ph.K;
    const bool has_norm = ph.norm_w != nullptr;
    float v[PBN][8], w[PBN <= 2 ? PBN : 1][8];
    bool valid[PBN];
#pragma unroll
    for (int k = 0; k < PBN; k++) {
        const int b = warp + (k_first + k) * MG_WARPS;
        valid[k] = b < nb;
        if (PBN <= 2 && has_norm) { // the norm weights are requested first: they do not depend on anybody
            const float4 *wb = (const float4 *)(ph.norm_w + (valid[k] ? b : 0) * 256 + lane * 8);
            const float4 w0 = wb[0], w1 = wb[1];
            w[k][0] = w0.x; w[k][1] = w0.y; w[k][2] = w0.z; w[k][3] = w0.w;
            w[k][4] = w1.x; w[k][5] = w1.y; w[k][6] = w1.z; w[k][7] = w1.w;
        }
    }
    if (ph.src_p < 0) { // plain fp32 row written before the launch
#pragma unroll
        for (int k = 0; k < PBN; k++) {
            const int b = warp + (k_first + k) * MG_WARPS;
            // Loads are unconditional (block index clamped): a load under a condition into a register that otherwise keeps
            // its value becomes load + select, and the select waits for the load -- the PBN round trips would serialise.
            const float4 *xb = (const float4 *)(ph.x + (valid[k] ? b : 0) * 256 + lane * 8);
            const float4 a0 = __ldcg(xb), a1 = __ldcg(xb + 1);
            v[k][0] = a0.x; v[k][1] = a0.y; v[k][2] = a0.z; v[k][3] = a0.w;
            v[k][4] = a1.x; v[k][5] = a1.y; v[k][6] = a1.z; v[k][7] = a1.w;
        }
    } else {
        // LL words written by phase src_p, possibly still on their way: every warp polls its own slice until all of it
        // carries the producing phase's tag (all loads of a round in flight together).  Measured alternatives: one warp
        // polling a sample of the row for the CTA 1.99 ms per token, every warp a sample of its slice first 1.85, one
        // thread polling a relaxed phase counter (B200LLM_LL_PROBE=2) 1.88, this 1.67.
        unsigned spins = 0;
        while (true) {
            bool ok = true;
#pragma unroll
            for (int k = 0; k < PBN; k++) {
                const int b = warp + (k_first + k) * MG_WARPS;
                ok = ll_ld8(ph.x_ll + (valid[k] ? b : 0) * 256 + lane * 8, tag_in, v[k]) && ok;
            }
            if (__all_sync(0xffffffffu, ok)) break;
            ll_backoff(spins);
        }
    }
#pragma unroll
    for (int k = 0; k < PBN; k++) // blocks past the row's end are carried as zeros
#pragma unroll
        for (int i = 0; i < 8; i++) v[k][i] = valid[k] ? v[k][i] : 0.f;
    if (has_norm) {
        double sum = 0.0;
#pragma unroll
        for (int k = 0; k < PBN; k++)
#pragma unroll
            for (int i = 0; i < 8; i++) sum += (double)__fmul_rn(v[k][i], v[k][i]);
        sum = warp_sum_d(sum);
        if (lane == 0) s_part[warp] = sum;
        __syncthreads();
        double tot = 0.0;
#pragma unroll
        for (int i = 0; i < MG_WARPS; i++) tot += s_part[i]; // every thread: same order, same result
        // K a power of two: tot * 2^-k is the correctly rounded quotient, without the division subroutine
        const float mean = (K & (K - 1)) == 0 ? (float)(tot * (1.0 / (double)K)) : (float)(tot / (double)K);
        const float scale = __fdiv_rn(1.0f, __fsqrt_rn(__fadd_rn(mean, eps)));
#pragma unroll
        for (int k = 0; k < PBN; k++) {
            const int b = warp + (k_first + k) * MG_WARPS;
            if (PBN > 2 && valid[k]) {
                const float4 *wb = (const float4 *)(ph.norm_w + b * 256 + lane * 8);
                const float4 w0 = wb[0], w1 = wb[1];
                const float wv[8] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w};
#pragma unroll
                for (int i = 0; i < 8; i++) v[k][i] = __fmul_rn(__fmul_rn(v[k][i], scale), wv[i]);
            } else if (PBN <= 2) {
#pragma unroll
                for (int i = 0; i < 8; i++) v[k][i] = __fmul_rn(__fmul_rn(v[k][i], scale), w[PBN <= 2 ? k : 0][i]);
            }
            if (write_xn && valid[k]) { // the normalised row itself (embeddings)
                float *o = xn_out + b * 256 + lane * 8;
                *(float4 *)o = make_float4(v[k][0], v[k][1], v[k][2], v[k][3]);
                *(float4 *)(o + 4) = make_float4(v[k][4], v[k][5], v[k][6], v[k][7]);
            }
        }
    }
    if (ph.fmt == ACT_Q8_K) {
        // [qs: K x int8 (chunk-swizzled)][d: K/256 x f32][bsums: K/16 x int16]
        constexpr int QB = PBN < 4 ? PBN : 4; // blocks quantised together (interleaved dependency chains)
#pragma unroll
        for (int k0 = 0; k0 < PBN; k0 += QB) {
            uint32_t amax[QB], key[QB];
#pragma unroll
            for (int kk = 0; kk < QB; kk++) {
                const int k = k0 + kk;
                float m = 0.f;
#pragma unroll
                for (int i = 0; i < 8; i++) m = fmaxf(m, fabsf(v[k][i]));
                amax[kk] = __reduce_max_sync(0xffffffffu, __float_as_uint(m)); // non-negative floats order like their bits
            }
#pragma unroll
            for (int kk = 0; kk < QB; kk++) { // the FIRST element (in index order) with the largest |x|: only its sign matters
                const int k = k0 + kk;
                uint32_t li = 0xffffffffu;
#pragma unroll
                for (int i = 7; i >= 0; i--)
                    if (__float_as_uint(fabsf(v[k][i])) == amax[kk]) li = ((uint32_t)(lane * 8 + i) << 1) | (__float_as_uint(v[k][i]) >> 31);
                key[kk] = __reduce_min_sync(0xffffffffu, li);
            }
#pragma unroll
            for (int kk = 0; kk < QB; kk++) {
                const int k = k0 + kk;
                const float mx = __uint_as_float(amax[kk] | (key[kk] << 31)); // +-amax: the value of that element
                const bool zero = amax[kk] == 0u;
                const float iscale = zero ? 0.f : __fdiv_rn(-127.f, mx);
                int sum = 0;
                int q[8];
#pragma unroll
                for (int i = 0; i < 8; i++) {
                    int vi = __float2int_rn(__fmul_rn(iscale, v[k][i])); // nearest_int(): round-half-even
                    vi = min(127, vi);
                    q[i] = vi;
                    sum += vi;
                }
                sum += __shfl_xor_sync(0xffffffffu, sum, 1);
                const int b = warp + (k_first + k) * MG_WARPS;
                if (valid[k]) {
                    int2 packed;
                    packed.x = (q[0] & 0xFF) | ((q[1] & 0xFF) << 8) | ((q[2] & 0xFF) << 16) | ((q[3] & 0xFF) << 24);
                    packed.y = (q[4] & 0xFF) | ((q[5] & 0xFF) << 8) | ((q[6] & 0xFF) << 16) | ((q[7] & 0xFF) << 24);
                    *(int2 *)(xs + mg_swz(b * 16 + (lane >> 1)) * 16 + (lane & 1) * 8) = packed;
                    if ((lane & 1) == 0) ((int16_t *)(xs + K + nb * 4))[b * 16 + (lane >> 1)] = (int16_t)sum;
                    if (lane == 0) ((float *)(xs + K))[b] = zero ? 0.f : __fdiv_rn(1.f, iscale);
                }
            }
        }
    } else {
        // ACT_Q8_0: [qs: K x int8 (chunk-swizzled)][d: K/32 x f32 (value of the fp16-rounded scale)]; 32 elements = 4 lanes
#pragma unroll
        for (int k = 0; k < PBN; k++) {
            float amax = 0.f;
#pragma unroll
            for (int i = 0; i < 8; i++) amax = fmaxf(amax, fabsf(v[k][i]));
            amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 1));
            amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, 2));
            const float d = __fdiv_rn(amax, 127.f);
            const float id = d != 0.f ? __fdiv_rn(1.f, d) : 0.f;
            int q[8];
#pragma unroll
            for (int i = 0; i < 8; i++) q[i] = (int)roundf(__fmul_rn(v[k][i], id));
            const int b = warp + (k_first + k) * MG_WARPS;
            if (valid[k]) {
                int2 packed;
                packed.x = (q[0] & 0xFF) | ((q[1] & 0xFF) << 8) | ((q[2] & 0xFF) << 16) | ((q[3] & 0xFF) << 24);
                packed.y = (q[4] & 0xFF) | ((q[5] & 0xFF) << 8) | ((q[6] & 0xFF) << 16) | ((q[7] & 0xFF) << 24);
                *(int2 *)(xs + mg_swz(b * 16 + (lane >> 1)) * 16 + (lane & 1) * 8) = packed;
                if ((lane & 3) == 0) ((float *)(xs + K))[b * 8 + (lane >> 2)] = __half2float(__float2half_rn(d));
            }
        }
    }
    __syncthreads();
}

// Tensor parallelism: the input row of a phase is the sum of reduce_n partial rows, one from every shard (they arrive in
// this GPU's memory as LL words stored by the peers), plus the residual.  Every CTA assembles a slice of it -- the partial
// rows are added in rank order, so all shards compute the same bits -- and publishes the slice in x_ll, with the tag the
// phase's prologue waits for.
__device__ __noinline__ void mg_reduce_partials(const MegaPhase &ph, uint32_t tag_in, int cta, int n_cta, int tid) {
    const int per = (ph.reduce_len + n_cta - 1) / n_cta;
    for (int k = tid; k < per; k += MG_THREADS) {
        const int e = cta * per + k;
        if (e >= ph.reduce_len) break;
        float pv[8];
        unsigned spins = 0;
        while (true) {
            bool ok = true;
#pragma unroll
            for (int j = 0; j < 8; j++)
                if (j < ph.reduce_n) {
                    const unsigned long long w = ll_ld_sys(ph.part_ll + (size_t)j * ph.reduce_len + e);
                    ok = ok && ll_tag(w) == tag_in;
                    pv[j] = ll_val(w);
                }
            if (ok) break;
            ll_backoff(spins);
        }
        float sum = ph.red_res_ll ? ll_val(ll_ld(ph.red_res_ll + e)) : ph.red_residual ? __ldcg(ph.red_residual + e) : 0.f;
#pragma unroll
        for (int j = 0; j < 8; j++)
            if (j < ph.reduce_n) sum += pv[j];
        ll_st(ph.x_ll_w + e, sum, tag_in);
    }
}

// ---- epilogue of a mat-vec phase: combine the K-split partials in a fixed order (deterministic), then the phase's tail op.
// Outputs that a later phase consumes leave as LL words tagged with this phase: no fence and no barrier follow.
__device__ __noinline__ void mg_epilogue(const MegaPhase &ph, const MegaArgs &a, const float *partial, const int *pbase, int u, int u_end, int pos,
                                         uint32_t tag_out, float res_pre, int tid) {
    const int S = ph.S, n_r = u_end - u;
    auto dot_sum = [&](int dr) {
        float s = 0.f;
        for (int sg = 0; sg < S; sg++) s += partial[dr * S + sg];
        return s;
    };
    if (ph.kind == PH_SWIGLU) {
        for (int r = tid; r < n_r; r += MG_THREADS) {
            const float g = dot_sum(r), up = dot_sum(n_r + r);
            ll_st(ph.out_ll + u + r, __fmul_rn(__fdiv_rn(g, __fadd_rn(1.0f, expf(-g))), up), tag_out);
        }
    } else if (ph.kind == PH_PLAIN && ph.n_peers > 0) {
        // a partial row of a tensor-parallel model: one plain 64-bit store per shard, this one included
        // (measured and dropped: a system fence after the stores +0.2 ms per token, 16-byte stores and polling with an
        // atomic no change)
        for (int r = tid; r < n_r; r += MG_THREADS) {
            const float sv = dot_sum(r);
#pragma unroll
            for (int j = 0; j < 8; j++)
                if (j < ph.n_peers) ll_st_sys(ph.peer_part[j] + u + r, sv, tag_out);
        }
    } else if (ph.kind == PH_PLAIN) {
        float *outp = a.out_override && !ph.out_ll ? a.out_override : ph.out;
        for (int r = tid; r < n_r; r += MG_THREADS) {
            float s = dot_sum(r);
            if (r == tid) s += res_pre; // requested before the tile loop
            else if (ph.res_ll) s += ll_val(ll_ld(ph.res_ll + u + r));
            else if (ph.residual) s += __ldcg(ph.residual + u + r);
            if (ph.out_ll) ll_st(ph.out_ll + u + r, s, tag_out);
            else outp[u + r] = s;
        }
    } else { // PH_QKV: job 0 = q, 1 = k, 2 = v; rows (r, r+1) form a rotary pair inside one head
        const int64_t cell = ((((int64_t)pbase[pos >> a.page_shift]) << a.page_shift) + (pos & (a.page_size - 1))) * a.kvd;
        for (int pr = tid; pr < n_r / 2; pr += MG_THREADS) {
            int jb, r;
            mg_locate_row(ph, u + 2 * pr, jb, r);
            const float v0 = dot_sum(2 * pr), v1 = dot_sum(2 * pr + 1);
            if (jb == 2) { // V: into the cache for later tokens, and as LL words for this token's attention
                const __half2 hv = __halves2half2(__float2half_rn(v0), __float2half_rn(v1));
                *(__half2 *)(ph.v_pool + cell + r) = hv;
                ll_st2(a.vcur_ll + r, __low2float(hv), __high2float(hv), tag_out);
            } else {
                const float *cs = a.rope_table + (int64_t)pos * a.hd + (r % a.hd);
                const float c = cs[0], sn = cs[1];
                const float o0 = __fsub_rn(__fmul_rn(v0, c), __fmul_rn(v1, sn));
                const float o1 = __fadd_rn(__fmul_rn(v0, sn), __fmul_rn(v1, c));
                if (jb == 0) {
                    ll_st2(ph.out_ll + r, o0, o1, tag_out);
                } else {
                    const __half2 hk2 = __halves2half2(__float2half_rn(o0), __float2half_rn(o1));
                    *(__half2 *)(ph.k_pool + cell + r) = hk2;
                    ll_st2(a.kcur_ll + r, __low2float(hk2), __high2float(hk2), tag_out);
                }
            }
        }
    }
}

// ------------------------------------------------------------------------------------------------ the kernel
// Tile header, the first 16 bytes of every tile of a warp's stream (written by MegaStreams::build):
//   w0: ggml type | rows << 8 | row pitch << 16        w1: first row's index in the phase's partial-sum rows
//   w2: bytes of the tile TWO ahead in this stream (0: none) -- what the freed stage is refilled with
//   w3: bit 0: the warp's last tile of this phase; bits 8..: the next phase in which the warp has tiles
template <bool DBG>
__global__ void __launch_bounds__(MG_THREADS, 1)
decode_mega_kernel(const __grid_constant__ MegaArgs a) {
    extern __shared__ __align__(128) uint8_t smem[];
    const MgSmem lay = mg_smem_layout(a.geo);
    const uint32_t stage_bytes = (uint32_t)a.geo.stage_bytes;
    uint8_t *xs = smem + lay.xs;
    float *partial = (float *)(smem + lay.part);
    double *s_part = (double *)(smem + lay.red);
    uint64_t *bars = (uint64_t *)(smem + lay.bar);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int cta = blockIdx.x, n_cta = gridDim.x;

    MegaPhase *sph = (MegaPhase *)(smem + lay.ph);
    constexpr int PH_WORDS = (int)(sizeof(MegaPhase) / 4);
    static_assert(PH_WORDS <= MG_THREADS, "phase descriptor copy");
    static_assert(MG_PH_SLOT % 16 == 0, "phase slots");
    // Phase p's descriptor sits in slot p % 3.  It is copied in during phase p - 2, right after that phase's mid-phase
    // __syncthreads (every thread is then past phase p - 3, the slot's previous tenant), and first read after phase
    // p - 1's mid-phase __syncthreads.
    auto stage_phase = [&](int p) {
        if (p < a.n_phases && tid < PH_WORDS) ((uint32_t *)((uint8_t *)sph + (p % MG_PH_WIN) * MG_PH_SLOT))[tid] = ((const uint32_t *)&a.phases[p])[tid];
    };
    stage_phase(0);
    stage_phase(1);
    const int pos = a.pos[0], sl = a.slot[0];
    int *pbase = (int *)(smem + lay.pt); // page of every 32-position block of this sequence (static during the step)
    for (int i = tid; a.page_table && i < (pos + a.page_size) / a.page_size && i < MG_MAX_PAGES; i += MG_THREADS) pbase[i] = a.page_table[(int64_t)sl * a.max_pages + i];
    if (tid == 0) {
        for (int i = 0; i < MG_WARPS * MG_STAGES; i++) mg_mbar_init(bars + i, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    // ---- this warp's ring and its weight stream: the first two tiles are requested right away
    const uint32_t ring = mg_s32(smem) + warp * (MG_STAGES * stage_bytes);
    const uint32_t bar0 = mg_s32(bars + warp * MG_STAGES);
    const uint8_t *src; // next tile to request into the ring
    int next_p; // the next phase in which this warp has tiles
    {
        const MegaStreamStart st = a.starts[cta * MG_WARPS + warp];
        src = st.src;
        next_p = st.first_p;
        if (lane == 0) {
            if (st.size0) { mg_expect_tx(bar0, st.size0); mg_bulk(ring, src, st.size0, bar0); }
            if (st.size1) { mg_expect_tx(bar0 + 8, st.size1); mg_bulk(ring + stage_bytes, src + st.size0, st.size1, bar0 + 8); }
        }
        src += st.size0 + st.size1;
    }
    uint32_t stg = 0, parity = 0; // ring position of the next tile to consume

    for (int p = 0; p < a.n_phases; p++) {
        const MegaPhase &ph = *(const MegaPhase *)((const uint8_t *)sph + (p % MG_PH_WIN) * MG_PH_SLOT);
        const uint32_t tag_in = a.tag_base + (uint32_t)ph.src_p + 1u, tag_out = a.tag_base + (uint32_t)p + 1u;
        long long t0 = 0, t1 = 0, t2 = 0, t_wait = 0, t_comp = 0, t_epi = 0, n_tiles = 0, tp0 = 0, tp3 = 0, tl_end = 0;
        long long g0 = 0, g1 = 0, g2 = 0;
        if (DBG && tid == 0) { t0 = clock64(); g0 = mg_gtime(); }
        int u = 0, u_end = 0;
        if (ph.kind == PH_ATT) {
            // a head is shared by `parts` CTAs (att_parts * n_head <= n_cta whenever att_parts > 1): att_parts for long
            // contexts, 2 below 192 positions, where the second split costs more in exchanges than it saves in rows
            // (Llama-3-8B: 1.658 against 1.672 ms per token at 128 positions, 1.82 against 1.74 at 512; every CTA reads
            // the same position, so all of them take the same branch)
            const int parts = (a.att_parts > 2 && pos + 1 < 192) ? 2 : a.att_parts, parts_log = parts > 2 ? a.att_parts_log : (parts == 2 ? 1 : 0);
            const int n_att = a.n_head * parts, n_comb = parts > 1 ? n_cta - n_att : 0; // CTAs left over collect the parts' rows
            for (int hp = cta; hp < n_att; hp += n_cta) {
                const int h = hp >> parts_log, part = hp & (parts - 1), hk = h / (a.n_head / a.n_head_kv);
                mg_attention_ll(ph.x_ll + (int64_t)h * a.hd, a.kcur_ll, a.vcur_ll, tag_in, pos + 1, pbase, ph.k_pool, ph.v_pool, a.page_shift,
                                a.kvd, hk, a.hd, a.att_scale, ph.out_ll + (int64_t)h * a.hd, (float *)xs, tid, part, parts,
                                a.att_scores + (int64_t)h * a.att_scores_stride, a.att_part_out + (int64_t)h * parts * a.hd, tag_out, n_comb <= 0);
            }
            if (n_comb > 0 && cta >= n_att)
                for (int h = cta - n_att; h < a.n_head; h += n_comb)
                    mg_attention_combine(a.att_part_out + (int64_t)h * parts * a.hd, parts, a.hd, ph.out_ll + (int64_t)h * a.hd, tag_out, tid);
            __syncthreads(); // the mid-phase sync of a phase without rows
            stage_phase(p + 2);
        } else {
            mg_cta_range(ph, cta, n_cta, u, u_end);
            if (u >= u_end) {
                __syncthreads();
                stage_phase(p + 2);
            }
        }
        if (a.ll_probe == 2 && ph.kind != PH_ATT && ph.src_p >= 0 && u < u_end) {
            // wait quietly until every CTA has stored the outputs of the producing phase (see mg_hint_arrive)
            if (tid == 0) mg_hint_wait(a.hint, a.hint_base + ((unsigned)ph.src_p + 1u) * (unsigned)n_cta);
            __syncthreads();
        }
        if (ph.kind != PH_ATT && ph.reduce_n > 0) mg_reduce_partials(ph, tag_in, cta, n_cta, tid);
        if (ph.kind != PH_ATT && ph.xq_ll) {
            // A long activation row is quantised ONCE per GPU (block b by warp 0 of CTA b, as soon as its inputs are
            // there), not once per CTA; the prologue below then copies the 16 KB image.
            const int nbq = (int)(ph.K >> 8);
            if (warp == 0)
                for (int b = cta; b < nbq; b += n_cta) mg_quant_block_ll(ph.x_ll, tag_in, b, ph.K, nbq, lane, ph.xq_ll, tag_out);
        }
        if (u < u_end) {
            const int64_t K = ph.K;
            const int nb = (int)(K >> 8);
            const int S = ph.S, rpp = ph.rpp, np = ph.np;
            // ---- prologue: (RMSNorm ->) quantise the activation row into shared memory
            if (DBG && tid == 0) tp0 = clock64();
            if (ph.xq_ll) { // quantised once per GPU (above): copy the image of the activation row
                mg_copy_xq(ph.xq_ll, tag_out, K, nb, xs, tid);
            } else {
                float *wx = ph.xn_out && cta == 0 ? (a.xn_override ? a.xn_override : ph.xn_out) : nullptr;
                if (nb <= MG_WARPS) mg_prologue<1>(ph, tag_in, a.eps, xs, s_part, nb, 0, warp, lane, wx);
                else if (nb <= 2 * MG_WARPS) mg_prologue<2>(ph, tag_in, a.eps, xs, s_part, nb, 0, warp, lane, wx);
                else // 4 blocks per warp at a time; rows longer than 32 blocks have no norm (mega_phase_prepare)
                    for (int k0 = 0; k0 * MG_WARPS < nb; k0 += 4) mg_prologue<4>(ph, tag_in, a.eps, xs, s_part, nb, k0, warp, lane, wx);
            }
            if (DBG && tid == 0) tp3 = clock64();
            if (ph.kind == PH_QKV) // the attention phase comes next: start its K/V rows on their way into L2
                for (int h = cta; h < a.n_head; h += n_cta)
                    mg_prefetch_kv(pos, pbase, ph.k_pool, ph.v_pool, a.page_shift, a.kvd, h / (a.n_head / a.n_head_kv), a.hd, tid);
            // ---- this lane's slice of the quantised row -> registers, for the whole phase
            const int q = warp & (S - 1);
            const uint32_t xs_a = mg_s32(xs);
            int xr[32];
            float dx = 0.f;
            int bs[8];
            MgLane L;
            if (ph.fmt == ACT_Q8_K) {
                const int lpr = ph.lpr;
                L.sr = rpp > 1 ? lane >> ph.lpr_log : 0;
                const int it = lane - L.sr * lpr;
                L.on = it < lpr && L.sr < rpp;
                L.kb = L.on ? it >> 1 : 0;
                L.h = it & 1;
                L.head = L.on && it == 0;
                L.span = rpp > 1 ? lpr : 32;
                const int blk = q * (nb >> ph.s_log) + L.kb;
                const int c0 = blk * 16 + 8 * L.h; // first 16-byte chunk of this lane's 128 activations
#pragma unroll
                for (int i = 0; i < 8; i++) {
                    const int4 t = mg_lds128(xs_a + mg_swz(c0 + i) * 16);
                    xr[4 * i] = t.x; xr[4 * i + 1] = t.y; xr[4 * i + 2] = t.z; xr[4 * i + 3] = t.w;
                }
                dx = ((const float *)(xs + K))[blk];
                const int *bsp = (const int *)(xs + K + nb * 4 + (blk * 16 + 8 * L.h) * 2); // 8 x int16
#pragma unroll
                for (int i = 0; i < 4; i++) {
                    const int w2 = bsp[i];
                    bs[2 * i] = (int)(int16_t)(w2 & 0xFFFF);
                    bs[2 * i + 1] = w2 >> 16;
                }
            } else { // ACT_Q8_0 (1 byte per weight): the activations are read from shared memory next to the weights
                L.sr = 0; L.kb = 0; L.h = 0; L.on = true; L.head = lane == 0; L.span = 32;
            }
            const int bs32[4] = {bs[0] + bs[1], bs[2] + bs[3], bs[4] + bs[5], bs[6] + bs[7]};
            const uint32_t q6d = (uint32_t)((nb >> ph.s_log) * Q6_BLK + 2 * L.kb); // Q6_K: this lane's fp16 scale inside a row-split
            // the residual of this thread's first output row is requested now, not in the epilogue (a round trip under load)
            float res_pre = 0.f;
            if (ph.kind == PH_PLAIN && ph.n_peers == 0 && tid < u_end - u) {
                if (ph.res_ll) res_pre = ll_val(ll_ld(ph.res_ll + u + tid));
                else if (ph.residual) res_pre = __ldcg(ph.residual + u + tid);
            }
            if (DBG && tid == 0) { t1 = clock64(); g1 = mg_gtime(); }

            // ---- this warp's tiles of the phase
            if (next_p == p) {
                while (true) {
                    const uint32_t st = ring + stg * stage_bytes, bar = bar0 + stg * 8;
                    long long tw0 = 0, tw1 = 0;
                    if (DBG && tid == 0) tw0 = clock64();
                    mg_wait(bar, parity);
                    if (DBG && tid == 0) tw1 = clock64();
                    const int4 th = mg_lds128(st);
                    const int type = th.x & 0xFF, n = (th.x >> 8) & 0xFF, rb = (int)((uint32_t)th.x >> 16);
                    const uint32_t td = st + MG_TILE_HDR;
                    float *pout = partial + (size_t)th.y * S + q;
                    if (type == B2_T_Q4_K) {
                        if (rpp == 1) { // 2 rows at a time: independent dp4a chains, one merged reduction
                            const uint32_t base = td + L.kb * 144;
#pragma unroll 1
                            for (int r0 = 0; r0 < n; r0 += 2) {
                                float v0 = mg_dot_q4k(base + r0 * rb, L.h, xr, dx, bs32);
                                float v1 = mg_dot_q4k(base + min(r0 + 1, n - 1) * rb, L.h, xr, dx, bs32);
                                if (!L.on) { v0 = 0.f; v1 = 0.f; }
                                const float tot = mg_reduce2(v0, v1, lane);
                                const int kk = r0 + ((lane >> 4) & 1);
                                if ((lane & 15) == 0 && kk < n) pout[(size_t)kk * S] = tot;
                            }
                        } else {
#pragma unroll 1
                            for (int rr = 0; rr < n; rr += rpp) {
                                const int row = rr + L.sr;
                                const bool on = L.on && row < n;
                                float v = mg_dot_q4k(td + (on ? row : 0) * rb + L.kb * 144, L.h, xr, dx, bs32);
                                v = mg_row_sum(on ? v : 0.f, L.span);
                                if (on && L.head) pout[(size_t)row * S] = v;
                            }
                        }
                    } else if (type == B2_T_Q6_K) {
                        // row-split = [208-byte blocks][their fp16 super-block scales]
                        if (rpp == 1) { // 2 rows at a time
                            const uint32_t base = td + L.kb * Q6_BLK;
#pragma unroll 1
                            for (int k = 0; k < n; k += 2) {
                                const uint32_t o0 = k * rb, o1 = min(k + 1, n - 1) * rb;
                                const uint32_t d0 = mg_lds_u16(td + o0 + q6d), d1 = mg_lds_u16(td + o1 + q6d);
                                float v0 = mg_dot_q6k(base + o0, L.h, xr, dx, bs, d0);
                                float v1 = mg_dot_q6k(base + o1, L.h, xr, dx, bs, d1);
                                if (!L.on) { v0 = 0.f; v1 = 0.f; }
                                const float tot = mg_reduce2(v0, v1, lane);
                                const int kk = k + ((lane >> 4) & 1);
                                if ((lane & 15) == 0 && kk < n) pout[(size_t)kk * S] = tot;
                            }
                        } else {
#pragma unroll 1
                            for (int rr = 0; rr < n; rr += rpp) {
                                const int row = rr + L.sr;
                                const bool on = L.on && row < n;
                                const uint32_t o = (on ? row : 0) * rb;
                                const uint32_t d16 = mg_lds_u16(td + o + q6d);
                                float v = mg_dot_q6k(td + o + L.kb * Q6_BLK, L.h, xr, dx, bs, d16);
                                v = mg_row_sum(on ? v : 0.f, L.span);
                                if (on && L.head) pout[(size_t)row * S] = v;
                            }
                        }
                    } else if (type == B2_T_Q5_K) {
#pragma unroll 1
                        for (int rr = 0; rr < n; rr += rpp) {
                            const int row = rr + L.sr;
                            const bool on = L.on && row < n;
                            float v = mg_dot_q5k(td + (on ? row : 0) * rb + L.kb * 176, L.h, xr, dx, bs32);
                            v = mg_row_sum(on ? v : 0.f, L.span);
                            if (on && L.head) pout[(size_t)row * S] = v;
                        }
                    } else { // Q8_0: row-split = [int8 K/S][fp16 scales K/S/32]
                        const int ks = (int)(K >> ph.s_log);
                        const int xc = (q * ks) >> 4;                               // first 16-byte chunk of this K-split's activations
                        const float *xdq = (const float *)(xs + K) + (q * ks) / 32; // and their block scales
#pragma unroll 1
                        for (int row = 0; row < n; row += 2) { // 2 rows at a time: independent chains, one merged reduction
                            const uint32_t rp0 = td + row * rb, rp1 = td + min(row + 1, n - 1) * rb;
                            float acc0 = 0.f, acc1 = 0.f;
#pragma unroll 4
                            for (int j = 0; j < np; j++) { // 16-byte pieces lane + 32 j: half a 32-element block each
                                const int piece = lane + 32 * j;
                                const int4 xv = mg_lds128(xs_a + mg_swz(xc + piece) * 16);
                                const int4 w0 = mg_lds128(rp0 + piece * 16), w1 = mg_lds128(rp1 + piece * 16);
                                const float dxj = xdq[piece >> 1];
                                const float d0 = h2f((uint16_t)mg_lds_u16(rp0 + ks + (piece >> 1) * 2)) * dxj;
                                const float d1 = h2f((uint16_t)mg_lds_u16(rp1 + ks + (piece >> 1) * 2)) * dxj;
                                int s0 = __dp4a(w0.x, xv.x, 0), s1 = __dp4a(w1.x, xv.x, 0);
                                s0 = __dp4a(w0.y, xv.y, s0); s1 = __dp4a(w1.y, xv.y, s1);
                                s0 = __dp4a(w0.z, xv.z, s0); s1 = __dp4a(w1.z, xv.z, s1);
                                s0 = __dp4a(w0.w, xv.w, s0); s1 = __dp4a(w1.w, xv.w, s1);
                                // the two halves of a block are scaled separately ((a + b) d = a d + b d up to one rounding)
                                acc0 += (float)s0 * d0;
                                acc1 += (float)s1 * d1;
                            }
                            const float tot = mg_reduce2(acc0, acc1, lane);
                            const int kk = row + ((lane >> 4) & 1);
                            if ((lane & 15) == 0 && kk < n) pout[(size_t)kk * S] = tot;
                        }
                    }
                    if (DBG && tid == 0) { t_wait += tw1 - tw0; t_comp += clock64() - tw1; n_tiles++; }
                    // the stage is free: refill it with the tile MG_STAGES ahead in the stream (possibly a later phase's)
                    __syncwarp();
                    if (th.z) {
                        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                        if (lane == 0) {
                            mg_expect_tx(bar, (uint32_t)th.z);
                            mg_bulk(st, src, (uint32_t)th.z, bar);
                        }
                        src += (uint32_t)th.z;
                    }
                    if (++stg == MG_STAGES) { stg = 0; parity ^= 1; }
                    if (th.w & 1) {
                        next_p = (int)((uint32_t)th.w >> 8);
                        // this warp is done with the phase and its ring holds the next two tiles: what it would idle through
                        // (the hand-over) HBM can spend on pulling the tiles after those into L2
                        if (a.wait_prefetch && lane == 0 && th.z) mg_prefetch_l2(src, (uint32_t)a.wait_prefetch);
                        break;
                    }
                }
            }
            if (DBG && tid == 0) tl_end = clock64();
            __syncthreads(); // the mid-phase sync: every warp's partial sums are in shared memory
            stage_phase(p + 2);
            if (DBG && tid == 0) { t_epi = clock64(); g2 = mg_gtime(); }

            mg_epilogue(ph, a, partial, pbase, u, u_end, pos, tag_out, res_pre, tid);
        }
        if (a.ll_probe == 2 && p + 1 < a.n_phases) { // this CTA's outputs of phase p are stored (or it had none)
            __syncthreads();
            if (tid == 0) mg_hint_arrive(a.hint);
        }
        if (DBG && tid == 0) {
            t2 = clock64();
            // per (CTA, phase): [0..3] global-timer ns at phase start / input row ready (prologue + slice done) / this warp past the
            // mid-phase sync / phase end; [4..7] warp 0's tile wait cycles, tile compute cycles, tiles, epilogue cycles
            long long *dd = a.dbg + ((int64_t)cta * a.n_phases + p) * 16;
            dd[0] = g0; dd[1] = g1 ? g1 : g0; dd[2] = g2 ? g2 : (g1 ? g1 : g0); dd[3] = mg_gtime();
            dd[4] = t_wait; dd[5] = t_comp; dd[6] = n_tiles; dd[7] = t_epi ? t2 - t_epi : 0;
            dd[8] = tp0 ? tp0 - t0 : 0; dd[9] = tp3 ? tp3 - tp0 : 0; dd[10] = t1 ? t1 - tp3 : 0; dd[11] = tl_end ? tl_end - t1 : 0; dd[12] = u_end - u;
        }
    }
}

void mega_fill_knobs(MegaArgs &a) { // read at every launch on purpose: the tests flip them inside one process
    const char *ns = getenv("B200LLM_LL_POLL_NS"), *pr = getenv("B200LLM_LL_PROBE");
    a.ll_poll_ns = ns ? atoi(ns) : 32;
    a.ll_probe = pr ? atoi(pr) : 0; // measured on Llama-3-8B: 0 -> 1.67 ms per token, 2 -> 1.88
    const char *wp = getenv("B200LLM_MEGA_WAIT_PREFETCH");
    a.wait_prefetch = wp ? std::min(64 << 10, std::max(0, atoi(wp))) / 16 * 16 : 0;
}

void launch_decode_mega(const MegaArgs &a, int n_sm, cudaStream_t s) {
    const MgSmem lay = mg_smem_layout(a.geo);
    B2_CHECK(lay.total <= (uint32_t)MG_SMEM_MAX && a.geo.stage_bytes % 16 == 0 && a.geo.xs_bytes % 16 == 0, B2_EINVAL, "decode kernel: shared-memory geometry");
    static PerDeviceOnce configured;
    configured.run([&] {
        B2_CUDA(cudaFuncSetAttribute(decode_mega_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, MG_SMEM_MAX));
        B2_CUDA(cudaFuncSetAttribute(decode_mega_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, MG_SMEM_MAX));
        const unsigned v = (unsigned)a.ll_poll_ns;
        B2_CUDA(cudaMemcpyToSymbol(g_ll_poll_ns, &v, sizeof(v)));
    });
    void *args[] = {(void *)&a};
    const void *fn = a.dbg ? (const void *)decode_mega_kernel<true> : (const void *)decode_mega_kernel<false>;
    B2_CUDA(cudaLaunchCooperativeKernel(fn, dim3(n_sm), dim3(MG_THREADS), args, lay.total, s));
    g_launch_count++;
}

// Shared memory per model: the activation row takes what the longest row needs (or the attention scratch of the longest
// context, whichever is larger), the 16 ring stages share the rest.
MegaGeo mega_geometry(const MegaPhase *phases, int n_phases, int n_ctx_slot, int hd) {
    MegaGeo g;
    int64_t row = 0;
    for (int p = 0; p < n_phases; p++)
        if (phases[p].kind != PH_ATT) row = std::max(row, mg_act_row_bytes_host(phases[p].fmt, phases[p].K));
    const int64_t att = (int64_t)sizeof(float) * ((int64_t)hd + 32 + (int64_t)MG_WARPS * hd + 16 + n_ctx_slot + 8) - (int64_t)MG_MAX_PART * (int64_t)sizeof(float);
    int64_t xs = std::max<int64_t>(std::max(row, att), 4096);
    xs = (xs + 255) / 256 * 256;
    g.xs_bytes = (int)xs;
    MegaGeo probe = g;
    probe.stage_bytes = 0;
    const int64_t fixed = mg_smem_layout(probe).total;
    g.stage_bytes = (int)((MG_SMEM_MAX - fixed) / (MG_WARPS * MG_STAGES) / 16 * 16);
    if (g.stage_bytes > 16384 + MG_TILE_HDR) g.stage_bytes = 16384 + MG_TILE_HDR;
    return g;
}

bool mega_phase_prepare(MegaPhase &ph, int n_cta, const MegaGeo &geo) {
    if (ph.kind == PH_ATT) return true;
    if (n_cta < 1) return false;
    ph.units_q = ph.units / n_cta;
    ph.units_r = ph.units % n_cta;
    int S, lpr, rpp, np;
    if (!mg_split(ph.fmt, ph.K, ph.xq_ll != nullptr, S, lpr, rpp, np)) return false;
    if (mg_act_row_bytes_host(ph.fmt, ph.K) > geo.xs_bytes || geo.stage_bytes <= MG_TILE_HDR) return false;
    if (ph.norm_w && ph.K / 256 > 4 * MG_WARPS) return false; // the prologue norms at most 4 blocks per warp in one go
    ph.S = S; ph.lpr = lpr; ph.rpp = rpp; ph.np = np;
    ph.s_log = 0;
    while ((1 << ph.s_log) < S) ph.s_log++;
    ph.lpr_log = 0;
    while ((1 << ph.lpr_log) < lpr) ph.lpr_log++; // only used when lpr is a power of two (rpp > 1)
    int64_t rows = 0;
    for (int j = 0; j < ph.n_jobs; j++) {
        ph.j0[j] = (int)rows;
        if (act_format_for(ph.job[j].type) != ph.fmt) return false;
        if (!mg_job_geo(ph.job[j].type, ph.K, S, rpp, geo.stage_bytes - MG_TILE_HDR, ph.rb[j], ph.R[j])) return false;
        if (ph.job[j].rows % 2 != 0 || ph.job[j].rows >= (1 << 22)) return false;
        rows += ph.job[j].rows;
    }
    // the per-CTA share of the rows x K-splits must fit the partial buffer
    return (rows / n_cta + 4) * S <= MG_MAX_PART;
}
bool mega_phase_supported(const MegaPhase &ph) {
    MegaPhase t = ph;
    return mega_phase_prepare(t, 128);
}

size_t mega_attention_ctx_limit(int hd, const MegaGeo &geo) {
    // q + reduction scratch + per-warp partials + scores live in the activation-row and partial regions; the page
    // table of the sequence has MG_MAX_PAGES entries (page size >= 16 assumed by the caller)
    const size_t avail = (size_t)geo.xs_bytes + MG_MAX_PART * sizeof(float);
    const size_t fixed = sizeof(float) * ((size_t)hd + 32 + (size_t)MG_WARPS * hd + 16);
    if (fixed >= avail) return 0;
    return std::min((size_t)((avail - fixed) / 4) - 8, (size_t)MG_MAX_PAGES * 16);
}

// ------------------------------------------------------------------------------------------------ weight streams
// One entry per tile: where its row-splits come from (GGUF-layout rows on the device) and where the tile goes.
struct MgTileSrc {
    const uint8_t *src0, *src1; // first row's two pieces (piece 1: Q6_K fp16 super-block scales / Q8_0 block scales)
    uint64_t dst;               // byte offset of the tile in the stream buffer
    uint32_t hdr[4];
    int32_t stride;             // source row stride
    int32_t len0, len1;         // bytes of the pieces (len0 a multiple of 16, len1 of 2)
    int32_t rb, n, pad;
};
static_assert(sizeof(MgTileSrc) == 64, "tile table entry");

__global__ void __launch_bounds__(256) mega_scatter_kernel(const MgTileSrc *__restrict__ tab, uint8_t *__restrict__ data) {
    const MgTileSrc t = tab[blockIdx.x];
    uint8_t *dst = data + t.dst;
    if (threadIdx.x == 0) *(uint4 *)dst = make_uint4(t.hdr[0], t.hdr[1], t.hdr[2], t.hdr[3]);
    dst += MG_TILE_HDR;
    const int c0 = t.len0 >> 4;
    for (int i = threadIdx.x; i < t.n * c0; i += blockDim.x) {
        const int r = i / c0, k = i - r * c0;
        ((int4 *)(dst + (size_t)r * t.rb))[k] = ld_stream((const int4 *)(t.src0 + (size_t)r * t.stride) + k);
    }
    const int c1 = t.len1 >> 1;
    for (int i = threadIdx.x; i < t.n * c1; i += blockDim.x) {
        const int r = i / c1, k = i - r * c1;
        ((uint16_t *)(dst + (size_t)r * t.rb + t.len0))[k] = ((const uint16_t *)(t.src1 + (size_t)r * t.stride))[k];
    }
}

void MegaStreams::free() {
    if (data) cudaFree(data);
    if (starts) cudaFree(starts);
    data = nullptr;
    starts = nullptr;
    bytes = 0;
    n_tiles = 0;
}

// Plan the streams (pure host arithmetic): the tile table in stream order, the per-warp starts, the first tile's offset
// of every stream; returns the bytes of the stream buffer.
// This warp's rows of matrix j of a phase: [r0, r1) in the matrix's own row numbering, and the offset dr_off that turns a
// row number into an index of the CTA's partial-sum rows.  Empty when the warp has none.
static bool mg_warp_rows(const MegaPhase &ph, int j, int cta, int n_cta, int warp, int &r0, int &r1, int &dr_off) {
    int u, u_end;
    mg_cta_range(ph, cta, n_cta, u, u_end);
    if (u >= u_end) return false;
    const int g_log = 3 - ph.s_log, g = warp >> ph.s_log; // MG_WARPS = S K-splits x 2^g_log row groups: warp = (row group g, K-split q)
    int lo, hi;
    if (ph.kind == PH_SWIGLU) { // gate rows [u, u_end) and the same rows of up
        lo = u; hi = u_end;
        dr_off = j * (u_end - u) - u;
    } else {
        const int j0 = ph.j0[j];
        lo = std::max(u, j0) - j0;
        hi = std::min(u_end, j0 + (int)ph.job[j].rows) - j0;
        dr_off = j0 - u;
    }
    if (hi <= lo) return false;
    const int n = hi - lo;
    r0 = lo + ((n * g) >> g_log);
    r1 = lo + ((n * (g + 1)) >> g_log);
    return r1 > r0;
}

// Plan the streams (pure host arithmetic): the tile table in stream order, the per-warp starts, the first tile's offset
// of every stream; returns the bytes of the stream buffer.
// Tiles per warp are EQUALISED per matrix: a warp's ring holds two tiles, so a phase lasts as many memory round trips as
// the warp with the most tiles needs -- one warp with 7 tiles among warps with 6 keeps the whole GPU waiting for one more
// round trip at the end of the phase.  Every warp therefore cuts its rows into the same number of tiles (the number the
// warp with the most rows needs), in units of row pairs.
static const bool g_taper = getenv("B200LLM_MEGA_TAPER") && atoi(getenv("B200LLM_MEGA_TAPER")) != 0; // B200LLM_MEGA_TAPER=1 (measured: 1.672 ms per token with it, 1.652 without: off by default)

static uint64_t mega_plan_streams(const MegaPhase *phases, int n_phases, int n_cta, const MegaGeo &geo, std::vector<MgTileSrc> &tab,
                                  std::vector<MegaStreamStart> &st, std::vector<uint64_t> &first_off) {
    tab.clear();
    st.assign((size_t)n_cta * MG_WARPS, MegaStreamStart{});
    first_off.assign(st.size(), 0);
    // tiles per warp of every (phase, matrix)
    std::vector<int> n_tiles_of((size_t)n_phases * 3, 0);
    for (int p = 0; p < n_phases; p++) {
        const MegaPhase &ph = phases[p];
        if (ph.kind == PH_ATT) continue;
        for (int j = 0; j < ph.n_jobs; j++) {
            int most = 0;
            for (int cta = 0; cta < n_cta; cta++)
                for (int warp = 0; warp < MG_WARPS; warp++) {
                    int r0, r1, dr_off;
                    if (mg_warp_rows(ph, j, cta, n_cta, warp, r0, r1, dr_off)) most = std::max(most, r1 - r0);
                }
            const int unit = ph.rpp == 1 ? 2 : ph.rpp, upt = std::max(1, ph.R[j] / unit); // units (row pairs / passes) per tile at most
            n_tiles_of[(size_t)p * 3 + j] = ((most + unit - 1) / unit + upt - 1) / upt;
        }
    }
    uint64_t off = 0;
    for (int cta = 0; cta < n_cta; cta++)
        for (int warp = 0; warp < MG_WARPS; warp++) {
            off = (off + 127) / 128 * 128;
            const size_t first = tab.size();
            long last_prev = -1;
            int first_p = n_phases;
            for (int p = 0; p < n_phases; p++) {
                const MegaPhase &ph = phases[p];
                if (ph.kind == PH_ATT) continue;
                const size_t before = tab.size();
                const int S = ph.S, q = warp & (S - 1);
                const int64_t nb = ph.K >> 8, nbs = nb >> ph.s_log, ks = ph.K >> ph.s_log;
                for (int j = 0; j < ph.n_jobs; j++) {
                    const FusedJob &job = ph.job[j];
                    int r0, r1, dr_off;
                    if (!mg_warp_rows(ph, j, cta, n_cta, warp, r0, r1, dr_off)) continue;
                    int64_t off0 = 0, len0 = 0, off1 = 0, len1 = 0;
                    if (job.type == B2_T_Q4_K) { off0 = q * nbs * 144; len0 = nbs * 144; }
                    else if (job.type == B2_T_Q5_K) { off0 = q * nbs * 176; len0 = nbs * 176; }
                    else if (job.type == B2_T_Q6_K) { off0 = q * nbs * Q6_BLK; len0 = nbs * Q6_BLK; off1 = nb * Q6_BLK + 2 * q * nbs; len1 = 2 * nbs; }
                    else { off0 = q * ks; len0 = ks; off1 = ph.K + q * (ks / 16); len1 = ks / 16; }
                    const int rb = ph.rb[j], R = ph.R[j];
                    B2_CHECK(len0 % 16 == 0 && len0 + len1 <= rb && MG_TILE_HDR + (int64_t)R * rb <= geo.stage_bytes, B2_EINVAL, "mega streams: tile geometry");
                    const int unit = ph.rpp == 1 ? 2 : ph.rpp, upt = std::max(1, R / unit);
                    const int n_rows = r1 - r0, n_units = (n_rows + unit - 1) / unit;
                    int T = std::min(n_tiles_of[(size_t)p * 3 + j], n_units);
                    // Tapered: when the warp's share of the LAST matrix of the phase is several tiles long, its last two tiles
                    // are one unit each.  The final round trips of a phase are what every other SM ends up waiting for; with small
                    // final tiles they deliver few bytes, so the warps of the GPU finish within a fraction of a microsecond of
                    // each other instead of a whole tile time apart.
                    const bool taper = g_taper && j == ph.n_jobs - 1 && T >= 4 && n_units >= T + 0 && (n_units - 2 + upt - 1) / upt + 2 <= T + 1;
                    std::vector<int> sizes;
                    if (taper) {
                        const int body = n_units - 2, Tb = std::max(1, std::min(T + 1 - 2, body));
                        for (int t = 0; t < Tb; t++) sizes.push_back(body / Tb + (t < body % Tb ? 1 : 0));
                        sizes.push_back(1);
                        sizes.push_back(1);
                    } else {
                        for (int t = 0; t < T; t++) sizes.push_back(n_units / T + (t < n_units % T ? 1 : 0));
                    }
                    T = (int)sizes.size();
                    int r = r0;
                    for (int t = 0; t < T; t++) {
                        const int units = sizes[(size_t)t];
                        const int nr = std::min(units * unit, r1 - r);
                        B2_CHECK(nr > 0 && nr <= R, B2_EINVAL, "mega streams: tile rows");
                        MgTileSrc ts;
                        const uint8_t *row = job.W + (int64_t)r * job.row_stride;
                        ts.src0 = row + off0;
                        ts.src1 = row + off1;
                        ts.dst = 0;
                        ts.hdr[0] = (uint32_t)job.type | ((uint32_t)nr << 8) | ((uint32_t)rb << 16);
                        ts.hdr[1] = (uint32_t)(dr_off + r);
                        ts.hdr[2] = 0;
                        ts.hdr[3] = 0;
                        ts.stride = (int32_t)job.row_stride;
                        ts.len0 = (int32_t)len0; ts.len1 = (int32_t)len1;
                        ts.rb = rb; ts.n = nr; ts.pad = p;
                        tab.push_back(ts);
                        r += nr;
                    }
                    B2_CHECK(r == r1, B2_EINVAL, "mega streams: rows left over");
                }
                if (tab.size() > before) {
                    if (first_p == n_phases) first_p = p;
                    if (last_prev >= 0) tab[(size_t)last_prev].hdr[3] |= (uint32_t)p << 8;
                    tab.back().hdr[3] |= 1u;
                    last_prev = (long)tab.size() - 1;
                }
            }
            if (last_prev >= 0) tab[(size_t)last_prev].hdr[3] |= (uint32_t)n_phases << 8;
            const size_t n_t = tab.size() - first;
            for (size_t i = first; i < tab.size(); i++) {
                tab[i].dst = off;
                off += MG_TILE_HDR + (uint64_t)tab[i].n * tab[i].rb;
            }
            auto tile_bytes = [&](size_t i) { return (uint32_t)(MG_TILE_HDR + tab[i].n * tab[i].rb); };
            for (size_t i = first; i + 2 < tab.size(); i++) tab[i].hdr[2] = tile_bytes(i + 2);
            MegaStreamStart &ss = st[(size_t)cta * MG_WARPS + warp];
            ss.first_p = first_p;
            ss.size0 = n_t > 0 ? tile_bytes(first) : 0;
            ss.size1 = n_t > 1 ? tile_bytes(first + 1) : 0;
            ss.total = n_t > 0 ? (uint32_t)(off - tab[first].dst) : 0;
            first_off[(size_t)cta * MG_WARPS + warp] = n_t > 0 ? tab[first].dst : 0;
        }
    return off + 128;
}

// Replay every warp's walk over its stream the way the kernel does it (sizes from the start entry and the headers'
// look-ahead field, phase hand-over from the last-tile flag) and check that each (row, K-split) of each phase is
// produced exactly once, from the right source bytes.  Throws on the first inconsistency.
static void mega_plan_verify(const MegaPhase *phases, int n_phases, int n_cta, const MegaGeo &geo, const std::vector<MgTileSrc> &tab,
                             const std::vector<MegaStreamStart> &st, const std::vector<uint64_t> &first_off, uint64_t total) {
    auto fail = [](const std::string &m) { throw Error(B2_EINVAL, "mega stream plan: " + m); };
    std::vector<std::vector<uint8_t>> cover((size_t)n_phases * n_cta);
    for (int p = 0; p < n_phases; p++)
        for (int cta = 0; cta < n_cta; cta++) {
            if (phases[p].kind == PH_ATT) continue;
            int u, u_end;
            mg_cta_range(phases[p], cta, n_cta, u, u_end);
            const int n_r = (u_end - u) * (phases[p].kind == PH_SWIGLU ? 2 : 1);
            if ((int64_t)n_r * phases[p].S > MG_MAX_PART) fail("partial sums do not fit");
            cover[(size_t)p * n_cta + cta].assign((size_t)n_r * phases[p].S, 0);
        }
    size_t idx = 0;
    for (int cta = 0; cta < n_cta; cta++)
        for (int warp = 0; warp < MG_WARPS; warp++) {
            const MegaStreamStart &ss = st[(size_t)cta * MG_WARPS + warp];
            uint32_t pend[2] = {ss.size0, ss.size1};
            uint64_t src = first_off[(size_t)cta * MG_WARPS + warp];
            uint64_t issued = src + ss.size0 + ss.size1; // where the next refill reads
            int next_p = ss.first_p;
            for (int p = 0; p < n_phases; p++) {
                const MegaPhase &ph = phases[p];
                if (ph.kind == PH_ATT || next_p != p) continue;
                int u, u_end;
                mg_cta_range(ph, cta, n_cta, u, u_end);
                if (u >= u_end) fail("tiles in a phase without rows");
                const int q = warp & (ph.S - 1);
                while (true) {
                    if (idx >= tab.size()) fail("stream runs past the table");
                    const MgTileSrc &t = tab[idx];
                    const uint32_t sz = (uint32_t)(MG_TILE_HDR + t.n * t.rb);
                    if (t.pad != p) fail("tile of another phase");
                    if (pend[0] != sz || t.dst != src) fail("tile size / offset chain broken");
                    if (sz > (uint32_t)geo.stage_bytes || sz % 16 || t.dst % 16 || t.dst + sz > total) fail("tile does not fit");
                    const int type = t.hdr[0] & 0xFF, n = (t.hdr[0] >> 8) & 0xFF, rb = (int)(t.hdr[0] >> 16);
                    if (n != t.n || rb != t.rb || n <= 0) fail("header fields");
                    auto &cv = cover[(size_t)p * n_cta + cta];
                    for (int r = 0; r < n; r++) {
                        const size_t c = ((size_t)t.hdr[1] + r) * ph.S + q;
                        if (c >= cv.size()) fail("partial index out of range");
                        if (cv[c]++) fail("row-split produced twice");
                        // the source row this partial index stands for
                        int dr = (int)t.hdr[1] + r, jb = 0, row;
                        if (ph.kind == PH_SWIGLU) { jb = dr / (u_end - u); row = u + dr % (u_end - u); }
                        else { row = u + dr; while (jb + 1 < ph.n_jobs && row >= (int)ph.job[jb].rows) { row -= (int)ph.job[jb].rows; jb++; } }
                        if (ph.job[jb].type != type) fail("tile type");
                        const int64_t nb = ph.K >> 8, nbs = nb >> ph.s_log, ks = ph.K >> ph.s_log;
                        const int64_t off0 = type == B2_T_Q4_K ? q * nbs * 144 : type == B2_T_Q5_K ? q * nbs * 176 : type == B2_T_Q6_K ? q * nbs * Q6_BLK : q * ks;
                        if (t.src0 + (int64_t)r * t.stride != ph.job[jb].W + (int64_t)row * ph.job[jb].row_stride + off0) fail("source row");
                    }
                    src += sz;
                    pend[0] = pend[1];
                    pend[1] = t.hdr[2];
                    if (t.hdr[2]) {
                        if (issued != src + pend[0]) fail("refill source pointer");
                        issued += t.hdr[2];
                    }
                    idx++;
                    if (t.hdr[3] & 1u) { next_p = (int)(t.hdr[3] >> 8); break; }
                }
            }
            if (next_p != n_phases || pend[0] || pend[1]) fail("stream does not end cleanly");
        }
    if (idx != tab.size()) fail("tiles left over");
    for (auto &cv : cover)
        for (uint8_t c : cv)
            if (c != 1) fail("row-split not produced");
}

int64_t mega_plan_check(const MegaPhase *phases, int n_phases, int n_cta, int64_t *n_tiles, const MegaGeo &geo) {
    std::vector<MgTileSrc> tab;
    std::vector<MegaStreamStart> st;
    std::vector<uint64_t> first_off;
    const uint64_t total = mega_plan_streams(phases, n_phases, n_cta, geo, tab, st, first_off);
    mega_plan_verify(phases, n_phases, n_cta, geo, tab, st, first_off, total);
    if (n_tiles) *n_tiles = (int64_t)tab.size();
    return (int64_t)total;
}

void MegaStreams::build(const MegaPhase *phases, int n_phases, int n_cta, cudaStream_t s, const MegaGeo &geo) {
    free();
    std::vector<MgTileSrc> tab;
    std::vector<MegaStreamStart> st;
    std::vector<uint64_t> first_off;
    const uint64_t off = mega_plan_streams(phases, n_phases, n_cta, geo, tab, st, first_off) - 128;
    bytes = (size_t)off + 128 + (64 << 10); // (slack: the L2 prefetch at a phase end may reach past the last stream)
    n_tiles = (int64_t)tab.size();
    B2_CUDA(cudaMalloc((void **)&data, bytes));
    B2_CUDA(cudaMalloc((void **)&starts, st.size() * sizeof(MegaStreamStart)));
    for (size_t i = 0; i < st.size(); i++) st[i].src = data + first_off[i];
    B2_CUDA(cudaMemsetAsync(data, 0, bytes, s));
    B2_CUDA(cudaMemcpyAsync(starts, st.data(), st.size() * sizeof(MegaStreamStart), cudaMemcpyHostToDevice, s));
    // scatter in slices of the table, so that the staging copy of the table stays small
    const size_t SLICE = (size_t)1 << 18;
    MgTileSrc *dtab = nullptr;
    if (!tab.empty()) {
        B2_CUDA(cudaMalloc((void **)&dtab, std::min(SLICE, tab.size()) * sizeof(MgTileSrc)));
        for (size_t i0 = 0; i0 < tab.size(); i0 += SLICE) {
            const size_t cnt = std::min(SLICE, tab.size() - i0);
            B2_CUDA(cudaMemcpyAsync(dtab, tab.data() + i0, cnt * sizeof(MgTileSrc), cudaMemcpyHostToDevice, s));
            mega_scatter_kernel<<<(unsigned)cnt, 256, 0, s>>>(dtab, data);
            B2_CUDA(cudaGetLastError());
            B2_CUDA(cudaStreamSynchronize(s)); // tab slice is pageable host memory and dtab is reused
        }
    }
    B2_CUDA(cudaStreamSynchronize(s));
    if (dtab) cudaFree(dtab);
}

} // namespace b2
