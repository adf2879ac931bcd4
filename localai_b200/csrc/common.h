// common.h -- shared host-side declarations for libb200llm (sm_100a only; no CPU fallback).
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <mutex>
#include <stdexcept>
#include <string>
#include <utility>

#include "../../include/b200llm.h"

namespace b2 {

struct Error : std::runtime_error {
    int code;
    Error(int c, const std::string &m) : std::runtime_error(m), code(c) {}
};

void set_last_error(const std::string &m);

#define B2_CUDA(expr)                                                                                   \
    do {                                                                                                \
        cudaError_t _e = (expr);                                                                        \
        if (_e != cudaSuccess)                                                                          \
            throw ::b2::Error(_e == cudaErrorMemoryAllocation ? B2_ENOMEM : B2_ECUDA,                   \
                              std::string(#expr) + ": " + cudaGetErrorString(_e) + " (" + __FILE__ +   \
                                  ":" + std::to_string(__LINE__) + ")");                               \
    } while (0)

#define B2_CHECK(cond, code, msg)                                  \
    do {                                                           \
        if (!(cond)) throw ::b2::Error((code), std::string(msg));  \
    } while (0)

// cudaFuncSetAttribute (opt-in shared memory above 48 KB) is PER DEVICE: one `static PerDeviceOnce` per launch site runs
// its set-up the first time the site is reached with each device current, whichever host thread that is.
struct PerDeviceOnce {
    std::mutex mu;
    uint64_t done = 0;
    template <class F>
    void run(F &&f) {
        int dev = 0;
        B2_CUDA(cudaGetDevice(&dev));
        std::lock_guard<std::mutex> lk(mu);
        if (dev < 64 && ((done >> dev) & 1)) return;
        f();
        if (dev < 64) done |= 1ull << dev;
    }
};

// Programmatic dependent launch (the kernels of a batched decode step; opt-in with B200LLM_PDL=1): the next kernel of
// the stream may be set up while this one still runs; it blocks in B2_PDL_WAIT() -- before its first access to anything an
// earlier kernel of the stream may have written or may still read -- until the previous kernel has completed and flushed.
// Every kernel launched this way waits before it exits, so completion is transitive along the stream.  Both instructions
// are no-ops in a kernel launched without the attribute.
#ifdef __CUDACC__
#define B2_PDL_TRIGGER() asm volatile("griddepcontrol.launch_dependents;" ::: "memory")
#define B2_PDL_WAIT() asm volatile("griddepcontrol.wait;" ::: "memory")
bool pdl_enabled();
template <typename... KArgs, typename... Args>
inline void launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t s, Args &&...args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = s;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = pdl_enabled() ? 1 : 0;
    B2_CUDA(cudaLaunchKernelEx(&cfg, kernel, KArgs(std::forward<Args>(args))...));
}
#endif

// ---- tensor types (ggml ids) and on-device layouts -------------------------------------------------
// Q4_K / Q5_K / F16 / F32 rows keep the GGUF layout (their blocks are 16-byte multiples).
// Q8_0 (34 B) and Q6_K (210 B) blocks are only 2-byte aligned on disk, so LoadModel repacks each ROW
// so that everything a kernel reads with 128-bit loads is 16-byte aligned (same byte count, rounded up
// to 16 per row): the fp16 block scales move to the end of the row.
//   Q8_0 row: [qs: cols x int8][d: nb x fp16]                             nb = cols/32
//   Q6_K row: [nb x {ql 128, qh 64, scales 16} = 208 B][d: nb x fp16]     nb = cols/256
struct DTensor {
    const uint8_t *data = nullptr; // device
    int type = -1;
    int64_t rows = 0, cols = 0;
    int64_t row_stride = 0; // bytes between consecutive rows on device
    int64_t bytes() const { return rows * row_stride; }
};

inline int64_t align_up(int64_t x, int64_t a) { return (x + a - 1) / a * a; }

inline int type_block_elems(int t) {
    switch (t) {
        case B2_T_F32: case B2_T_F16: return 1;
        case B2_T_Q8_0: case B2_T_Q4_0: case B2_T_Q4_1: case B2_T_Q5_0: case B2_T_Q5_1: case B2_T_IQ4_NL: return 32;
        case B2_T_Q4_K: case B2_T_Q5_K: case B2_T_Q6_K: case B2_T_Q2_K: case B2_T_Q3_K: case B2_T_IQ4_XS: return 256;
    }
    return 0;
}
inline int type_block_bytes(int t) {
    switch (t) {
        case B2_T_F32: return 4; case B2_T_F16: return 2; case B2_T_Q8_0: return 34;
        case B2_T_Q4_0: return 18; case B2_T_Q4_1: return 20; case B2_T_Q5_0: return 22; case B2_T_Q5_1: return 24;
        case B2_T_Q2_K: return 84; case B2_T_Q3_K: return 110; case B2_T_IQ4_NL: return 18; case B2_T_IQ4_XS: return 136;
        case B2_T_Q4_K: return 144; case B2_T_Q5_K: return 176; case B2_T_Q6_K: return 210;
    }
    return 0;
}
// bytes of one row in the GGUF file
inline int64_t gguf_row_bytes(int t, int64_t cols) { return cols / type_block_elems(t) * type_block_bytes(t); }
// bytes of one row on the device (after repack)
inline int64_t device_row_stride(int t, int64_t cols) { return align_up(gguf_row_bytes(t, cols), 16); }

// formats without an integer decode kernel: tensor-core path (fp16 tiles) only
inline bool type_tiles_only(int t) {
    return t == B2_T_Q4_0 || t == B2_T_Q4_1 || t == B2_T_Q5_0 || t == B2_T_Q5_1 || t == B2_T_Q2_K || t == B2_T_Q3_K ||
           t == B2_T_IQ4_NL || t == B2_T_IQ4_XS;
}

// activation formats (what llama.cpp converts src1 to before the dot: ggml type_traits.vec_dot_type)
enum ActFmt { ACT_F32 = 0, ACT_F16 = 1, ACT_Q8_0 = 2, ACT_Q8_K = 3, ACT_BF16 = 4 };
inline ActFmt act_format_for(int wtype) {
    switch (wtype) {
        case B2_T_F32: return ACT_F32;
        case B2_T_F16: return ACT_F16;
        case B2_T_Q8_0: return ACT_Q8_0;
        default: return ACT_Q8_K;
    }
}
// bytes of one quantised activation row on device (SoA):
//   Q8_K: [qs: K int8][d: K/256 f32][bsums: K/16 int16]
//   Q8_0: [qs: K int8][d: K/32 f32 (value of the fp16-rounded scale)]
inline int64_t act_row_bytes(ActFmt f, int64_t K) {
    switch (f) {
        case ACT_F32: return K * 4;
        case ACT_F16: case ACT_BF16: return align_up(K * 2, 16);
        case ACT_Q8_0: return align_up(K + K / 32 * 4, 16);
        case ACT_Q8_K: return align_up(K + K / 256 * 4 + K / 16 * 2, 16);
    }
    return 0;
}

} // namespace b2
