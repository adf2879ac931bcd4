// noalloc.cu -- b2_mul_mat_noalloc: the quantised mat-mul for callers that are not allowed to allocate.
//
// Who needs this: code that runs where device allocation is forbidden or impossible -- a host framework that
// pre-reserved all of HBM for its own pool (so cudaMalloc fails or fragments it), a real-time serving loop that
// must not take the allocator lock or trigger the implicit device synchronisation cudaMalloc/cudaFree can cause,
// and embedders whose rules simply say "no allocation after start-up".  Such a caller already owns device buffers
// for the weights, the activations and the result; the only thing the mat-mul needs beyond those is a small
// scratch area for the int8-quantised activation rows.  Here that scratch is a fixed array compiled into the
// library, so the call allocates nothing -- no cudaMalloc, no cudaFree, no library routine that allocates.
//
// Cost and limits: every GPU that loads the library carries the 128 KiB array until the process exits.  The call
// handles at most B2_NOALLOC_MAX_TOK activation rows of at most B2_NOALLOC_MAX_COLS elements; larger requests are
// rejected with B2_EINVAL rather than silently truncated.
//
// Sharing: there is ONE array per device and every call uses it.  It keeps whatever the previous call left behind,
// which is harmless: nothing in it has to start at zero, because the quantisation pass of each call rewrites every
// byte the mat-mul pass of that call reads.  But two calls that overlap on the same GPU (different streams, or
// different threads) would scribble over each other's activations -- callers must serialise them.
//
// Set-up: the first call on each device looks up the device address of the array (cudaGetSymbolAddress, with that
// device current) and remembers it.  Make that first call before starting any stream capture.
#include <mutex>

#include "kernels.h"

namespace b2 {

// scratch for the quantised activation rows; sized for the worst format (f16: 2 bytes per element)
constexpr int64_t NOALLOC_MAX_COLS = 16384;
constexpr int NOALLOC_MAX_TOK = 4;
constexpr size_t NOALLOC_SCRATCH_BYTES = (size_t)NOALLOC_MAX_TOK * NOALLOC_MAX_COLS * 2; // 128 KiB

static __device__ __align__(256) unsigned char g_noalloc_scratch[NOALLOC_SCRATCH_BYTES];

// In host code `g_noalloc_scratch` is only a handle for the runtime, not an address a kernel can use; the real
// device address is fetched once per device and cached here.
constexpr int NOALLOC_MAX_DEVICES = 64;
static std::mutex g_noalloc_mu;
static uint8_t *g_noalloc_ptr[NOALLOC_MAX_DEVICES] = {nullptr};

static uint8_t *noalloc_scratch_for_current_device() {
    int dev = -1;
    B2_CUDA(cudaGetDevice(&dev));
    B2_CHECK(dev >= 0 && dev < NOALLOC_MAX_DEVICES, B2_EINVAL, "b2_mul_mat_noalloc: device ordinal out of range");
    std::lock_guard<std::mutex> lk(g_noalloc_mu);
    if (!g_noalloc_ptr[dev]) {
        void *p = nullptr;
        B2_CUDA(cudaGetSymbolAddress(&p, g_noalloc_scratch)); // resolves against the CURRENT device's copy
        B2_CHECK(p != nullptr, B2_ECUDA, "b2_mul_mat_noalloc: scratch symbol has no device address");
        g_noalloc_ptr[dev] = (uint8_t *)p;
    }
    return g_noalloc_ptr[dev];
}

int mul_mat_noalloc(int wtype, const void *w_dev, int64_t rows, int64_t cols, const float *x_dev, int n_tok, float *out_dev,
                    cudaStream_t stream) {
    B2_CHECK(w_dev && x_dev && out_dev, B2_EINVAL, "b2_mul_mat_noalloc: null device pointer");
    B2_CHECK(wtype == B2_T_Q4_K || wtype == B2_T_Q5_K || wtype == B2_T_F16, B2_EUNSUPPORTED,
             "b2_mul_mat_noalloc: weight type must be Q4_K, Q5_K or F16 (types whose device layout is the GGUF layout)");
    B2_CHECK(rows > 0 && cols > 0 && cols % 256 == 0, B2_EINVAL, "b2_mul_mat_noalloc: cols must be a positive multiple of 256");
    B2_CHECK(cols <= NOALLOC_MAX_COLS, B2_EINVAL,
             "b2_mul_mat_noalloc: cols exceeds the built-in scratch (max " + std::to_string(NOALLOC_MAX_COLS) + ")");
    B2_CHECK(n_tok >= 1 && n_tok <= NOALLOC_MAX_TOK, B2_EINVAL,
             "b2_mul_mat_noalloc: n_tok exceeds the built-in scratch (max " + std::to_string(NOALLOC_MAX_TOK) + ")");
    const ActFmt f = act_format_for(wtype);
    B2_CHECK((size_t)n_tok * (size_t)act_row_bytes(f, cols) <= NOALLOC_SCRATCH_BYTES, B2_EINVAL,
             "b2_mul_mat_noalloc: request does not fit the built-in scratch");
    uint8_t *scratch = noalloc_scratch_for_current_device();
    DTensor W;
    W.data = (const uint8_t *)w_dev;
    W.type = wtype;
    W.rows = rows;
    W.cols = cols;
    W.row_stride = device_row_stride(wtype, cols);
    launch_quantize_act(f, x_dev, cols, n_tok, cols, scratch, stream);
    launch_gemv(W, scratch, f, n_tok, out_dev, rows, nullptr, 0, stream);
    return 0;
}

} // namespace b2

extern "C" int b2_mul_mat_noalloc(int wtype, const void *w_dev, int64_t rows, int64_t cols, const float *x_dev, int n_tok,
                                  float *out_dev, void *cuda_stream) {
    try {
        return b2::mul_mat_noalloc(wtype, w_dev, rows, cols, x_dev, n_tok, out_dev, (cudaStream_t)cuda_stream);
    } catch (const b2::Error &e) {
        b2::set_last_error(e.what());
        return e.code;
    } catch (const std::exception &e) {
        b2::set_last_error(e.what());
        return B2_EINVAL;
    }
}
