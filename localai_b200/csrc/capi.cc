// capi.cc -- extern "C" face of libb200llm.so (include/b200llm.h).
#include <cuda_fp16.h>

#include <cstring>
#include <memory>
#include <mutex>
#include <vector>

#include "stream_text.h"
#include "bert.h"
#include "engine.h"
#include "grammar.h"
#include "kernels.h"
#include "sampling.h"
#include "model.h"

namespace b2 {
static thread_local std::string g_err;
void set_last_error(const std::string &m) { g_err = m; }
} // namespace b2

using namespace b2;

#define B2_API_BEGIN try {
#define B2_API_END(failret)                       \
    }                                             \
    catch (const b2::Error &e) {                  \
        b2::set_last_error(e.what());             \
        return (failret == 0 ? e.code : failret); \
    }                                             \
    catch (const std::exception &e) {             \
        b2::set_last_error(e.what());             \
        return (failret == 0 ? B2_EINVAL : failret); \
    }

// ------------------------------------------------------------------------------------------------ kernel test hooks
namespace {
struct DevBuf {
    void *p = nullptr;
    explicit DevBuf(size_t n) { B2_CUDA(cudaMalloc(&p, n ? n : 16)); }
    ~DevBuf() { cudaFree(p); }
    template <typename T> T *as() { return (T *)p; }
};
int sm100_device_count();
void require_device() {
    B2_CHECK(sm100_device_count() > 0, B2_ENODEV, "no sm_100 CUDA device (this library has no CPU fallback)");
    B2_CUDA(cudaSetDevice(0));
}
DTensor upload_matrix(int type, const void *blocks, int64_t rows, int64_t cols, DevBuf &staging, DevBuf &dst) {
    DTensor t;
    t.type = type;
    t.rows = rows;
    t.cols = cols;
    t.row_stride = device_row_stride(type, cols);
    B2_CUDA(cudaMemcpy(staging.p, blocks, (size_t)(rows * gguf_row_bytes(type, cols)), cudaMemcpyHostToDevice));
    launch_repack(type, staging.as<uint8_t>(), dst.as<uint8_t>(), rows, cols, 0);
    t.data = dst.as<uint8_t>();
    return t;
}
} // namespace

extern "C" {

int b2_abi_version(void) { return B2_ABI_VERSION; }
const char *b2_last_error(void) { return b2::g_err.c_str(); }

int b2_device_count(void) { return sm100_device_count(); }

} // extern "C"
namespace {
int sm100_device_count() {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    int ok = 0;
    for (int i = 0; i < n; i++) {
        cudaDeviceProp p;
        if (cudaGetDeviceProperties(&p, i) == cudaSuccess && p.major == 10) ok++;
    }
    return ok;
}
} // namespace
extern "C" {

b2_engine *b2_create(const b2_load_opts *opts) {
    try {
        B2_CHECK(opts && opts->model_path, B2_EINVAL, "b2_create: null options / model path");
        b2_load_opts o;
        memset(&o, 0, sizeof(o));
        memcpy(&o, opts, std::min<size_t>(opts->struct_size ? opts->struct_size : sizeof(o), sizeof(o)));
        return reinterpret_cast<b2_engine *>(new Engine(o));
    } catch (const b2::Error &e) {
        b2::set_last_error(e.what());
    } catch (const std::exception &e) {
        b2::set_last_error(e.what());
    }
    return nullptr;
}

void b2_destroy(b2_engine *e) { delete reinterpret_cast<Engine *>(e); }

static Engine *E(b2_engine *e) {
    B2_CHECK(e != nullptr, B2_EINVAL, "null engine handle");
    return reinterpret_cast<Engine *>(e);
}

int b2_model_info(b2_engine *e, b2_model_info_t *out) {
    B2_API_BEGIN
    B2_CHECK(out, B2_EINVAL, "null out");
    E(e)->model_info(out);
    return 0;
    B2_API_END(0)
}

int b2_tokenize(b2_engine *e, const char *text, size_t n, int add_bos, int parse_special, int32_t *out, int cap) {
    B2_API_BEGIN
    std::vector<int32_t> ids = E(e)->tokenize(std::string(text ? text : "", text ? n : 0), add_bos != 0, parse_special != 0);
    if ((int)ids.size() > cap) return -(int)ids.size();
    if (!ids.empty()) memcpy(out, ids.data(), ids.size() * 4);
    return (int)ids.size();
    B2_API_END(0)
}

int b2_token_to_piece(b2_engine *e, int32_t token, char *out, int cap) {
    B2_API_BEGIN
    const std::string p = E(e)->token_to_piece(token);
    if ((int)p.size() > cap) return -(int)p.size();
    memcpy(out, p.data(), p.size());
    return (int)p.size();
    B2_API_END(0)
}

int64_t b2_submit(b2_engine *e, const b2_request *req) {
    B2_API_BEGIN
    B2_CHECK(req, B2_EINVAL, "null request");
    return E(e)->submit(*req);
    B2_API_END(0)
}

int b2_poll(b2_engine *e, int64_t id, b2_event *ev, int timeout_ms) {
    B2_API_BEGIN
    B2_CHECK(ev, B2_EINVAL, "null event");
    return E(e)->poll(id, ev, timeout_ms);
    B2_API_END(0)
}

int b2_cancel(b2_engine *e, int64_t id) {
    B2_API_BEGIN
    return E(e)->cancel(id);
    B2_API_END(0)
}

int b2_embed_result(b2_engine *e, int64_t id, float *out, int cap) {
    B2_API_BEGIN
    return E(e)->embed_result(id, out, cap);
    B2_API_END(0)
}

int b2_release(b2_engine *e, int64_t id) {
    B2_API_BEGIN
    return E(e)->release(id);
    B2_API_END(0)
}

int b2_metrics(b2_engine *e, b2_metrics_t *out) {
    B2_API_BEGIN
    B2_CHECK(out, B2_EINVAL, "null out");
    E(e)->metrics(out);
    return 0;
    B2_API_END(0)
}

int b2_eval(b2_engine *e, int slot, const int32_t *tokens, int n, int n_past, float *logits, int logits_all, float *embd) {
    B2_API_BEGIN
    return E(e)->eval(slot, tokens, n, n_past, logits, logits_all != 0, embd);
    B2_API_END(0)
}

int b2_eval_step(b2_engine *e, const int32_t *tokens, const int32_t *pos, const int32_t *slots, int n, float *logits) {
    B2_API_BEGIN
    return E(e)->eval_step(tokens, pos, slots, n, logits);
    B2_API_END(0)
}

int b2_decode_greedy(b2_engine *e, int slot, const int32_t *prompt, int n_prompt, int n_gen, int32_t *out, double *prefill_ms,
                     double *decode_ms) {
    B2_API_BEGIN
    return E(e)->decode_greedy(slot, prompt, n_prompt, n_gen, out, prefill_ms, decode_ms);
    B2_API_END(0)
}

int b2_debug_hidden(b2_engine *e, float *out, int cap_floats) {
    B2_API_BEGIN
    return E(e)->debug_hidden(out, cap_floats);
    B2_API_END(0)
}

int b2_bench_decode(b2_engine *e, int batch, int ctx_len, int n_steps, int n_warmup, double *ms_per_step, int64_t *launches) {
    B2_API_BEGIN
    return E(e)->bench_decode(batch, ctx_len, n_steps, n_warmup, ms_per_step, launches);
    B2_API_END(0)
}

// undocumented bench hook (not part of the ABI header): mean duration in ms of one decode_mega_kernel launch during the
// last b2_bench_decode with batch 1 (CUDA events on the engine's stream around every launch); 0 if that kernel did not run
double b2_bench_last_mega_ms(b2_engine *e) {
    try {
        return e ? E(e)->last_mega_ms() : 0.0;
    } catch (...) {
        return 0.0;
    }
}

int b2_tp_set_mode(b2_engine *e, int mode) {
    B2_API_BEGIN
    std::lock_guard<std::mutex> lk(E(e)->device_mutex());
    E(e)->model().tp_set_mode(mode);
    return 0;
    B2_API_END(0)
}

int b2_tp_info(b2_engine *e, b2_tp_info_t *out) {
    B2_API_BEGIN
    B2_CHECK(out != nullptr, B2_EINVAL, "b2_tp_info: no output");
    Model &m = E(e)->model();
    out->tp_size = m.tp_size;
    out->mode = m.tp_size > 1 ? m.tp_mode() : 0;
    out->n_allreduce_per_step = m.tp_size > 1 ? m.tp_n_allreduce : 0;
    out->allreduce_ms_per_step = E(e)->last_tp_allreduce_ms_;
    out->shard_stream_bytes = m.decode_stream_bytes;
    return 0;
    B2_API_END(0)
}

// undocumented profiling hook (not part of the ABI header): per-CTA, per-phase clock64 stamps of the decode mega-kernel
int b2_mega_debug(b2_engine *e, long long *out, int cap) {
    B2_API_BEGIN
    return E(e)->mega_debug(out, cap);
    B2_API_END(0)
}

// undocumented test hooks (not part of the ABI header; need no device): the request path's text and prompt rules
// (stream_text.h) driven piece by piece.  b2_stream_text: pieces = the tokens' bytes back to back with their lengths; returns
// the bytes a streaming client would have received in `streamed`, the final text in `final_text`, *n_used = pieces consumed
// (generation stops at a stop word).  Return value: 1 if a stop word ended it, 0 if not, negative on error.
extern "C" int b2_stream_text(const char *pieces, const int *lens, int n, const char *const *stops, int n_stops, char *streamed, int cap_s,
                              char *final_text, int cap_f, int *n_used) {
    B2_API_BEGIN
    std::vector<std::string> sw;
    for (int i = 0; i < n_stops; i++) sw.emplace_back(stops[i]);
    StreamText t;
    t.reset(&sw);
    std::string out;
    int used = 0, stopped = 0;
    size_t off = 0;
    for (int i = 0; i < n && !stopped; i++, used++) {
        StreamText::Step st = t.append(std::string(pieces + off, (size_t)lens[i]));
        off += (size_t)lens[i];
        out += st.release;
        stopped = st.stop_word ? 1 : 0;
    }
    B2_CHECK((int)out.size() < cap_s && (int)t.text().size() < cap_f, B2_ENOSPC, "b2_stream_text: buffers too small");
    memcpy(streamed, out.c_str(), out.size() + 1);
    memcpy(final_text, t.text().c_str(), t.text().size() + 1);
    if (n_used) *n_used = used;
    return stopped;
    B2_API_END(0)
}
// b2_partial_stop: where in `text` its longest suffix that is a prefix of `word` begins (-1: none) -- the hold-back rule
// of the streamed text (undocumented test hook, needs no device)
extern "C" int64_t b2_partial_stop(const char *text, size_t n, const char *word, size_t m) {
    B2_API_BEGIN
    const size_t at = StreamText::prefix_overlap(std::string(text, n), std::string(word, m));
    return at == std::string::npos ? (int64_t)-1 : (int64_t)at;
    B2_API_END(-2)
}
// b2_fit_request_prompt: what LoadModel's slot does to a prompt of n tokens for a request that asks for n_keep (raw, as the
// request carries it): returns the new length (undocumented test hook)
extern "C" int b2_fit_request_prompt(int32_t *tokens, int n, int n_ctx, int n_keep) {
    B2_API_BEGIN
    std::vector<int32_t> v(tokens, tokens + n);
    fit_prompt_to_context(v, n_ctx, effective_n_keep(n_keep, n, n_ctx));
    std::copy(v.begin(), v.end(), tokens);
    return (int)v.size();
    B2_API_END(-1)
}
// b2_utf8_incomplete: does the text end inside a multi-byte character (its bytes are then held back)? (undocumented test hook)
extern "C" int b2_utf8_incomplete(const char *text, size_t n) {
    B2_API_BEGIN
    return StreamText::ends_inside_utf8(std::string(text, n)) ? 1 : 0;
    B2_API_END(-1)
}
// b2_reusable_prefix: how many tokens of `prompt` a slot that holds `cached` need not evaluate again (undocumented test hook)
extern "C" int b2_reusable_prefix(const int32_t *cached, int n_cached, const int32_t *prompt, int n_prompt) {
    B2_API_BEGIN
    return reusable_prefix(std::vector<int32_t>(cached, cached + n_cached), std::vector<int32_t>(prompt, prompt + n_prompt));
    B2_API_END(-1)
}
// b2_fit_prompt: the truncation rule on a token list in place; returns the new length (tokens dropped iff it changed)
extern "C" int b2_fit_prompt(int32_t *tokens, int n, int n_ctx, int n_keep) {
    B2_API_BEGIN
    std::vector<int32_t> v(tokens, tokens + n);
    fit_prompt_to_context(v, n_ctx, n_keep);
    std::copy(v.begin(), v.end(), tokens);
    return (int)v.size();
    B2_API_END(0)
}

// undocumented test hook (not part of the ABI header; needs no device): replay the batched-decode GEMM's work split
// (stream-K ranges, split-K segment arithmetic, partial slots) for a launch of n_tiles 128-row tiles x nkb super-blocks.
// Returns the number of tiles whose K range is split over several CTAs, or a negative error code.
int64_t b2_qgemm_plan_check(int n_tiles, int nkb, int n_sm) {
    B2_API_BEGIN
    B2_CHECK(n_tiles > 0 && nkb > 0 && n_sm > 0, B2_EINVAL, "b2_qgemm_plan_check: bad arguments");
    return qgemm_plan_check(n_tiles, nkb, n_sm);
    B2_API_END(0)
}

// undocumented test hook (not part of the ABI header; needs no device): plan the decode kernel's weight streams for a
// phase list given as 9 int64 per phase {kind, K, n_jobs, type0, rows0, type1, rows1, type2, rows2} and replay every
// warp's walk over them.  Returns the stream bytes, or a negative error code when the planner is inconsistent.
int64_t b2_mega_plan_check(const int64_t *desc, int n_phases, int n_cta, int64_t *n_tiles) {
    B2_API_BEGIN
    B2_CHECK(desc && n_phases > 0 && n_cta > 0, B2_EINVAL, "b2_mega_plan_check: bad arguments");
    std::vector<MegaPhase> ph((size_t)n_phases);
    for (int p = 0; p < n_phases; p++) {
        const int64_t *d = desc + 9 * p;
        MegaPhase &m = ph[(size_t)p];
        m.kind = (int)d[0];
        if (m.kind == PH_ATT) continue;
        m.K = d[1];
        m.n_jobs = (int)d[2];
        B2_CHECK(m.n_jobs >= 1 && m.n_jobs <= 3, B2_EINVAL, "b2_mega_plan_check: 1..3 matrices per phase");
        int64_t rows = 0;
        for (int j = 0; j < m.n_jobs; j++) {
            const int type = (int)d[3 + 2 * j];
            m.job[j] = FusedJob{(const uint8_t *)(uintptr_t)(((uint64_t)p * 3 + j + 1) << 36), type, d[4 + 2 * j], device_row_stride(type, m.K)};
            rows += d[4 + 2 * j];
        }
        m.fmt = act_format_for(m.job[0].type);
        m.units = (int)(m.kind == PH_SWIGLU ? m.job[0].rows : rows / 2);
        if (m.kind == PH_PLAIN && m.K / 256 > 32 && m.K <= 32768) m.xq_ll = (uint2 *)16; // as Model::build_mega: quantised once per GPU
    }
    const MegaGeo geo = mega_geometry(ph.data(), n_phases, 2048, 128);
    for (int p = 0; p < n_phases; p++)
        if (ph[(size_t)p].kind != PH_ATT)
            B2_CHECK(mega_phase_prepare(ph[(size_t)p], n_cta, geo), B2_EUNSUPPORTED, "phase " + std::to_string(p) + ": no lane mapping for this shape/type");
    return mega_plan_check(ph.data(), n_phases, n_cta, n_tiles, geo);
    B2_API_END(0)
}

int b2_bench_gemv(b2_engine *e, int which, int n_tok, int n_rounds, double *ms_per_launch, int64_t *bytes_per_launch,
                  int *ggml_type) {
    B2_API_BEGIN
    return E(e)->bench_gemv(which, n_tok, n_rounds, ms_per_launch, bytes_per_launch, ggml_type);
    B2_API_END(0)
}

int b2_bench_prefill(b2_engine *e, int n_tokens, int n_iters, int n_warmup, double *ms_per_iter, int64_t *launches) {
    B2_API_BEGIN
    return E(e)->bench_prefill(n_tokens, n_iters, n_warmup, ms_per_iter, launches);
    B2_API_END(0)
}


int b2_dequantize(int type, const void *blocks, int64_t n_elems, float *out) {
    B2_API_BEGIN
    require_device();
    const int be = type_block_elems(type);
    B2_CHECK(be > 0, B2_EUNSUPPORTED, "unsupported ggml type");
    const int64_t row_elems = 256; // every supported type packs 256 elements into whole blocks
    B2_CHECK(n_elems > 0 && n_elems % row_elems == 0, B2_EINVAL, "n_elems must be a positive multiple of 256");
    const int64_t rows = n_elems / row_elems;
    DevBuf staging((size_t)(rows * gguf_row_bytes(type, row_elems))), dst((size_t)(rows * device_row_stride(type, row_elems))),
        o((size_t)n_elems * 4);
    DTensor W = upload_matrix(type, blocks, rows, row_elems, staging, dst);
    launch_dequant_rows(W, nullptr, 0, (int)rows, o.as<float>(), row_elems, 0);
    B2_CUDA(cudaMemcpy(out, o.p, (size_t)n_elems * 4, cudaMemcpyDeviceToHost));
    return 0;
    B2_API_END(0)
}

int b2_quantize_act(int wtype, const float *x, int64_t n_elems, void *out, size_t out_cap) {
    B2_API_BEGIN
    require_device();
    B2_CHECK(n_elems > 0 && n_elems % 256 == 0, B2_EINVAL, "n_elems must be a positive multiple of 256");
    const ActFmt f = act_format_for(wtype);
    const int64_t K = n_elems;
    const int64_t rb = act_row_bytes(f, K);
    DevBuf dx((size_t)K * 4), dq((size_t)rb);
    B2_CUDA(cudaMemcpy(dx.p, x, (size_t)K * 4, cudaMemcpyHostToDevice));
    launch_quantize_act(f, dx.as<float>(), K, 1, K, dq.as<uint8_t>(), 0);
    std::vector<uint8_t> soa((size_t)rb);
    B2_CUDA(cudaMemcpy(soa.data(), dq.p, (size_t)rb, cudaMemcpyDeviceToHost));
    uint8_t *o = (uint8_t *)out;
    if (f == ACT_Q8_0) { // -> block_q8_0 { fp16 d; int8 qs[32]; }
        const int64_t nb = K / 32;
        B2_CHECK(out_cap >= (size_t)nb * 34, B2_ENOSPC, "output buffer too small");
        const float *d = (const float *)(soa.data() + K);
        for (int64_t b = 0; b < nb; b++) {
            const __half h = __float2half_rn(d[b]);
            memcpy(o + b * 34, &h, 2);
            memcpy(o + b * 34 + 2, soa.data() + b * 32, 32);
        }
        return (int)(nb * 34);
    }
    if (f == ACT_Q8_K) { // -> block_q8_K { float d; int8 qs[256]; int16 bsums[16]; }
        const int64_t nb = K / 256;
        B2_CHECK(out_cap >= (size_t)nb * 292, B2_ENOSPC, "output buffer too small");
        const uint8_t *d = soa.data() + K;
        const uint8_t *bs = soa.data() + K + nb * 4;
        for (int64_t b = 0; b < nb; b++) {
            memcpy(o + b * 292, d + b * 4, 4);
            memcpy(o + b * 292 + 4, soa.data() + b * 256, 256);
            memcpy(o + b * 292 + 260, bs + b * 32, 32);
        }
        return (int)(nb * 292);
    }
    const size_t n = f == ACT_F16 ? (size_t)K * 2 : (size_t)K * 4;
    B2_CHECK(out_cap >= n, B2_ENOSPC, "output buffer too small");
    memcpy(o, soa.data(), n);
    return (int)n;
    B2_API_END(0)
}

int b2_mul_mat(int wtype, const void *blocks, int64_t rows, int64_t cols, const float *x, int n_tok, float *out, int path) {
    B2_API_BEGIN
    require_device();
    B2_CHECK(type_block_elems(wtype) > 0, B2_EUNSUPPORTED, "unsupported ggml type");
    B2_CHECK(rows > 0 && cols > 0 && n_tok > 0 && cols % (type_tiles_only(wtype) ? std::max(64, type_block_elems(wtype)) : 256) == 0, B2_EINVAL,
             "bad shape (cols must be a multiple of 256; 64 for the legacy 32-element block types)");
    B2_CHECK(!type_tiles_only(wtype) || path == 2, B2_EUNSUPPORTED, "Q4_0 / Q4_1 / Q5_0 / Q5_1 / Q2_K / Q3_K / IQ4_NL / IQ4_XS run on the tensor-core path (2) only");
    B2_CHECK(path >= 0 && path <= 6, B2_EUNSUPPORTED, "unknown mat-mul path");
    DevBuf staging((size_t)(rows * gguf_row_bytes(wtype, cols))), dst((size_t)(rows * device_row_stride(wtype, cols)));
    DTensor W = upload_matrix(wtype, blocks, rows, cols, staging, dst);
    const ActFmt f = act_format_for(wtype);
    const int64_t rb = act_row_bytes(f, cols);
    DevBuf dx((size_t)n_tok * cols * 4), dq((size_t)n_tok * rb), dout((size_t)n_tok * rows * 4);
    B2_CUDA(cudaMemcpy(dx.p, x, (size_t)n_tok * cols * 4, cudaMemcpyHostToDevice));
    const uint8_t *xq = dq.as<uint8_t>();
    if (f == ACT_F32) xq = dx.as<uint8_t>();
    else launch_quantize_act(f, dx.as<float>(), cols, n_tok, cols, dq.as<uint8_t>(), 0);
    if (path == 2) { // tensor-core GEMM (fp16 operands, fp32 accumulate)
        DevBuf wt(gemm_w_tiles_bytes(rows, cols)), xt(gemm_x_tiles_bytes(n_tok, cols)), sk((size_t)32 << 20);
        launch_gemm_prepare_x(dx.as<float>(), cols, n_tok, cols, xt.p, 0);
        launch_gemm_tc(W, xt.p, n_tok, dout.as<float>(), rows, nullptr, 0, wt.p, false, sk.as<float>(), (size_t)32 << 20, 0);
        B2_CUDA(cudaDeviceSynchronize());
    } else if (path == 6) { // batched-decode GEMM that dequantises inside the kernel (fp16 operands, fp32 accumulate)
        B2_CHECK(qgemm_type_supported(wtype) && n_tok <= QGEMM_MAX_TOKENS, B2_EUNSUPPORTED, "qgemm path: Q4_K / Q5_K / Q6_K / Q8_0, at most 64 tokens");
        int n_sm = 0;
        B2_CUDA(cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, 0));
        DevBuf qt(qgemm_w_bytes(wtype, rows, cols)), xt(qgemm_x_bytes(n_tok, cols)), part(qgemm_partial_bytes(n_sm)), cnt(QGEMM_COUNTERS * 4);
        B2_CUDA(cudaMemset(cnt.p, 0, QGEMM_COUNTERS * 4));
        launch_qgemm_repack(W, qt.p, 0);
        launch_qgemm_prepare_x(dx.as<float>(), cols, n_tok, cols, xt.p, 0);
        QGemmJob j{&W, qt.p, dout.as<float>(), rows};
        for (int rep = 0; rep < 2; rep++) // twice: the second launch runs on the counters the first one left behind
            launch_qgemm_multi(&j, 1, xt.p, n_tok, nullptr, 0, part.as<float>(), qgemm_partial_bytes(n_sm), cnt.as<unsigned>(), QGEMM_COUNTERS, n_sm, 0);
        B2_CUDA(cudaDeviceSynchronize());
    } else if (path == 3 || path == 5) { // fused decode kernel: 3 = quantise in the prologue, 5 = pre-quantised rows
        FusedGemv a;
        if (path == 3) { a.x = dx.as<float>(); a.x_stride = cols; }
        else a.xq_pre = xq;
        a.K = cols; a.fmt = f; a.n_jobs = 1;
        a.job[0] = FusedJob{W.data, W.type, W.rows, W.row_stride};
        a.epi = EPI_PLAIN; a.out = dout.as<float>(); a.out_stride = rows;
        launch_fused_gemv(a, n_tok, 0);
    } else if (path == 4) { // persistent decode kernel with a single mat-vec phase
        B2_CHECK(n_tok == 1 && rows % 2 == 0, B2_EINVAL, "mega path: one activation row, even row count");
        MegaPhase ph;
        ph.kind = PH_PLAIN; ph.K = cols; ph.fmt = f; ph.n_jobs = 1;
        ph.job[0] = FusedJob{W.data, W.type, W.rows, W.row_stride};
        ph.x = dx.as<float>(); ph.out = dout.as<float>(); ph.units = (int)(rows / 2);
        int n_sm = 0;
        B2_CUDA(cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, 0));
        const MegaGeo geo = mega_geometry(&ph, 1, 64, 64);
        B2_CHECK(mega_phase_prepare(ph, n_sm, geo), B2_EUNSUPPORTED, "mega path: shape/type not supported");
        MegaStreams ms;
        ms.build(&ph, 1, n_sm, 0, geo);
        struct Guard { MegaStreams &m; ~Guard() { m.free(); } } guard{ms};
        DevBuf dph(sizeof(MegaPhase)), dps(64);
        B2_CUDA(cudaMemcpy(dph.p, &ph, sizeof(ph), cudaMemcpyHostToDevice));
        B2_CUDA(cudaMemset(dps.p, 0, 64));
        MegaArgs a;
        a.geo = geo;
        mega_fill_knobs(a);
        a.hint = dps.as<unsigned>() + 8; // (a single phase never waits on it)
        a.phases = dph.as<MegaPhase>(); a.n_phases = 1; a.starts = ms.starts; a.pos = dps.as<int32_t>(); a.slot = dps.as<int32_t>() + 4;
        a.page_size = 32; a.page_shift = 5; a.max_pages = 1; a.hd = 64; a.n_head = 1; a.n_head_kv = 1;
        launch_decode_mega(a, n_sm, 0);
        B2_CUDA(cudaDeviceSynchronize());
    } else {
        launch_gemv(W, xq, f, n_tok, dout.as<float>(), rows, nullptr, 0, 0);
    }
    B2_CUDA(cudaMemcpy(out, dout.p, (size_t)n_tok * rows * 4, cudaMemcpyDeviceToHost));
    return 0;
    B2_API_END(0)
}

extern "C" int b2_qgemm_debug(long long *out, int n) { return b2::qgemm_debug_read(out, n); } // profiling builds only (not in the header)

struct b2_bert {
    std::unique_ptr<BertModel> m;
};

b2_bert *b2_bert_load(const char *model_path, int device, int vocab_only) {
    try {
        B2_CHECK(model_path, B2_EINVAL, "b2_bert_load: no model path");
        B2_CHECK(vocab_only || sm100_device_count() > 0, B2_ENODEV, "no sm_100 CUDA device (this library has no CPU fallback)");
        auto h = std::make_unique<b2_bert>();
        h->m.reset(new BertModel(model_path, device, vocab_only != 0));
        return h.release();
    } catch (const std::exception &e) {
        b2::set_last_error(e.what());
        return nullptr;
    }
}

void b2_bert_free(b2_bert *b) { delete b; }

int b2_bert_info(b2_bert *b, b2_bert_info_t *out) {
    B2_API_BEGIN
    B2_CHECK(b && out, B2_EINVAL, "b2_bert_info: bad arguments");
    const BertModel &m = *b->m;
    *out = b2_bert_info_t{m.n_layer, m.n_embd, m.n_head, m.n_ff, m.n_ctx, m.n_vocab, m.weight_bytes};
    return B2_OK;
    B2_API_END(0)
}

int b2_bert_tokenize(b2_bert *b, const char *text, size_t text_len, int32_t *out, int cap) {
    B2_API_BEGIN
    B2_CHECK(b && (text || text_len == 0), B2_EINVAL, "b2_bert_tokenize: bad arguments");
    const std::vector<int32_t> ids = b->m->tokenize(std::string(text ? text : "", text_len));
    if ((int)ids.size() > cap || !out) return -(int)ids.size();
    memcpy(out, ids.data(), ids.size() * 4);
    return (int)ids.size();
    B2_API_END(0)
}

int b2_bert_embed_tokens(b2_bert *b, const int32_t *ids, int n_ids, float *out, int cap) {
    B2_API_BEGIN
    B2_CHECK(b && ids && out && n_ids > 0, B2_EINVAL, "b2_bert_embed_tokens: bad arguments");
    B2_CHECK(cap >= b->m->n_embd, B2_ENOSPC, "b2_bert_embed_tokens: the output needs n_embd floats");
    const int32_t offs[2] = {0, n_ids};
    b->m->embed(ids, offs, 1, out);
    return b->m->n_embd;
    B2_API_END(0)
}

int b2_bert_embed_text(b2_bert *b, const char *text, size_t text_len, float *out, int cap) {
    B2_API_BEGIN
    B2_CHECK(b && out && (text || text_len == 0), B2_EINVAL, "b2_bert_embed_text: bad arguments");
    B2_CHECK(cap >= b->m->n_embd, B2_ENOSPC, "b2_bert_embed_text: the output needs n_embd floats");
    const std::vector<int32_t> ids = b->m->tokenize(std::string(text ? text : "", text_len));
    const int32_t offs[2] = {0, (int32_t)ids.size()};
    b->m->embed(ids.data(), offs, 1, out);
    return b->m->n_embd;
    B2_API_END(0)
}

int b2_bert_embed_batch(b2_bert *b, const int32_t *ids, const int32_t *offsets, int n_seq, float *out) {
    B2_API_BEGIN
    B2_CHECK(b && ids && offsets && out && n_seq > 0, B2_EINVAL, "b2_bert_embed_batch: bad arguments");
    b->m->embed(ids, offsets, n_seq, out);
    return n_seq;
    B2_API_END(0)
}

int b2_bert_hidden(b2_bert *b, const int32_t *ids, int n_ids, float *out) {
    B2_API_BEGIN
    B2_CHECK(b && ids && out && n_ids > 0, B2_EINVAL, "b2_bert_hidden: bad arguments");
    b->m->hidden(ids, n_ids, out);
    return n_ids;
    B2_API_END(0)
}

int b2_grammar_match(const char *gbnf, const char *text, size_t text_len, size_t *n_accepted, int *can_stop) {
    B2_API_BEGIN
    B2_CHECK(gbnf && (text || text_len == 0), B2_EINVAL, "b2_grammar_match: bad arguments");
    Grammar g(grammar_parse(gbnf));
    size_t i = 0;
    while (i < text_len) { // one code point at a time, so that the count stops at the first one that does not fit
        size_t n = 1;
        while (i + n < text_len && ((uint8_t)text[i + n] >> 6) == 2) n++;
        if (!g.accept_bytes(std::string(text + i, n))) break;
        i += n;
    }
    if (n_accepted) *n_accepted = i;
    if (can_stop) *can_stop = g.can_stop() ? 1 : 0;
    return B2_OK;
    B2_API_END(0)
}

int b2_grammar_token_mask(const char *gbnf, const char *accepted, size_t accepted_len, const char *const *pieces, const size_t *piece_len,
                          const uint8_t *is_eog, int n_vocab, uint32_t *mask_bits) {
    B2_API_BEGIN
    B2_CHECK(gbnf && pieces && piece_len && mask_bits && n_vocab > 0, B2_EINVAL, "b2_grammar_token_mask: bad arguments");
    Grammar g(grammar_parse(gbnf));
    B2_CHECK(accepted_len == 0 || g.accept_bytes(std::string(accepted, accepted_len)), B2_EINVAL,
             "b2_grammar_token_mask: the accepted text does not fit the grammar");
    GrammarVocab v;
    v.cps.resize(n_vocab); v.tails.resize(n_vocab); v.is_eog.resize(n_vocab); v.is_empty.resize(n_vocab); v.pieces.resize(n_vocab);
    for (int t = 0; t < n_vocab; t++) {
        v.pieces[t].assign(pieces[t] ? pieces[t] : "", pieces[t] ? piece_len[t] : 0);
        v.is_eog[t] = is_eog ? is_eog[t] : 0;
        v.is_empty[t] = v.pieces[t].empty() || v.pieces[t][0] == '\0';
        Utf8Tail tail;
        utf8_decode_piece(v.pieces[t], tail, v.cps[t]);
        v.tails[t] = tail;
    }
    g.token_mask(v, mask_bits);
    for (int t = 0; t < n_vocab; t++) // the single-candidate check must agree with the mask
        B2_CHECK(g.token_fits(v, t) == (bool)((mask_bits[t >> 5] >> (t & 31)) & 1u), B2_ESTATE, "b2_grammar_token_mask: mask and check disagree");
    return B2_OK;
    B2_API_END(0)
}

int b2_sample(int device, const float *logits, int n_vocab, const b2_sample_params *p, int n_draws, int32_t *out_tokens, float *out_mu) {
    B2_API_BEGIN
    B2_CHECK(logits && p && out_tokens && n_vocab > 0 && n_draws > 0, B2_EINVAL, "b2_sample: bad arguments");
    B2_CHECK(p->struct_size == sizeof(b2_sample_params), B2_EINVAL, "b2_sample: struct_size mismatch");
    B2_CHECK(device >= 0 && sm100_device_count() > device, B2_ENODEV, "b2_sample: no such sm_100 device");
    B2_CUDA(cudaSetDevice(device));
    DevBuf src((size_t)n_vocab * 4), row((size_t)n_vocab * 4), spec(sizeof(SampleSpec)), tok(4), mu(4);
    B2_CUDA(cudaMemcpy(src.p, logits, (size_t)n_vocab * 4, cudaMemcpyHostToDevice));
    SampleSpec sp;
    sp.temperature = p->temperature; sp.top_k = p->top_k; sp.top_p = p->top_p > 0 ? p->top_p : 1.f; sp.min_p = p->min_p;
    sp.ban_token = p->ban_token; sp.nl_token = p->nl_token;
    sp.repeat_penalty = p->repeat_penalty > 0 ? p->repeat_penalty : 1.f;
    sp.frequency_penalty = p->frequency_penalty; sp.presence_penalty = p->presence_penalty;
    sp.n_prev = std::min(std::max(p->n_prev, 0), SAMPLE_MAX_PREV);
    for (int k = 0; k < sp.n_prev; k++) sp.prev[k] = p->prev[p->n_prev - sp.n_prev + k];
    sp.tfs_z = (p->tfs_z > 0.f && p->tfs_z < 1.f) ? p->tfs_z : 1.f;
    sp.typical_p = (p->typical_p > 0.f && p->typical_p < 1.f) ? p->typical_p : 1.f;
    sp.mirostat = p->mirostat; sp.mirostat_tau = p->mirostat_tau; sp.mirostat_eta = p->mirostat_eta; sp.mirostat_mu = p->mirostat_mu;
    uint64_t rng = (uint64_t)p->seed;
    for (int i = 0; i < n_draws; i++) {
        sp.rng = rng;
        rng = rng * 6364136223846793005ULL + 1442695040888963407ULL; // the scheduler's per-token advance
        B2_CUDA(cudaMemcpy(row.p, src.p, (size_t)n_vocab * 4, cudaMemcpyDeviceToDevice));
        B2_CUDA(cudaMemcpy(spec.p, &sp, sizeof(sp), cudaMemcpyHostToDevice));
        launch_sample(row.as<float>(), n_vocab, 1, n_vocab, spec.as<SampleSpec>(), sp.has_penalties(), tok.as<int32_t>(), mu.as<float>(), 0,
                      sp.temperature <= 0.f);
        B2_CUDA(cudaMemcpy(&out_tokens[i], tok.p, 4, cudaMemcpyDeviceToHost));
        B2_CUDA(cudaMemcpy(&sp.mirostat_mu, mu.p, 4, cudaMemcpyDeviceToHost));
        if (out_mu) out_mu[i] = sp.mirostat_mu;
    }
    return 0;
    B2_API_END(0)
}

} // extern "C"
