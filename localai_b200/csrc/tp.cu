// tp.cu -- tensor-parallel decode for models that are too slow (or too big) on one GPU: BASELINE config 5, Llama-3-70B
// Q4_K_M over 2 / 4 / 8 B200s.  The reference reaches several GPUs through llama.cpp's tensor_split / main_gpu
// (backend/backend.proto:175-176 -> backend/cpp/llama/grpc-server.cpp:2240-2262); this is the B200-native replacement.
//
// Partitioning: every shard (one per GPU, all in this process) holds
//   * its rows of wq | wk | wv | ffn_gate | ffn_up | output   (column-parallel: n_head / tp heads, n_head_kv / tp KV heads),
//   * its K-slice of attn_output | ffn_down                     (row-parallel, on 256-weight block boundaries),
//   * its KV heads' cache, and replicas of the norms, the embedding table and the residual stream x.
// A decode step therefore needs one exchange after attn_output and one after ffn_down per layer: the sum over the shards
// of a [n_embd] partial row.  Two implementations:
//   TP_NCCL   the per-layer fused kernels (kernels_fused.cu) with ncclAllReduce(sum, f32) between them -- what BASELINE's
//             north_star literally names; one host thread per shard issues that shard's launches.  ~7 launches and 2
//             collectives per layer: launch- and latency-bound, the baseline the fused path is measured against.
//   TP_FUSED  one persistent decode kernel per GPU (kernels_mega.cu).  The attn_output / ffn_down phases store their
//             partial rows straight into every peer's receive buffer with ordinary 64-bit stores over NVLink ({value,
//             tag} words, so no fence and no flag round trip); the next phase's prologue adds the tp partial rows in rank
//             order (deterministic) plus the residual.  No collective library on the data path, no kernel boundary.
//             Peer memory is only ever touched by plain loads / stores: TMA multicast loads here read only this GPU's
//             memory, since on B200 a TMA multicast load from another GPU's memory hung the GPU.
// Prompts are fed through the same decode path (<= 4 tokens per step in TP_NCCL, 1 in TP_FUSED); logits of the wanted rows are
// gathered on shard 0, where the sampler and the scheduler run unchanged.
#include <dlfcn.h>
#include <nccl.h>

#include <condition_variable>
#include <functional>
#include <thread>

#include "model.h"

namespace b2 {

// NCCL is bound at run time: a process that never asks for tensor parallelism needs no libnccl, and one that already
// carries a copy (PyTorch bundles its own) shares it instead of loading a second one.
struct NcclApi {
    ncclResult_t (*CommInitAll)(ncclComm_t *, int, const int *) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*AllReduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
    const char *(*GetErrorString)(ncclResult_t) = nullptr;
    ncclResult_t (*GetVersion)(int *) = nullptr;
    void *h = nullptr;
    void load() {
        if (h) return;
        h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD | RTLD_GLOBAL);
        if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
        if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
        B2_CHECK(h != nullptr, B2_EUNSUPPORTED, std::string("tensor parallelism (TP_NCCL) needs libnccl.so.2: ") + (dlerror() ? dlerror() : ""));
        auto sym = [&](const char *n) {
            void *p = dlsym(h, n);
            B2_CHECK(p != nullptr, B2_EUNSUPPORTED, std::string("libnccl: missing symbol ") + n);
            return p;
        };
        CommInitAll = (decltype(CommInitAll))sym("ncclCommInitAll");
        CommDestroy = (decltype(CommDestroy))sym("ncclCommDestroy");
        AllReduce = (decltype(AllReduce))sym("ncclAllReduce");
        GetErrorString = (decltype(GetErrorString))sym("ncclGetErrorString");
        GetVersion = (decltype(GetVersion))sym("ncclGetVersion");
    }
};
static NcclApi g_nccl;
#define B2_NCCL(expr)                                                                                                      \
    do {                                                                                                                   \
        ncclResult_t _r = (expr);                                                                                          \
        if (_r != ncclSuccess) throw ::b2::Error(B2_ECUDA, std::string(#expr) + ": " + g_nccl.GetErrorString(_r));          \
    } while (0)

struct TPGroup {
    std::vector<std::unique_ptr<Model>> peers; // shards 1 .. n-1 (shard 0 owns the group)
    std::vector<Model *> shard;                // all of them, by rank
    std::vector<ncclComm_t> comm;
    // one worker thread per shard: it owns that device's launches, so the shards' launch streams run in parallel
    std::vector<std::thread> workers;
    std::mutex mu;
    std::condition_variable cv, cv_done;
    std::function<void(Model &)> job;
    uint64_t gen = 0;
    int pending = 0;
    bool stop = false;
    std::exception_ptr err;
    // all-reduce timing (shard 0): event pairs around every collective of the last forward
    std::vector<cudaEvent_t> ev;
    size_t ev_n = 0;
    bool timing = false;

    void start() {
        for (size_t r = 0; r < shard.size(); r++)
            workers.emplace_back([this, r] {
                Model &m = *shard[r];
                cudaSetDevice(m.device);
                uint64_t seen = 0;
                while (true) {
                    std::function<void(Model &)> f;
                    {
                        std::unique_lock<std::mutex> lk(mu);
                        cv.wait(lk, [&] { return stop || gen != seen; });
                        if (stop) return;
                        seen = gen;
                        f = job;
                    }
                    try {
                        f(m);
                    } catch (...) {
                        std::lock_guard<std::mutex> lk(mu);
                        if (!err) err = std::current_exception();
                    }
                    {
                        std::lock_guard<std::mutex> lk(mu);
                        if (--pending == 0) cv_done.notify_all();
                    }
                }
            });
    }
    // run f(shard) for every shard on that shard's thread; returns when all are done (launches issued, not completed)
    void run_all(std::function<void(Model &)> f) {
        std::unique_lock<std::mutex> lk(mu);
        job = std::move(f);
        pending = (int)shard.size();
        err = nullptr;
        gen++;
        cv.notify_all();
        cv_done.wait(lk, [&] { return pending == 0; });
        if (err) std::rethrow_exception(err);
    }
    ~TPGroup() {
        {
            std::lock_guard<std::mutex> lk(mu);
            stop = true;
        }
        cv.notify_all();
        for (auto &t : workers) t.join();
        for (auto c : comm)
            if (c && g_nccl.CommDestroy) g_nccl.CommDestroy(c);
        for (auto e : ev) cudaEventDestroy(e);
        peers.clear();
    }
};

void Model::tp_init(const GGUFFile &f, const b2_load_opts &o) {
    tp_ = new TPGroup();
    TPGroup &g = *tp_;
    g.shard.push_back(this);
    for (int r = 1; r < tp_size; r++) {
        g.peers.emplace_back(new Model(f, o, r, tp_size));
        g.peers.back()->tp_root_ = this;
        g.shard.push_back(g.peers.back().get());
    }
    // every shard reads and writes every other shard's exchange buffers
    for (int a = 0; a < tp_size; a++) {
        B2_CUDA(cudaSetDevice(g.shard[a]->device));
        for (int b = 0; b < tp_size; b++) {
            if (a == b) continue;
            int ok = 0;
            B2_CUDA(cudaDeviceCanAccessPeer(&ok, g.shard[a]->device, g.shard[b]->device));
            B2_CHECK(ok, B2_EUNSUPPORTED, "tensor parallelism needs peer access between the GPUs (NVLink / NVSwitch)");
            const cudaError_t e = cudaDeviceEnablePeerAccess(g.shard[b]->device, 0);
            if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) B2_CUDA(e);
            cudaGetLastError();
        }
    }
    g_nccl.load();
    std::vector<int> devs;
    for (auto *m : g.shard) devs.push_back(m->device);
    g.comm.assign((size_t)tp_size, nullptr);
    B2_NCCL(g_nccl.CommInitAll(g.comm.data(), tp_size, devs.data()));
    B2_CUDA(cudaSetDevice(device));
    tp_n_allreduce = 2 * hp.n_layer;
    g.start();
    tp_build_fused();
    int mode = TP_FUSED;
    if (getenv("B200LLM_TP_MODE")) mode = atoi(getenv("B200LLM_TP_MODE")) ? TP_FUSED : TP_NCCL;
    if (!tp_fused_ready_) mode = TP_NCCL;
    for (auto *m : g.shard) m->tp_mode_ = mode;
}

void Model::tp_destroy() {
    delete tp_;
    tp_ = nullptr;
}

void Model::tp_set_mode(int mode) {
    B2_CHECK(tp_size > 1, B2_ESTATE, "not a tensor-parallel model");
    B2_CHECK(mode == TP_NCCL || (mode == TP_FUSED && tp_fused_ready_), B2_EUNSUPPORTED, "tensor-parallel mode not available for this model");
    B2_CHECK(tp_ != nullptr, B2_ESTATE, "tp_set_mode: call it on shard 0");
    for (auto *m : tp_->shard) m->tp_mode_ = mode; // every shard must take the same path: they meet in the exchange
}

// xp_ -> x_ : the sum of the shards' partial rows (rows [t0, t0 + n) of the current batch)
void Model::tp_allreduce(int t0, int n) {
    TPGroup &g = tp_root_ ? *tp_root_->tp_ : *tp_;
    const size_t off = (size_t)t0 * hp.n_embd, count = (size_t)n * hp.n_embd;
    const bool timed = tp_rank == 0 && g.timing && g.ev_n + 2 <= g.ev.size();
    if (timed) B2_CUDA(cudaEventRecord(g.ev[g.ev_n++], stream));
    B2_NCCL(g_nccl.AllReduce(xp_ + off, x_ + off, count, ncclFloat32, ncclSum, g.comm[(size_t)tp_rank], stream));
    if (timed) B2_CUDA(cudaEventRecord(g.ev[g.ev_n++], stream));
}

static FusedJob tp_job_of(const DTensor &W) { return FusedJob{W.data, W.type, W.rows, W.row_stride}; }

// One layer on this shard, TP_NCCL: [norm + quantise + QKV + RoPE + KV store] [attention over my heads] [wo K-slice -> partial]
// all-reduce [norm + quantise + gate|up + SwiGLU] [down K-slice -> partial] all-reduce.  Shard 0's partial carries the residual.
void Model::run_layer_tp(const LayerW &L, const KVLayout &kv, int n, int t0) {
    const int d = hp.n_embd, hd = hp.head_dim;
    float *x = x_ + (size_t)t0 * d, *xp = xp_ + (size_t)t0 * d;
    FusedGemv a;
    a.x = x; a.x_stride = d; a.norm_w = (const float *)L.attn_norm.data; a.eps = hp.rms_eps; a.K = d;
    a.fmt = act_format_for(L.wq.type);
    a.n_jobs = 3; a.job[0] = tp_job_of(L.wq); a.job[1] = tp_job_of(L.wk); a.job[2] = tp_job_of(L.wv);
    a.epi = EPI_QKV; a.out = q_ + (size_t)t0 * dq_l; a.out_stride = dq_l; a.pos = pos_ + t0; a.slot = slot_ + t0; a.rope_table = rope_table_;
    a.kv = kv; a.hd = hd;
    launch_fused_gemv(a, n, stream);
    launch_attention(q_ + (size_t)t0 * dq_l, pos_ + t0, slot_ + t0, n, nh_l, nkv_l, hd, kv, att_ + (size_t)t0 * dq_l, n_ctx_slot, stream);
    FusedGemv o;
    o.x = att_ + (size_t)t0 * dq_l; o.x_stride = dq_l; o.K = dq_l; o.fmt = act_format_for(L.wo.type);
    o.n_jobs = 1; o.job[0] = tp_job_of(L.wo); o.epi = EPI_PLAIN; o.out = xp; o.out_stride = d;
    if (tp_rank == 0) { o.residual = x; o.res_stride = d; }
    launch_fused_gemv(o, n, stream);
    tp_allreduce(t0, n);
    FusedGemv gt;
    gt.x = x; gt.x_stride = d; gt.norm_w = (const float *)L.ffn_norm.data; gt.eps = hp.rms_eps; gt.K = d;
    gt.fmt = act_format_for(L.gate.type);
    gt.n_jobs = 2; gt.job[0] = tp_job_of(L.gate); gt.job[1] = tp_job_of(L.up); gt.epi = EPI_SWIGLU; gt.out = gate_ + (size_t)t0 * ff_l; gt.out_stride = ff_l;
    launch_fused_gemv(gt, n, stream);
    FusedGemv w;
    w.x = gate_ + (size_t)t0 * ff_l; w.x_stride = ff_l; w.K = ff_l; w.fmt = act_format_for(L.down.type);
    w.n_jobs = 1; w.job[0] = tp_job_of(L.down); w.epi = EPI_PLAIN; w.out = xp; w.out_stride = d;
    if (tp_rank == 0) { w.residual = x; w.res_stride = d; }
    launch_fused_gemv(w, n, stream);
    tp_allreduce(t0, n);
}

// Forward of the tokens [t0, t0 + n) of the current batch on THIS shard (its worker thread, its device current).
// out_row0 >= 0: the logits of all n tokens go to rows out_row0 .. of logits_ (this shard's vocabulary slice).
void Model::tp_step(int n, int t0, int out_row0) {
    const int d = hp.n_embd;
    if (tp_mode_ == TP_FUSED && n == 1 && tp_fused_ready_) {
        run_decode_mega_tp(t0, out_row0);
        return;
    }
    launch_dequant_rows(tok_embd, tok_ + t0, 0, n, x_ + (size_t)t0 * d, d, stream);
    for (int l = 0; l < hp.n_layer; l++) {
        KVLayout kv{k_pool_ + (size_t)l * pool_layer_stride_, v_pool_ + (size_t)l * pool_layer_stride_, page_table_, page_size, max_pages, kvd_l};
        run_layer_tp(layers[l], kv, n, t0);
    }
    if (out_row0 >= 0) { // LM head: my vocabulary slice of every wanted row
        FusedGemv h;
        h.x = x_ + (size_t)t0 * d; h.x_stride = d; h.norm_w = (const float *)output_norm.data; h.eps = hp.rms_eps;
        h.xn_out = xn_out_ + (size_t)out_row0 * d; h.K = d; h.fmt = act_format_for(output.type);
        h.n_jobs = 1; h.job[0] = tp_job_of(output); h.epi = EPI_PLAIN;
        h.out = logits_ + (size_t)out_row0 * hp.n_vocab + (size_t)tp_rank * vocab_l; h.out_stride = hp.n_vocab;
        launch_fused_gemv(h, n, stream);
    }
}

// The whole batch on all shards; logits of the wanted rows end up in shard 0's logits_ (full vocabulary rows).
void Model::forward_tp(const Batch &b) {
    TPGroup &g = *tp_;
    std::vector<char> want((size_t)b.n, b.n_out < 0 ? 1 : 0);
    int n_out = b.n_out < 0 ? b.n : b.n_out;
    for (int i = 0; b.n_out > 0 && i < b.n_out; i++) {
        B2_CHECK(b.out_rows[i] >= 0 && b.out_rows[i] < b.n && (i == 0 || b.out_rows[i] > b.out_rows[i - 1]), B2_EINVAL, "output rows must be ascending");
        want[(size_t)b.out_rows[i]] = 1;
    }
    for (auto *m : g.shard) {
        B2_CUDA(cudaSetDevice(m->device));
        m->ensure_logits_rows(std::max(n_out, 1));
    }
    B2_CUDA(cudaSetDevice(device));
    g.ev_n = 0;
    const int max_chunk = tp_mode_ == TP_FUSED ? 1 : 4; // the fused decode kernels take one token; the per-layer kernels quantise <= 4 rows per CTA
    g.run_all([&](Model &m) {
        if (b.tokens) B2_CUDA(cudaMemcpyAsync(m.tok_, b.tokens, (size_t)b.n * 4, cudaMemcpyHostToDevice, m.stream));
        B2_CUDA(cudaMemcpyAsync(m.pos_, b.pos, (size_t)b.n * 4, cudaMemcpyHostToDevice, m.stream));
        B2_CUDA(cudaMemcpyAsync(m.slot_, b.slot, (size_t)b.n * 4, cudaMemcpyHostToDevice, m.stream));
        int row = 0;
        for (int t0 = 0; t0 < b.n;) {
            // a chunk = tokens that all want logits or all do not (the LM head runs on whole chunks)
            int n = 1;
            while (n < max_chunk && t0 + n < b.n && want[(size_t)(t0 + n)] == want[(size_t)t0]) n++;
            m.tp_step(n, t0, want[(size_t)t0] ? row : -1);
            if (want[(size_t)t0]) row += n;
            t0 += n;
        }
        if (m.tp_rank != 0 && n_out > 0) { // my vocabulary slice of the wanted rows -> shard 0
            Model &root = *m.tp_root_;
            B2_CUDA(cudaMemcpy2DAsync(root.logits_ + (size_t)m.tp_rank * m.vocab_l, (size_t)hp.n_vocab * 4, m.logits_ + (size_t)m.tp_rank * m.vocab_l,
                                      (size_t)hp.n_vocab * 4, (size_t)m.vocab_l * 4, (size_t)n_out, cudaMemcpyDeviceToDevice, m.stream));
        }
        B2_CUDA(cudaEventRecord(m.tp_done_, m.stream));
    });
    for (auto *m : g.shard)
        if (m != this) B2_CUDA(cudaStreamWaitEvent(stream, m->tp_done_, 0)); // shard 0's stream continues (sampling) after every slice has landed
}

// ------------------------------------------------------------------------------------------------ TP_FUSED
static FusedJob tpj(const DTensor &W) { return FusedJob{W.data, W.type, W.rows, W.row_stride}; }

// Phase lists of the persistent decode kernels: the list of Model::build_mega with this shard's rows, and the two
// row-parallel mat-muls (attn_output, ffn_down) handing their partial rows to every shard.
void Model::tp_build_fused() {
    TPGroup &g = *tp_;
    tp_fused_ready_ = false;
    if (no_mega_ || unfused_) return;
    const int d = hp.n_embd, hd = hp.head_dim;
    const unsigned c0 = getenv("B200LLM_MEGA_COUNTER0") ? (unsigned)strtoul(getenv("B200LLM_MEGA_COUNTER0"), nullptr, 0) : 0u;
    // ---- buffers of every shard first: the lists hold peers' addresses
    for (auto *mp : g.shard) {
        Model &m = *mp;
        B2_CUDA(cudaSetDevice(m.device));
        cudaDeviceGetAttribute(&m.n_sm_, cudaDevAttrMultiProcessorCount, m.device);
        m.mega_att_parts_ = 1;
        while (m.mega_att_parts_ < 4 && m.nh_l * m.mega_att_parts_ * 2 <= m.n_sm_) m.mega_att_parts_ *= 2;
        m.mega_tag_base_ = c0;
        const size_t kmax = (size_t)std::max(d, m.ff_l);
        const size_t n_xq = kmax / 4 + kmax / 256 + kmax / 32 + 8;
        const size_t sizes[11] = {(size_t)d, (size_t)m.dq_l, (size_t)m.dq_l, (size_t)m.ff_l, (size_t)m.kvd_l, (size_t)m.kvd_l, n_xq,
                                  (size_t)m.nh_l * m.n_ctx_slot, (size_t)m.nh_l * m.mega_att_parts_ * hd, (size_t)tp_size * d, (size_t)tp_size * d};
        uint2 **bufs[11] = {&m.x_ll_, &m.q_ll_, &m.att_ll_, &m.gate_ll_, &m.kcur_ll_, &m.vcur_ll_, &m.xq_ll_, &m.mega_att_scores_, &m.mega_att_part_,
                            &m.tp_part_[0], &m.tp_part_[1]};
        size_t most = 0;
        for (size_t n : sizes) most = std::max(most, n);
        std::vector<uint2> init(most, make_uint2(0u, c0));
        for (int i = 0; i < 11; i++) {
            *bufs[i] = (uint2 *)m.dalloc(sizes[i] * sizeof(uint2));
            B2_CUDA(cudaMemcpyAsync(*bufs[i], init.data(), sizes[i] * sizeof(uint2), cudaMemcpyHostToDevice, m.stream));
        }
        m.mega_hint_ = (unsigned *)m.dalloc(256);
        m.mega_hint_base_ = c0;
        {
            std::vector<unsigned> h0(64, c0);
            B2_CUDA(cudaMemcpyAsync(m.mega_hint_, h0.data(), 256, cudaMemcpyHostToDevice, m.stream));
        }
        B2_CUDA(cudaStreamSynchronize(m.stream));
    }
    const bool dist_quant = !(getenv("B200LLM_NO_DIST_QUANT") && atoi(getenv("B200LLM_NO_DIST_QUANT")));
    bool ok = true;
    for (auto *mp : g.shard) {
        Model &m = *mp;
        B2_CUDA(cudaSetDevice(m.device));
        const int r = m.tp_rank;
        std::vector<MegaPhase> ph;
        int prev = -1;   // the phase whose output is the residual stream x (-1: the embedding row, plain fp32 in x_)
        int prev_buf = 0; // which receive buffer holds prev's partial rows
        auto consume_x = [&](MegaPhase &p) { // p reads x: plain before the first exchange, else assembled from the shards' partial rows
            p.x = m.x_; p.x_ll = m.x_ll_; p.src_p = prev;
            if (prev >= 0) {
                p.reduce_n = tp_size; p.reduce_len = d; p.part_ll = m.tp_part_[prev_buf]; p.x_ll_w = m.x_ll_;
            }
        };
        for (int l = 0; l < hp.n_layer && ok; l++) {
            const LayerW &L = m.layers[(size_t)l];
            __half *kp = m.k_pool_ + (size_t)l * m.pool_layer_stride_, *vp = m.v_pool_ + (size_t)l * m.pool_layer_stride_;
            if (act_format_for(L.wq.type) != act_format_for(L.wk.type) || act_format_for(L.wq.type) != act_format_for(L.wv.type) || L.gate.type != L.up.type) { ok = false; break; }
            MegaPhase q;
            q.kind = PH_QKV; q.K = d; q.fmt = act_format_for(L.wq.type); q.n_jobs = 3;
            q.job[0] = tpj(L.wq); q.job[1] = tpj(L.wk); q.job[2] = tpj(L.wv);
            consume_x(q);
            if (prev >= 0) q.red_res_ll = m.x_ll_; // x (after attn_output of the previous layer) + the ffn_down partial rows
            q.norm_w = (const float *)L.attn_norm.data; q.out_ll = m.q_ll_; q.k_pool = kp; q.v_pool = vp;
            q.units = (int)((L.wq.rows + L.wk.rows + L.wv.rows) / 2);
            const int iq = (int)ph.size();
            ph.push_back(q);
            MegaPhase at;
            at.kind = PH_ATT; at.x_ll = m.q_ll_; at.src_p = iq; at.out_ll = m.att_ll_; at.k_pool = kp; at.v_pool = vp;
            const int iat = (int)ph.size();
            ph.push_back(at);
            MegaPhase o; // attn_output, my K-slice: a partial row for everybody
            o.kind = PH_PLAIN; o.K = m.dq_l; o.fmt = act_format_for(L.wo.type); o.n_jobs = 1; o.job[0] = tpj(L.wo);
            o.x_ll = m.att_ll_; o.src_p = iat; o.units = (int)(L.wo.rows / 2);
            o.n_peers = tp_size;
            for (int j = 0; j < tp_size; j++) o.peer_part[j] = g.shard[(size_t)j]->tp_part_[0] + (size_t)r * d;
            const int io = (int)ph.size();
            ph.push_back(o);
            MegaPhase gt;
            gt.kind = PH_SWIGLU; gt.K = d; gt.fmt = act_format_for(L.gate.type); gt.n_jobs = 2; gt.job[0] = tpj(L.gate); gt.job[1] = tpj(L.up);
            gt.x = m.x_; gt.x_ll = m.x_ll_; gt.src_p = io;
            gt.reduce_n = tp_size; gt.reduce_len = d; gt.part_ll = m.tp_part_[0]; gt.x_ll_w = m.x_ll_;
            if (prev < 0) gt.red_residual = m.x_; else gt.red_res_ll = m.x_ll_;
            gt.norm_w = (const float *)L.ffn_norm.data; gt.out_ll = m.gate_ll_; gt.units = (int)L.gate.rows;
            const int ig = (int)ph.size();
            ph.push_back(gt);
            MegaPhase w; // ffn_down, my K-slice
            w.kind = PH_PLAIN; w.K = m.ff_l; w.fmt = act_format_for(L.down.type); w.n_jobs = 1; w.job[0] = tpj(L.down);
            w.x_ll = m.gate_ll_; w.src_p = ig; w.units = (int)(L.down.rows / 2);
            w.n_peers = tp_size;
            for (int j = 0; j < tp_size; j++) w.peer_part[j] = g.shard[(size_t)j]->tp_part_[1] + (size_t)r * d;
            if (dist_quant && w.fmt == ACT_Q8_K && w.K / 256 > 32 && w.K <= 32768) w.xq_ll = m.xq_ll_;
            prev = (int)ph.size();
            prev_buf = 1;
            ph.push_back(w);
        }
        if (!ok) break;
        MegaPhase h; // LM head: my vocabulary rows
        h.kind = PH_PLAIN; h.K = d; h.fmt = act_format_for(m.output.type); h.n_jobs = 1; h.job[0] = tpj(m.output);
        consume_x(h);
        if (prev >= 0) h.red_res_ll = m.x_ll_;
        h.norm_w = (const float *)m.output_norm.data; h.out = m.logits_ + (size_t)r * m.vocab_l; h.xn_out = m.xn_out_; h.units = (int)(m.output.rows / 2);
        ph.push_back(h);
        m.mega_geo_ = mega_geometry(ph.data(), (int)ph.size(), m.n_ctx_slot, hd);
        for (auto &p : ph)
            if (!mega_phase_prepare(p, m.n_sm_, m.mega_geo_)) { ok = false; break; }
        if (!ok) break;
        if ((size_t)m.n_ctx_slot > mega_attention_ctx_limit(hd, m.mega_geo_) || hd % 64 != 0 || hd > 128 || m.page_size < 16 || (m.page_size & (m.page_size - 1)) ||
            (uint64_t)m.pool_layer_stride_ >= (1ull << 31) || m.kvd_l % 2 != 0) { ok = false; break; }
        size_t free_b = 0, total_b = 0;
        if (cudaMemGetInfo(&free_b, &total_b) != cudaSuccess || free_b < (size_t)m.decode_stream_bytes + m.shadow_reserve_) { ok = false; break; }
        m.mega_streams_.build(ph.data(), (int)ph.size(), m.n_sm_, m.stream, m.mega_geo_);
        m.mega_phases_ = (MegaPhase *)m.dalloc(ph.size() * sizeof(MegaPhase));
        B2_CUDA(cudaMemcpyAsync(m.mega_phases_, ph.data(), ph.size() * sizeof(MegaPhase), cudaMemcpyHostToDevice, m.stream));
        B2_CUDA(cudaStreamSynchronize(m.stream));
        m.mega_n_phases_ = (int)ph.size();
        int coop = 0;
        cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, m.device);
        if (!coop) ok = false;
        if (getenv("B200LLM_MEGA_DBG")) m.mega_dbg_ = (long long *)m.dalloc((size_t)m.n_sm_ * m.mega_n_phases_ * 16 * sizeof(long long));
    }
    B2_CUDA(cudaSetDevice(device));
    for (auto *mp : g.shard) mp->tp_fused_ready_ = ok;
    tp_fused_ready_ = ok;
}

// One decode step of token t0 on this shard through its persistent kernel (the other shards' kernels run concurrently on
// their GPUs and meet this one through the partial-row buffers).
void Model::run_decode_mega_tp(int t0, int out_row0) {
    const int d = hp.n_embd;
    launch_dequant_rows(tok_embd, tok_ + t0, 0, 1, x_, d, stream);
    MegaArgs a;
    a.phases = mega_phases_; a.n_phases = mega_n_phases_; a.starts = mega_streams_.starts;
    a.pos = pos_ + t0; a.slot = slot_ + t0; a.rope_table = rope_table_; a.page_table = page_table_;
    a.page_size = page_size; a.max_pages = max_pages; a.kvd = kvd_l;
    a.n_head = nh_l; a.n_head_kv = nkv_l; a.hd = hp.head_dim;
    a.eps = hp.rms_eps; a.att_scale = 1.0f / sqrtf((float)hp.head_dim);
    a.page_shift = 0;
    while ((1 << a.page_shift) < page_size) a.page_shift++;
    a.att_parts = mega_att_parts_;
    a.att_parts_log = 0;
    while ((1 << a.att_parts_log) < a.att_parts) a.att_parts_log++;
    a.att_scores = mega_att_scores_; a.att_scores_stride = n_ctx_slot;
    a.att_part_out = mega_att_part_;
    a.kcur_ll = kcur_ll_; a.vcur_ll = vcur_ll_;
    a.tag_base = mega_tag_base_;
    mega_tag_base_ += (unsigned)mega_n_phases_;
    a.geo = mega_geo_;
    mega_fill_knobs(a);
    a.hint = mega_hint_;
    a.hint_base = mega_hint_base_;
    if (a.ll_probe == 2) mega_hint_base_ += (unsigned)n_sm_ * (unsigned)(mega_n_phases_ - 1);
    a.dbg = mega_dbg_;
    // the logits row of this step (unwanted rows go to a spare row past the last one the caller can ask for)
    const int row = out_row0 >= 0 ? out_row0 : logits_rows_cap_ - 1;
    a.out_override = logits_ + (size_t)row * hp.n_vocab + (size_t)tp_rank * vocab_l;
    a.xn_override = xn_out_ + (size_t)std::max(out_row0, 0) * d;
    launch_decode_mega(a, n_sm_, stream);
}

void Model::forward_device_inputs_tp(int n) {
    TPGroup &g = *tp_;
    B2_CUDA(cudaEventRecord(tp_in_, stream)); // shard 0's token / position buffers are final (the sampler wrote them)
    g.run_all([&](Model &m) {
        if (m.tp_rank != 0) {
            Model &root = *m.tp_root_;
            B2_CUDA(cudaStreamWaitEvent(m.stream, root.tp_in_, 0));
            B2_CUDA(cudaMemcpyPeerAsync(m.tok_, m.device, root.tok_, root.device, (size_t)n * 4, m.stream));
            B2_CUDA(cudaMemcpyPeerAsync(m.pos_, m.device, root.pos_, root.device, (size_t)n * 4, m.stream));
            B2_CUDA(cudaMemcpyPeerAsync(m.slot_, m.device, root.slot_, root.device, (size_t)n * 4, m.stream));
        }
        for (int t = 0; t < n; t++) m.tp_step(1, t, t);
        if (m.tp_rank != 0) {
            Model &root = *m.tp_root_;
            B2_CUDA(cudaMemcpy2DAsync(root.logits_ + (size_t)m.tp_rank * m.vocab_l, (size_t)hp.n_vocab * 4, m.logits_ + (size_t)m.tp_rank * m.vocab_l,
                                      (size_t)hp.n_vocab * 4, (size_t)m.vocab_l * 4, (size_t)n, cudaMemcpyDeviceToDevice, m.stream));
        }
        B2_CUDA(cudaEventRecord(m.tp_done_, m.stream));
    });
    for (auto *m : g.shard)
        if (m != this) B2_CUDA(cudaStreamWaitEvent(stream, m->tp_done_, 0));
}

void Model::tp_timing(bool on) {
    TPGroup &g = *tp_;
    g.timing = on;
    g.ev_n = 0;
    if (on) {
        B2_CUDA(cudaSetDevice(device));
        while (g.ev.size() < (size_t)(4 * tp_n_allreduce + 8)) {
            cudaEvent_t e;
            B2_CUDA(cudaEventCreate(&e));
            g.ev.push_back(e);
        }
    }
}

// ms inside the collectives since tp_timing(true) / the last call (shard 0's stream; call after a sync)
double Model::tp_allreduce_ms() {
    TPGroup &g = *tp_;
    double tot = 0;
    for (size_t i = 0; i + 1 < g.ev_n; i += 2) {
        float ms = 0;
        B2_CUDA(cudaEventElapsedTime(&ms, g.ev[i], g.ev[i + 1]));
        tot += ms;
    }
    g.ev_n = 0;
    return tot;
}

void Model::tp_sync() {
    for (auto *m : tp_->shard) {
        B2_CUDA(cudaSetDevice(m->device));
        B2_CUDA(cudaStreamSynchronize(m->stream));
    }
    B2_CUDA(cudaSetDevice(device));
}

} // namespace b2
