#!/usr/bin/env python3
"""bench.py -- BASELINE.json's metric on BASELINE.json's config, one JSON line on stdout.

Workload (config.workload): configs[1] = Llama-3-8B-arch GGUF Q4_K_M, random-init, batch-1 greedy decode
(the north-star number) at a short context.  A "step" is one decode step (one token for every active sequence).

  value     decode tok/s with everything resident in HBM (b2_bench_decode: CUDA events on the engine's stream)
  e2e       the same metric through the request path a LocalAI backend uses (b2_submit / b2_poll, i.e.
            PredictStream): host prompt ids in, token text out, per-step H2D/D2H inside the timed region;
            tok/s = n_decoded / t_generation as the reference defines it (grpc-server.cpp:305-318)
  roofline  the dominant kernel IS the step: one launch of decode_mega_kernel per token.  achieved = (every weight
            byte once + the KV rows read) / the mean duration of that kernel's launches, measured live with CUDA events
            recorded around every launch on the engine's stream; peak = MEASURED_PEAKS.json hbm_gbs; traffic = the
            dram bytes of one launch from the committed ncu capture (profiles/r2_mega_traffic.json)
  cpu_baseline  the oracle (C restatement of the llama.cpp CPU path, scalar loops under OpenMP -- NOT llama.cpp's
            AVX2 / AVX-512 build, which cannot be built offline) on a bounded sample of the same workload, same GGUF
  parity_checked  the first 16 tokens of the e2e request equal the oracle's greedy continuation of the same prompt
  extra     the other BASELINE configs on the same box, each with its own roofline: TinyLlama Q8_0 b=1 (configs[0]),
            64 concurrent streams (configs[2]), 2048-token prefill, BERT-base F16 embeddings (configs[3]); at N > 1
            also `pool`: ONE process driving a ReplicaPool over the N devices (the deployment shape)

N > 1 (torchrun): independent per-GPU replicas, no data-path collective ("replicas only");
value = tokens of all ranks / max-over-ranks time.

--impl reference times the CPU implementation alone (the reference's llama.cpp is not buildable here, so it is
the oracle port), rank 0 only.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))

METRIC = "Llama-3-8B Q4_K_M decode tok/s (b=1)"
UNIT = "tok/s"


def mega_traffic():
    """dram__bytes_read.sum + dram__bytes_write.sum of ONE decode_mega_kernel launch (llama3-8b q4_k_m, ctx ~128), as
    tools/ncu_traffic.py extracted it from the committed ncu --set full capture; None if no capture is committed."""
    p = os.path.join(ROOT, "profiles", "r2_mega_traffic.json")
    try:
        with open(p) as f:
            d = json.load(f)
        return float(d["dram_bytes_read"]) + float(d["dram_bytes_write"]), d.get("source")
    except Exception:
        return None, None


def log(*a):
    print(*a, file=sys.stderr, flush=True)


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""

    def __init__(self, index: int):
        self.index, self.rows, self.proc = index, [], None

    def __enter__(self):
        q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def __exit__(self, *a):
        if self.proc:
            time.sleep(0.15)
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:
                self.proc.kill()

    def summary(self):
        sm, mx, reasons = [], 0.0, set()
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx = max(mx, float(r[1]))
            except Exception:
                continue
            for name, v in zip(["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"], r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx or None, "reasons": sorted(reasons),
                "samples": len(sm)}


def ensure_model(arch: str, preset: str, seed: int, rank: int, barrier) -> tuple[str, dict]:
    import make_gguf as mg

    d = os.environ.get("B200LLM_BENCH_DIR", "/dev/shm/b200llm_bench" if os.path.isdir("/dev/shm") else "/tmp/b200llm_bench")
    os.makedirs(d, exist_ok=True)
    path = os.path.join(d, f"{arch}-{preset}-seed{seed}.gguf")
    meta = path + ".json"
    if rank == 0 and not (os.path.exists(path) and os.path.exists(meta)):
        t0 = time.time()
        info = mg.write_gguf(path + ".tmp", mg.ARCHS[arch], preset, seed=seed, quiet=True)
        os.replace(path + ".tmp", path)
        with open(meta, "w") as f:
            json.dump(info, f)
        log(f"[bench] wrote {path} ({info['file_bytes'] / 1e9:.2f} GB) in {time.time() - t0:.1f}s")
    barrier()
    with open(meta) as f:
        return path, json.load(f)


def cpu_sample(path: str, n_prompt: int, n_decode: int):
    """Oracle port timed on the host cores: (decode tok/s, prompt tok/s, threads)."""
    import numpy as np

    from oracle import oracle as O

    O.set_num_threads(os.cpu_count() or 1)
    m = O.OracleModel(path, n_ctx=n_prompt + n_decode + 8)
    prompt = (np.arange(n_prompt) * 7919 % 4000 + 3).astype(np.int32)
    t0 = time.time()
    lg = m.eval(prompt)
    t1 = time.time()
    tok = int(lg[0].argmax())
    for i in range(n_decode):
        lg = m.eval([tok], n_past=n_prompt + i)
        tok = int(lg[0].argmax())
    t2 = time.time()
    return n_decode / (t2 - t1), n_prompt / (t1 - t0), O.num_threads()


def run_reference(args, rank: int, emit):
    if rank != 0:
        return
    path, info = ensure_model(args.arch, args.preset, args.seed, 0, lambda: None)
    import numpy as np

    from oracle import oracle as O

    O.set_num_threads(os.cpu_count() or 1)  # all host cores, also under torchrun (which exports OMP_NUM_THREADS=1)
    n_prompt = args.ctx  # the same context the GPU arm decodes at
    m = O.OracleModel(path, n_ctx=args.ctx + args.steps + 64)
    prompt = (np.arange(n_prompt) * 7919 % 4000 + 3).astype(np.int32)
    lg = m.eval(prompt)
    tok, pos = int(lg[0].argmax()), n_prompt
    budget_s = float(os.environ.get("B200LLM_REF_BUDGET_S", "150"))
    t_w = time.time()
    for _ in range(max(1, min(args.warmup, 2))):
        lg = m.eval([tok], n_past=pos)
        tok, pos = int(lg[0].argmax()), pos + 1
    per = (time.time() - t_w) / max(1, min(args.warmup, 2))
    steps = max(1, min(args.steps, int(budget_s / max(per, 1e-3))))
    t0 = time.time()
    for _ in range(steps):
        lg = m.eval([tok], n_past=pos)
        tok, pos = int(lg[0].argmax()), pos + 1
    dt = time.time() - t0
    v = steps / dt
    cores = O.num_threads()
    line = {
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": steps,
        "warmup": max(1, min(args.warmup, 2)), "ms_per_step": dt / steps * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "int8 x q4/q6 block dot, f32 accumulate", "data": "synthetic",
        "config": {"workload": f"{args.arch} {args.preset} random-init GGUF, batch-1 greedy decode at ctx {args.ctx}",
                   "note": "CPU oracle port (reference llama.cpp is not buildable offline); steps bounded by a time budget"},
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "port",
                         "sample": f"{steps} greedy decode steps after an {n_prompt}-token prompt, OpenMP {cores} threads"},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)


def eng_pieces(path: str, ids):
    """token ids -> the bytes the engine would stream for them (host tokenizer only)."""
    from localai_b200 import capi

    with capi.Engine(path, vocab_only=True) as e:
        return [e.token_to_piece(int(t)) for t in ids]


def bench_tinyllama(args, rank, local_rank, world, ranks, peak):
    """BASELINE configs[0]: TinyLlama-1.1B-arch Q8_0, batch-1 greedy decode after a 128-token prompt."""
    import numpy as np

    from localai_b200 import capi

    path, _ = ensure_model("tinyllama-1.1b", "q8_0", 0, rank, ranks.barrier)
    with capi.Engine(path, n_ctx=512, n_batch=128, n_parallel=1, device=local_rank) as e:
        ms, launches = e.bench_decode(1, 128, 64, 5)
        k_ms = ranks.max(e.last_mega_ms())
        ms = ranks.max(ms)
        step_bytes = e.info.decode_stream_bytes + e.info.kv_bytes_per_pos * 128
        prompt = (np.arange(128) * 7919 % 4000 + 3).astype(np.int32)
        got, _, _ = e.decode_greedy(prompt, 16)
    out = {"metric": "TinyLlama-1.1B Q8_0 decode tok/s (b=1)", "tok_s": world * 1e3 / ms, "ms_per_step": ms, "launches_per_step": launches,
           "roofline": {"bound": "hbm", "kernel": "decode_mega_kernel", "achieved": step_bytes / ((k_ms or ms) * 1e-3) / 1e9, "peak": peak,
                        "unit": "GB/s", "frac": step_bytes / ((k_ms or ms) * 1e-3) / 1e9 / peak, "bytes_per_launch": step_bytes,
                        "us_per_launch": (k_ms or ms) * 1e3}}
    if rank == 0 and world == 1:
        try:
            from oracle import oracle as O

            O.set_num_threads(os.cpu_count() or 1)
            t0 = time.time()
            want, _, _ = O.OracleModel(path, n_ctx=256).generate_greedy(prompt, 16)
            out["cpu_tok_s"] = (128 + 16) / (time.time() - t0)
            out["parity_checked"] = bool(got.tolist() == want.tolist())
        except Exception as ex:  # noqa: BLE001
            out["parity_checked"] = False
            out["parity_error"] = str(ex)
    return out


def bench_bert_extra(rank, local_rank, world, ranks):
    """BASELINE configs[3]: BERT-base F16 embeddings of 512-token texts: 64 and 16 texts per call, and one text per call (the
    shape of the reference API, backend.proto:141)."""
    import numpy as np

    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import make_bert_gguf as mb
    from localai_b200 import capi

    d = os.environ.get("B200LLM_BENCH_DIR", "/dev/shm/b200llm_bench" if os.path.isdir("/dev/shm") else "/tmp/b200llm_bench")
    os.makedirs(d, exist_ok=True)
    path = os.path.join(d, "bert_base_f16_seed0.gguf")
    if rank == 0 and not os.path.exists(path):
        mb.write_bert_gguf(path + ".tmp", mb.BERT_ARCHS["bert-base"], seed=0)
        os.replace(path + ".tmp", path)
    ranks.barrier()
    a = mb.BERT_ARCHS["bert-base"]
    n_tok = 512
    flops_seq = 2.0 * a.n_layer * (4 * a.n_embd * a.n_embd + 2 * a.n_embd * a.n_ff) * n_tok + 4.0 * n_tok * n_tok * a.n_embd * a.n_layer
    with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
        tf_peak = float(json.load(f).get("bf16_tflops_sustained", 1400.0))
    rng = np.random.default_rng(rank)
    out = {"metric": "BERT-base F16 embeddings seq/s (512-token texts)"}
    with capi.Bert(path, device=local_rank) as b:
        L = capi.lib()
        for per_call, n_calls in ((64, 8), (16, 16), (1, 64)):
            ids = np.ascontiguousarray(np.concatenate([np.concatenate([[2], rng.integers(5, 30000, n_tok - 2), [3]]) for _ in range(per_call)]), dtype=np.int32)
            offs = (np.arange(per_call + 1) * n_tok).astype(np.int32)
            vec = np.empty((per_call, b.n_embd), dtype=np.float32)
            for _ in range(3):
                assert L.b2_bert_embed_batch(b._h, ids.ctypes.data, offs.ctypes.data, per_call, vec.ctypes.data) == per_call
            ranks.barrier()
            t0 = time.perf_counter()
            for _ in range(n_calls):
                L.b2_bert_embed_batch(b._h, ids.ctypes.data, offs.ctypes.data, per_call, vec.ctypes.data)
            dt = ranks.max(time.perf_counter() - t0)
            tf = flops_seq * per_call * n_calls / dt / 1e12
            out["per_call_%d" % per_call] = {"seq_s": world * per_call * n_calls / dt, "ms_per_call": dt * 1e3 / n_calls,
                                             "roofline": {"bound": "tensor", "achieved": tf, "peak": tf_peak, "unit": "TFLOP/s", "frac": tf / tf_peak}}
        if rank == 0 and world == 1:
            try:
                from oracle.bert_ref import BertRef

                want = BertRef(path).embed_tokens(ids[:n_tok])
                got = b.embed_tokens(ids[:n_tok])
                cos = float((got * want).sum() / np.sqrt((got * got).sum() * (want * want).sum()))
                out["parity_checked"] = bool(cos >= 0.9995)
                out["cosine_vs_oracle"] = cos
            except Exception as ex:  # noqa: BLE001
                out["parity_checked"] = False
                out["parity_error"] = str(ex)
    return out


def bench_pool(path, rank, world, ranks):
    """N > 1: ONE process (rank 0) drives a ReplicaPool over all N devices -- one engine per GPU, least-loaded dispatch --
    with 64 x N concurrent greedy streams submitted from the host; the other ranks wait at the barrier."""
    import numpy as np

    out = None
    ranks.host_barrier()
    if rank == 0:
        from localai_b200.replicas import ReplicaPool

        n_streams, n_predict = 64 * world, 128
        pool = ReplicaPool(path, devices=list(range(world)), n_ctx=64 * 256, n_batch=512, n_parallel=64)
        try:
            rng = np.random.default_rng(5)
            prompts = [rng.integers(3, 4000, 16).astype(np.int32) for _ in range(n_streams)]
            # warm-up: the same wave of requests once, short (captures the decode graphs of the batch sizes the wave goes through)
            warm = [pool.submit(prompt_ids=p, n_predict=8, temperature=0.0, ignore_eos=True, stream=False) for p in prompts]
            for h in warm:
                list(pool.events(h))
            pool.assigned = [0] * len(pool.devices)
            t0 = time.perf_counter()
            hs = [pool.submit(prompt_ids=p, n_predict=n_predict, temperature=0.0, ignore_eos=True, stream=False) for p in prompts]
            n_tok = 0
            for h in hs:
                n_tok += list(pool.events(h))[-1]["n_decoded"]
            dt = time.perf_counter() - t0
            out = {"tok_s": n_tok / dt, "streams": n_streams, "n_predict": n_predict, "wall_s": dt, "assigned_per_replica": list(pool.assigned),
                   "path": "one process, ReplicaPool over %d devices, b2_submit / b2_poll with host buffers (prompt + decode inside the timed region)" % world}
        finally:
            pool.close()
    ranks.host_barrier()
    return out



def bench_70b_one_gpu(peak):
    """Llama-3-70B-arch Q4_K_M batch-1 decode on ONE GPU (the baseline of the tensor-parallel scaling curve)."""
    from localai_b200 import capi

    path, _ = ensure_model("llama3-70b", "q4_k_m", 0, 0, lambda: None)
    with capi.Engine(path, n_ctx=512, n_batch=32, n_parallel=1, device=0) as e:
        ms, launches = e.bench_decode(1, 128, 32, 3)
        b = e.info.decode_stream_bytes + e.info.kv_bytes_per_pos * 128
        return {"metric": "Llama-3-70B Q4_K_M decode tok/s (b=1)", "tp": 1,
                "one_gpu": {"tok_s": 1e3 / ms, "ms_per_step": ms, "hbm_frac": b / (ms * 1e-3) / 1e9 / peak, "launches_per_step": launches}}


def bench_tp70b(rank, world, ranks, peak, with_one_gpu=True):
    """BASELINE configs[4]: Llama-3-70B-arch Q4_K_M, batch-1 decode, ONE engine sharded over the N devices (tensor
    parallel: column-parallel wq|wk|wv|gate|up|output, row-parallel wo|down, the partial rows summed after attn_output
    and ffn_down) -- driven by rank 0 alone, the other ranks wait at the barrier.  Both exchange modes are timed: per-layer
    kernels + ncclAllReduce (what north_star names; the time inside the collectives is reported separately) and the
    persistent kernels that store their partial rows into the peers' memory.  The one-GPU number is the scaling baseline."""
    out = None
    ranks.host_barrier()
    if rank == 0:
        from localai_b200 import capi

        path, _ = ensure_model("llama3-70b", "q4_k_m", 0, 0, lambda: None)
        out = {"metric": "Llama-3-70B Q4_K_M decode tok/s (b=1), tensor-parallel", "tp": world}
        ctx, steps = 128, 32
        if with_one_gpu:
            try:
                with capi.Engine(path, n_ctx=512, n_batch=32, n_parallel=1, device=0) as e:
                    ms, _ = e.bench_decode(1, ctx, steps, 3)
                    b = e.info.decode_stream_bytes + e.info.kv_bytes_per_pos * ctx
                    out["one_gpu"] = {"tok_s": 1e3 / ms, "ms_per_step": ms, "hbm_frac": b / (ms * 1e-3) / 1e9 / peak}
            except Exception as ex:  # noqa: BLE001
                out["one_gpu"] = {"error": str(ex)}
        with capi.Engine(path, n_ctx=512, n_batch=32, n_parallel=1, device=0, tp_size=world) as e:
            for mode, name in ((capi.B2_TP_NCCL, "nccl"), (capi.B2_TP_FUSED, "fused")):
                try:
                    e.tp_set_mode(mode)
                    ms, launches = e.bench_decode(1, ctx, steps, 3)
                    info = e.tp_info()
                    b = info["shard_stream_bytes"] + e.info.kv_bytes_per_pos * ctx  # per GPU
                    out[name] = {"tok_s": 1e3 / ms, "ms_per_step": ms, "hbm_frac_per_gpu": b / (ms * 1e-3) / 1e9 / peak,
                                 "n_allreduce": info["n_allreduce_per_step"], "launches_per_step": launches}
                    if mode == capi.B2_TP_NCCL:
                        out[name]["allreduce_us_per_token"] = info["allreduce_ms_per_step"] * 1e3
                except Exception as ex:  # noqa: BLE001
                    out[name] = {"error": str(ex)}
    ranks.host_barrier()
    return out


def run_bert(args, rank, local_rank, world, ranks, emit, warmup):
    """BASELINE config 4: BERT-base F16, texts of 512 WordPiece tokens, through b2_bert_embed_batch (host ids in, host
    vectors out -- the call is synchronous, so every timed step holds its own H2D and D2H copies).  A step = one call
    with 64 sequences (32768 tokens: the activations of a step are 100-400 MB per tensor, nothing the L2 could keep)."""
    import numpy as np

    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import make_bert_gguf as mb
    from localai_b200 import capi

    d = os.environ.get("B200LLM_BENCH_DIR", "/dev/shm/b200llm_bench" if os.path.isdir("/dev/shm") else "/tmp/b200llm_bench")
    os.makedirs(d, exist_ok=True)
    path = os.path.join(d, "bert_base_f16_seed0.gguf")
    if rank == 0 and not os.path.exists(path):
        mb.write_bert_gguf(path + ".tmp", mb.BERT_ARCHS["bert-base"], seed=0)
        os.replace(path + ".tmp", path)
    ranks.barrier()
    per_call, n_tok = 64, 512
    rng = np.random.default_rng(rank)
    b = capi.Bert(path, device=local_rank)
    ids = np.ascontiguousarray(np.concatenate([np.concatenate([[2], rng.integers(5, 300, n_tok - 2), [3]]) for _ in range(per_call)]), dtype=np.int32)
    offs = (np.arange(per_call + 1) * n_tok).astype(np.int32)
    out = np.empty((per_call, b.n_embd), dtype=np.float32)
    L = capi.lib()

    def step():
        rc = L.b2_bert_embed_batch(b._h, ids.ctypes.data, offs.ctypes.data, per_call, out.ctypes.data)
        assert rc == per_call, rc

    for _ in range(warmup):
        step()
    ranks.barrier()
    with ClockSampler(local_rank) as clk:
        t0 = time.perf_counter()
        for _ in range(args.steps):
            step()
        dt = time.perf_counter() - t0
    ranks.barrier()
    dt = ranks.max(dt)
    seqs = ranks.sum(float(per_call * args.steps))
    if rank != 0:
        b.close()
        return 0
    a = mb.BERT_ARCHS["bert-base"]
    lin = 2.0 * a.n_layer * (4 * a.n_embd * a.n_embd + 2 * a.n_embd * a.n_ff) * n_tok
    att = 4.0 * n_tok * n_tok * a.n_embd * a.n_layer
    flops_seq = lin + att
    with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
        tf_peak = float(json.load(f).get("bf16_tflops_sustained", 1392.2))
    achieved = flops_seq * per_call * args.steps / dt / 1e12  # per GPU: every rank runs the same number of texts
    cpu = None
    try:  # the numpy restatement on ONE 512-token text (about 2 s per text on a few cores)
        from oracle.bert_ref import BertRef

        ref = BertRef(path)
        t1 = time.perf_counter()
        ref.embed_tokens(ids[:n_tok])
        cpu_dt = time.perf_counter() - t1
        cpu = {"value": 1.0 / cpu_dt, "unit": "seq/s", "cores": os.cpu_count(), "kind": "port",
               "sample": "1 text of 512 tokens through oracle/bert_ref.py (numpy float32, BLAS threads)"}
    except Exception as e:  # noqa: BLE001
        log(f"[bench] cpu baseline skipped: {e}")
    v = seqs / dt
    launches_per_step = 2 + a.n_layer * 10 + 2
    emit({"metric": "BERT-base F16 embeddings seq/s (512-token texts)", "value": v, "unit": "seq/s", "n_gpus": world, "steps": args.steps,
          "warmup": warmup, "ms_per_step": dt * 1e3 / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
          "dtype": "fp16 GEMM operands (the file's F16 weights exactly), bf16 attention operands, f32 accumulate", "data": "synthetic",
          "config": {"workload": "bert-base f16 random-init GGUF, 64 texts x 512 tokens per b2_bert_embed_batch call", "replicas": world,
                     "l2": "inputs larger than L2: a step moves 25-100 MB per activation tensor, 24 tensors per layer pass"},
          "e2e": {"value": v, "unit": "seq/s", "h2d_bytes_per_step": int(ids.nbytes + offs.nbytes), "d2h_bytes_per_step": int(out.nbytes),
                  "path": "b2_bert_embed_batch (host ids -> host vectors, synchronous); value is this same measurement"},
          "gpu_launches": launches_per_step * args.steps, "gpu_launches_per_step": launches_per_step,
          "roofline": {"bound": "tensor", "kernel": "whole step (gemm_tcgen05_kernel x 4 + attention per layer)", "achieved": achieved,
                       "peak": tf_peak, "unit": "TFLOP/s", "frac": achieved / tf_peak, "traffic": None,
                       "flops_per_seq": flops_seq, "peak_source": "measured (MEASURED_PEAKS.json bf16_tflops_sustained)"},
          "cpu_baseline": cpu, "clocks": clk.summary()})
    b.close()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=256)
    ap.add_argument("--warmup", type=int, default=8)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--arch", default="llama3-8b")
    ap.add_argument("--preset", default="q4_k_m")
    ap.add_argument("--seed", type=int, default=0)
    ap.add_argument("--ctx", type=int, default=128, help="KV positions already in the cache for the decode legs")
    ap.add_argument("--prefill", type=int, default=2048)
    ap.add_argument("--batch", type=int, default=64, help="concurrent streams for the batched-decode extra")
    ap.add_argument("--no-extras", action="store_true", help="skip the b=64 / prefill / cpu legs")
    ap.add_argument("--with-70b", action="store_true", help="N = 1: also time Llama-3-70B Q4_K_M b=1 decode on the one GPU (a 42.5 GB file is generated)")
    ap.add_argument("--workload", default="decode", choices=["decode", "bert"],
                    help="decode: the headline (Llama-3-8B Q4_K_M b=1 decode); bert: BASELINE config 4, BERT-base F16 embeddings of 512-token texts")
    args = ap.parse_args()

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    # stdout carries exactly ONE JSON line: libraries that print there (NCCL's version banner) go to stderr instead
    sys.stdout.flush()
    json_fd = os.dup(1)
    os.dup2(2, 1)

    def emit(line: dict):
        os.write(json_fd, (json.dumps(line) + "\n").encode())

    if args.impl == "reference":
        run_reference(args, rank, emit)
        return 0

    warmup = max(args.warmup, 3)
    from localai_b200.replicas import Ranks

    ranks = Ranks(rank, world, local_rank)  # nccl; one independent replica per rank, no data-path collective
    barrier, max_over_ranks = ranks.barrier, ranks.max

    from localai_b200 import capi

    if capi.device_count() < 1:
        raise SystemExit("bench.py: no sm_100 device visible; there is no CPU fallback for the b200 arm")
    if args.workload == "bert":
        return run_bert(args, rank, local_rank, world, ranks, emit, warmup)
    path, info = ensure_model(args.arch, args.preset, args.seed, rank, barrier)
    n_par = max(1, args.batch)
    n_ctx_slot = max(args.prefill, args.ctx + args.steps + warmup) + 128
    t0 = time.time()
    n_batch = max(512, args.prefill)  # NBatch (ModelOptions): one chunk per prompt, weights are dequantised once per chunk
    eng = capi.Engine(path, n_ctx=n_ctx_slot * n_par, n_batch=n_batch, n_parallel=n_par, device=local_rank)
    log(f"[bench r{rank}] loaded in {time.time() - t0:.1f}s: {eng.info.name.decode()} weights {eng.info.weight_bytes_device / 1e9:.3f} GB")
    stream_bytes = eng.info.decode_stream_bytes
    kv_per_pos = eng.info.kv_bytes_per_pos
    peak, peak_src = peaks()

    # ---- value: device-resident batch-1 decode
    barrier()
    with ClockSampler(local_rank) as clk:
        ms, launches = eng.bench_decode(1, args.ctx, args.steps, warmup)
    ms = max_over_ranks(ms)
    barrier()
    value = world * 1e3 / ms
    step_bytes = stream_bytes + kv_per_pos * args.ctx
    clocks = clk.summary()

    # ---- e2e: the PredictStream path (host ids in, text out)
    import numpy as np

    prompt = (np.arange(args.ctx) * 7919 % 4000 + 3).astype(np.int32)
    rid = eng.submit(prompt_ids=prompt, n_predict=min(32, args.steps), temperature=0.0, ignore_eos=True, stream=True)
    list(eng.events(rid))  # warm-up request
    eng.release(rid)
    barrier()
    t0 = time.time()
    rid = eng.submit(prompt_ids=prompt, n_predict=args.steps, temperature=0.0, ignore_eos=True, stream=True)
    evs = list(eng.events(rid))
    t_wall = time.time() - t0
    eng.release(rid)
    fin = evs[-1]
    e2e_ms = max_over_ranks(fin["t_gen_ms"] / max(1, fin["n_decoded"] - 1))
    e2e = {"value": world * 1e3 / e2e_ms, "unit": UNIT, "h2d_bytes_per_step": 12 + 584, "d2h_bytes_per_step": 8,  # token/pos/slot + SampleSpec up; token id + mirostat mu down
           "prompt_ms": fin["t_prompt_ms"], "wall_s": t_wall, "n_decoded": fin["n_decoded"],
           "path": "b2_submit/b2_poll (PredictStream), greedy, host prompt ids"}

    # ---- roofline of the dominant kernel.  A b=1 decode step IS one launch of decode_mega_kernel (the persistent
    # kernel that walks every layer): its launches were timed one by one with CUDA events on the engine's stream
    # during the timed region above; its algorithmic bytes are the step's: every weight byte once + the KV rows read.
    mega_ms = max_over_ranks(eng.last_mega_ms())
    k_ms = mega_ms if mega_ms > 0 else ms  # (the fused per-layer kernels took the step: whole-step time)
    achieved = step_bytes / (k_ms * 1e-3) / 1e9
    traffic, traffic_src = mega_traffic()
    roofline = {"bound": "hbm", "kernel": "decode_mega_kernel (one launch per decode step; CUDA events around every launch)" if mega_ms > 0 else
                "fused decode kernels (%d launches per step), whole step" % launches,
                "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_src,
                "bytes_per_launch": step_bytes, "us_per_launch": k_ms * 1e3, "step_us": ms * 1e3, "kernel_share_of_step": k_ms / ms,
                "peak_source": peak_src, "frac_of_8tbs": step_bytes / (k_ms * 1e-3) / 8e12}

    extra = {}
    cpu = None
    parity_checked = None
    if not args.no_extras:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            tf_peak = float(json.load(f).get("bf16_tflops_sustained", 1400.0))
        i = eng.info
        kvd = i.n_head_kv * i.head_dim
        lin_params = i.n_layer * (2 * i.n_embd * i.n_embd + 2 * i.n_embd * kvd + 3 * i.n_embd * i.n_ff)  # mat-mul weights of the layers
        head_params = i.n_vocab * i.n_embd
        try:  # configs[2]: 64 concurrent streams, one step = one token for each
            bms, bl = eng.bench_decode(n_par, args.ctx, max(8, args.steps // 8), 3)
            bms = max_over_ranks(bms)
            b_bytes = stream_bytes + n_par * kv_per_pos * args.ctx
            b_flops = 2.0 * (lin_params + head_params) * n_par
            extra["decode_b%d" % n_par] = {
                "tok_s": world * n_par * 1e3 / bms, "ms_per_step": bms, "launches_per_step": bl,
                "kernel": "qgemm_kernel (2..64 tokens: GGUF blocks dequantised inside the tcgen05 GEMM, fp16 operand in TMEM; "
                          "B200LLM_NO_QGEMM=1 selects the fp16-tile GEMM)" if n_par <= 64 and not os.environ.get("B200LLM_NO_QGEMM") else "gemm_tcgen05_kernel (fp16 tiles)",
                "hbm": {"bound": "hbm", "achieved": b_bytes / (bms * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                        "frac": b_bytes / (bms * 1e-3) / 1e9 / peak},
                "tensor": {"bound": "tensor", "achieved": b_flops / (bms * 1e-3) / 1e12, "peak": tf_peak, "unit": "TFLOP/s",
                           "frac": b_flops / (bms * 1e-3) / 1e12 / tf_peak}}
        except capi.B2Error as ex:
            extra["decode_b%d" % n_par] = {"error": str(ex)}
        try:  # 2048-token prefill
            pms, pl = eng.bench_prefill(args.prefill, 2, 1)
            pms = max_over_ranks(pms)
            n = args.prefill
            p_flops = 2.0 * lin_params * n + 2.0 * n * n * i.n_embd * i.n_layer + 2.0 * head_params  # causal attention: half of 4 n^2 d
            extra["prefill_%d" % n] = {"tok_s": world * n * 1e3 / pms, "ms": pms, "launches": pl,
                                       "roofline": {"bound": "tensor", "achieved": p_flops / (pms * 1e-3) / 1e12, "peak": tf_peak,
                                                    "unit": "TFLOP/s", "frac": p_flops / (pms * 1e-3) / 1e12 / tf_peak}}
        except capi.B2Error as ex:
            extra["prefill_%d" % args.prefill] = {"error": str(ex)}
        if rank == 0 and world == 1:
            # ---- the checker, after every timed region: the oracle continues the e2e request's prompt greedily; its
            # first 16 tokens must be the text the engine streamed.  The same run is the CPU baseline sample.
            try:
                from oracle import oracle as O

                O.set_num_threads(os.cpu_count() or 1)
                n_chk = max(1, min(16, fin["n_decoded"]))
                n_dec = max(n_chk, min(args.steps, 20))
                m = O.OracleModel(path, n_ctx=args.ctx + n_dec + 8)
                t1 = time.time()
                lg = m.eval(prompt)
                t2 = time.time()
                ids = []
                for k in range(n_dec):
                    ids.append(int(lg[0].argmax()))
                    lg = m.eval([ids[-1]], n_past=args.ctx + k)
                t3 = time.time()
                want_text = b"".join(eng_pieces(path, ids[:n_chk]))
                got_text = b"".join(ev["text"] for ev in evs)
                parity_checked = bool(got_text[:len(want_text)] == want_text and len(want_text) > 0)
                cores = O.num_threads()
                cpu = {"value": n_dec / (t3 - t2), "unit": UNIT, "cores": cores, "kind": "port", "prompt_tok_s": args.ctx / (t2 - t1),
                       "sample": f"oracle (C restatement of the llama.cpp CPU path: scalar block dots under OpenMP, {cores} threads; NOT "
                                 f"llama.cpp's AVX2/AVX-512 build, which cannot be built offline): {args.ctx}-token prompt + {n_dec} greedy "
                                 f"decode steps at ctx {args.ctx} on the same GGUF"}
                del m
            except Exception as ex:  # the checker failing must not void the GPU measurement
                cpu = {"value": None, "unit": UNIT, "cores": os.cpu_count(), "kind": "port", "sample": f"failed: {ex}"}
                parity_checked = False
    eng.close()
    if not args.no_extras:
        try:
            extra["tinyllama_q8_0_b1"] = bench_tinyllama(args, rank, local_rank, world, ranks, peak)
        except Exception as ex:  # noqa: BLE001
            extra["tinyllama_q8_0_b1"] = {"error": str(ex)}
        try:
            extra["bert_base_256x512"] = bench_bert_extra(rank, local_rank, world, ranks)
        except Exception as ex:  # noqa: BLE001
            extra["bert_base_256x512"] = {"error": str(ex)}
        if world > 1:
            try:
                extra["pool"] = bench_pool(path, rank, world, ranks)
            except Exception as ex:  # noqa: BLE001
                extra["pool"] = {"error": str(ex)}
        if world > 1 or args.with_70b:
            try:
                extra["tp70b"] = bench_tp70b(rank, world, ranks, peak) if world > 1 else bench_70b_one_gpu(peak)
            except Exception as ex:  # noqa: BLE001
                extra["tp70b"] = {"error": str(ex)}
    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": warmup,
            "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "int8 x q4/q6 block dot (dp4a), f32 accumulate", "data": "synthetic",
            "config": {"workload": f"{args.arch} {args.preset} random-init GGUF, batch-1 greedy decode at ctx {args.ctx}",
                       "replicas": world, "l2": "inputs larger than L2: every step streams %.2f GB of weights (L2 = 126 MB)" % (stream_bytes / 1e9),
                       "decode_stream_bytes": stream_bytes, "kv_bytes_per_pos": kv_per_pos, "n_batch": n_batch,
                       "n_parallel": n_par},
            "e2e": e2e, "gpu_launches": int(launches) * args.steps, "gpu_launches_per_step": int(launches),
            "roofline": roofline, "cpu_baseline": cpu, "parity_checked": parity_checked, "clocks": clocks, "extra": extra,
        }
        emit(line)
    ranks.close()
    return 0


if __name__ == "__main__":
    sys.exit(main())
