#!/usr/bin/env python3
"""bench.py -- BASELINE.json's metric on BASELINE.json's config, one JSON line on stdout.

Workload (config.workload): configs[1] = Llama-3-8B-arch GGUF Q4_K_M, random-init, batch-1 greedy decode
(the north-star number) at a short context; the same run also reports batch-64 decode and the 2048-token
prefill as `extra`.  A "step" is one decode step (one token for every active sequence).

  value     decode tok/s with everything resident in HBM (b2_bench_decode: CUDA events on the engine's stream)
  e2e       the same metric through the request path a LocalAI backend uses (b2_submit / b2_poll, i.e.
            PredictStream): host prompt ids in, token text out, per-step H2D/D2H inside the timed region;
            tok/s = n_decoded / t_generation as the reference defines it (grpc-server.cpp:305-318)
  roofline  the dominant kernel (streaming dequant-GEMV on the Q4_K ffn_gate matrices) timed alone with CUDA
            events; achieved = rows x row_bytes / time against MEASURED_PEAKS.json hbm_gbs
  cpu_baseline  the oracle (C restatement of the llama.cpp CPU path, OpenMP) on a bounded sample, same GGUF

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
# dram__bytes_read.sum + dram__bytes_write.sum of one decode_mega_kernel launch (llama3-8b q4_k_m, ctx 128), from the
# ncu --set full capture summarised in profiles/ (None until a capture of the current kernel is committed)
MEGA_TRAFFIC_BYTES = 4.6731e9  # profiles/r1_mega_v6_ncu_summary.txt: 4.6391 GB read + 34.1 MB written


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


def run_bert(args, rank, local_rank, world, ranks, emit, warmup):
    """BASELINE config 4: BERT-base F16, texts of 512 WordPiece tokens, through b2_bert_embed_batch (host ids in, host
    vectors out -- the call is synchronous, so every timed step holds its own H2D and D2H copies).  A step = one call
    with 16 sequences (8192 tokens, larger than nothing L2 could keep: activations of a step are 25-100 MB per tensor)."""
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
    per_call, n_tok = 16, 512
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
          "dtype": "bf16 operands, f32 accumulate (weights F16 in the file)", "data": "synthetic",
          "config": {"workload": "bert-base f16 random-init GGUF, 16 texts x 512 tokens per b2_bert_embed_batch call", "replicas": world,
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
    # kernel that walks every layer), so its launch duration is the step time measured above with CUDA events on the
    # engine's stream, and its algorithmic bytes are the step's: every weight row once + the KV rows read.
    # `traffic` is dram__bytes_read.sum + dram__bytes_write.sum of one launch from profiles/r1_mega_ncu_raw.csv.
    achieved = step_bytes / (ms * 1e-3) / 1e9
    k_ms, k_bytes, k_type = eng.bench_gemv(4, 1, 8)  # the row-streaming loop alone on one ffn_gate matrix, for reference
    roofline = {"bound": "hbm", "kernel": "decode_mega_kernel (1 launch per decode step)" if launches == 1 else
                "fused decode kernels (%d launches per step), whole step" % launches,
                "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": MEGA_TRAFFIC_BYTES,
                "bytes_per_launch": step_bytes, "us_per_launch": ms * 1e3, "peak_source": peak_src,
                "frac_of_8tbs": step_bytes / (ms * 1e-3) / 8e12,
                "gemv_alone": {"kernel": f"gemv_kernel<type {k_type}, NT=1> on ffn_gate [{eng.info.n_ff}x{eng.info.n_embd}]",
                               "bytes_per_launch": k_bytes, "us_per_launch": k_ms * 1e3,
                               "achieved": k_bytes / (k_ms * 1e-3) / 1e9, "frac": k_bytes / (k_ms * 1e-3) / 1e9 / peak}}

    extra = {}
    cpu = None
    if not args.no_extras:
        try:
            bms, bl = eng.bench_decode(n_par, args.ctx, max(8, args.steps // 8), 3)
            bms = max_over_ranks(bms)
            extra["decode_b%d" % n_par] = {"tok_s": world * n_par * 1e3 / bms, "ms_per_step": bms, "launches_per_step": bl,
                                           "hbm_frac_of_measured": (stream_bytes + n_par * kv_per_pos * args.ctx) / (bms * 1e-3) / 1e9 / peak}
        except capi.B2Error as ex:
            extra["decode_b%d" % n_par] = {"error": str(ex)}
        try:
            pms, pl = eng.bench_prefill(args.prefill, 2, 1)
            pms = max_over_ranks(pms)
            extra["prefill_%d" % args.prefill] = {"tok_s": world * args.prefill * 1e3 / pms, "ms": pms, "launches": pl}
        except capi.B2Error as ex:
            extra["prefill_%d" % args.prefill] = {"error": str(ex)}
        if rank == 0 and world == 1:
            try:
                dv, pv, cores = cpu_sample(path, 8, 8)
                cpu = {"value": dv, "unit": UNIT, "cores": cores, "kind": "port", "prompt_tok_s": pv,
                       "sample": "oracle (C restatement of the llama.cpp CPU path, OpenMP): 8-token prompt + 8 greedy decode steps on the same GGUF"}
            except Exception as ex:  # the checker failing must not void the GPU measurement
                cpu = {"value": None, "unit": UNIT, "cores": os.cpu_count(), "kind": "port", "sample": f"failed: {ex}"}
    eng.close()
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
            "roofline": roofline, "cpu_baseline": cpu, "clocks": clocks, "extra": extra,
        }
        emit(line)
    ranks.close()
    return 0


if __name__ == "__main__":
    sys.exit(main())
