"""Tensor parallelism (BASELINE config 5: Llama-3-70B-arch Q4_K_M over 2 / 4 / 8 GPUs) on a
70B-shaped small model (the 70B type mix: Q4_K, Q5_K / Q6_K attn_v, Q6_K ffn_down on the "more bits" layers; 8 KV heads).

The sharded engine (b2_load_opts.tp_size) must compute what the one-GPU engine and the oracle compute: logits within the
stated tolerance (the shards add K-slice partial rows in rank order, so only the fp32 summation order differs from one
GPU), greedy ids identical -- in both exchange modes: per-layer kernels + ncclAllReduce, and the persistent kernels that
store their partial rows into the peers' memory.  Skipped below 2 devices."""
import os

import numpy as np
import pytest

from localai_b200 import capi
from oracle import oracle as O
from test_parity_gpu import assert_logits_close

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def mini70b(gguf_dir):
    import make_gguf as mg
    path = os.path.join(gguf_dir, "mini70b_q4_k_m.gguf")
    mg.write_gguf(path, mg.ARCHS["mini-70b"], "q4_k_m", seed=5, quiet=True)
    return path


def _tp_sizes():
    n = capi.device_count()
    return [t for t in (2, 4, 8) if t <= n]


@pytest.mark.parametrize("mode", [capi.B2_TP_NCCL, capi.B2_TP_FUSED])
def test_tensor_parallel_matches_one_gpu_and_oracle(mini70b, mode):
    sizes = _tp_sizes()
    if not sizes:
        pytest.skip("needs at least 2 sm_100 devices")
    ref = O.OracleModel(mini70b, n_ctx=128)
    prompt = np.random.default_rng(70).integers(0, 8000, 24).astype(np.int32)
    want_logits = ref.eval(prompt, logits_all=True)
    want_ids, _, margins = ref.generate_greedy(prompt, 32)
    with capi.Engine(mini70b, n_ctx=128, n_batch=32) as e1:
        one = e1.eval(prompt, logits_all=True)
    assert_logits_close(one, want_logits, "one GPU")
    for tp in sizes:
        with capi.Engine(mini70b, n_ctx=128, n_batch=32, tp_size=tp) as e:
            e.tp_set_mode(mode)
            info = e.tp_info()
            assert info["tp_size"] == tp and info["mode"] == mode and info["n_allreduce_per_step"] == 8
            got = e.eval(prompt, logits_all=True)
            assert_logits_close(got, want_logits, f"tp={tp} mode={mode} vs oracle")
            assert_logits_close(got, one, f"tp={tp} mode={mode} vs one GPU")
            ids, _, _ = e.decode_greedy(prompt, 32)
            assert ids.tolist() == want_ids.tolist(), f"tp={tp} mode={mode}"
            # the request path on a sharded engine: same text as on one GPU
            rid = e.submit(prompt_ids=prompt, n_predict=12, temperature=0.0, ignore_eos=True, stream=False)
            text = b"".join(ev["text"] for ev in e.events(rid))
            e.release(rid)
        with capi.Engine(mini70b, n_ctx=128, n_batch=32) as e1:
            rid = e1.submit(prompt_ids=prompt, n_predict=12, temperature=0.0, ignore_eos=True, stream=False)
            assert text == b"".join(ev["text"] for ev in e1.events(rid))
        print(f"tp={tp} mode={mode}: max |dlogit| vs one GPU {np.abs(got - one).max() / one.std():.4f} sigma, min greedy margin {margins.min():.2f}")


def test_tensor_parallel_bench_hooks(mini70b):
    sizes = _tp_sizes()
    if not sizes:
        pytest.skip("needs at least 2 sm_100 devices")
    with capi.Engine(mini70b, n_ctx=256, n_batch=32, tp_size=sizes[0]) as e:
        for mode in (capi.B2_TP_NCCL, capi.B2_TP_FUSED):
            e.tp_set_mode(mode)
            ms, launches = e.bench_decode(1, 64, 8, 3)
            info = e.tp_info()
            assert ms > 0
            if mode == capi.B2_TP_NCCL:
                assert 0 < info["allreduce_ms_per_step"] < ms
