import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu on the GPU box)")


def have_gpu() -> bool:
    try:
        from localai_b200 import capi
        return capi.device_count() > 0
    except Exception:
        return False


@pytest.fixture(scope="session")
def gguf_dir(tmp_path_factory):
    return str(tmp_path_factory.mktemp("gguf"))


@pytest.fixture(scope="session")
def model_files(gguf_dir):
    """Small random-init GGUF models covering every quant type and both tokenisers."""
    import make_gguf as mg

    specs = {
        "micro_mixed": ("micro", "mixed"),
        "micro_q8_0": ("micro", "q8_0"),
        "mini_q4_k_m": ("mini", "q4_k_m"),
        "mini_q5_k": ("mini", "q5_k"),
        "mini_f16": ("mini", "f16"),
        "mini_q4_0": ("mini", "q4_0"),
        "mini_q5_1": ("mini", "q5_1"),
        "mini_q2_k": ("mini", "q2_k"),
        "mini_q3_k": ("mini", "q3_k"),
        "mini_iq4_nl": ("mini", "iq4_nl"),
        "mini_iq4_xs": ("mini", "iq4_xs"),
        # RoPE scaling variants (linear, YaRN, per-dimension divisors a la Llama-3.1)
        "micro_linear": ("micro-linear", "mixed"),
        "micro_yarn": ("micro-yarn", "mixed"),
        "micro_l31": ("micro-l31", "mixed"),
        "micro_g8": ("micro-g8", "q4_k_m"),  # 8 query heads per KV head
        "mini_qwen2": ("mini-qwen2", "q4_k_m"),  # qwen2: QKV biases, NEOX rotary pairs, its own pre-tokeniser
        "mini_qwen2_g7": ("mini-qwen2-g7", "q4_k_m"),  # 7 query heads per KV head: no grouped attention kernel for it
        "mini_phi3": ("mini-phi3", "q4_k_m"),  # phi3: fused attn_qkv and gate|up tensors, NEOX rotary pairs
    }
    out = {}
    for name, (arch, preset) in specs.items():
        path = os.path.join(gguf_dir, name + ".gguf")
        mg.write_gguf(path, mg.ARCHS[arch], preset, seed=7, quiet=True)
        out[name] = path
    return out
