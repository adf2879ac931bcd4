#!/usr/bin/env python3
"""Second opinion on the oracle's forward pass: the same GGUF through HuggingFace transformers.

`transformers.modeling_gguf_pytorch_utils.load_gguf_checkpoint` dequantises the tensors with gguf-py and un-permutes
Q/K; `LlamaForCausalLM` (fp32, CPU, rotate-half RoPE, SDPA attention) is an independent statement of the llama graph:
RoPE pairing, GQA, norm placement, SwiGLU, the LM head.  The oracle (oracle/ggml_ref.c) restates llama.cpp's CPU path,
which quantises activations to Q8_K / Q8_0 before every integer dot and rounds K/V and the soft-max to f16; transformers
keeps everything in fp32.  So the F16 file must agree to fp16-rounding precision and the quantised files at the
int8-activation noise floor.  Writes tests/golden/transformers_crosscheck.json (the measured agreement), which
tests/test_oracle.py::test_oracle_matches_transformers re-measures and bounds.

Usage: python tools/crosscheck_transformers.py [--out tests/golden/transformers_crosscheck.json]
"""
import argparse
import json
import os
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))


def hf_llama_logits(path: str, ids, rope_scaling=None) -> np.ndarray:
    """logits [n_tok][n_vocab] of the GGUF at `path` from transformers' LlamaForCausalLM in fp32 on the CPU.
    rope_scaling: the HF config entry that says what the GGUF's rope.scaling.* metadata / rope_freqs.weight say (transformers'
    GGUF reader does not map them)"""
    import torch
    from transformers import LlamaConfig, LlamaForCausalLM
    from transformers.modeling_gguf_pytorch_utils import load_gguf_checkpoint

    cfg = load_gguf_checkpoint(path, return_tensors=False)["config"]
    torch.manual_seed(0)
    if cfg.get("model_type") == "qwen2":  # QKV biases, rotate-half RoPE on un-permuted weights
        from transformers import Qwen2Config, Qwen2ForCausalLM

        model = Qwen2ForCausalLM(Qwen2Config(**{k: v for k, v in cfg.items() if k != "model_type"})).to(torch.float32).eval()
    elif cfg.get("model_type") == "phi3":  # fused qkv_proj / gate_up_proj tensors, rotate-half RoPE
        from transformers import Phi3Config, Phi3ForCausalLM

        model = Phi3ForCausalLM(Phi3Config(**{k: v for k, v in cfg.items() if k != "model_type"}, pad_token_id=0)).to(torch.float32).eval()
    else:
        extra = {"rope_scaling": rope_scaling} if rope_scaling else {}
        model = LlamaForCausalLM(LlamaConfig(**{k: v for k, v in cfg.items() if k != "model_type"}, **extra)).to(torch.float32).eval()
    tensors = load_gguf_checkpoint(path, return_tensors=True, model_to_load=model)["tensors"]
    missing, unexpected = model.load_state_dict(tensors, strict=False)
    assert not missing and not unexpected, (missing, unexpected)
    with torch.no_grad():
        return model(torch.tensor(np.asarray(ids, dtype=np.int64))[None]).logits[0].numpy().astype(np.float32)


def hf_bert_hidden(path: str, ids) -> np.ndarray:
    """last hidden states [n_tok][n_embd] of a BERT GGUF (tools/make_bert_gguf.py naming) from transformers' BertModel."""
    import gguf
    import torch
    from transformers import BertConfig, BertModel

    r = gguf.GGUFReader(path)
    kv = {k: v for k, v in r.fields.items()}

    def val(key):
        f = kv[key]
        return f.parts[f.data[0]][0]

    t = {x.name: gguf.quants.dequantize(x.data, x.tensor_type).astype(np.float32).reshape([int(n) for n in reversed(x.shape)]) for x in r.tensors}
    n_layer, d = int(val("bert.block_count")), int(val("bert.embedding_length"))
    cfg = BertConfig(vocab_size=t["token_embd.weight"].shape[0], hidden_size=d, num_hidden_layers=n_layer,
                     num_attention_heads=int(val("bert.attention.head_count")), intermediate_size=int(val("bert.feed_forward_length")),
                     max_position_embeddings=t["position_embd.weight"].shape[0], type_vocab_size=t["token_types.weight"].shape[0],
                     layer_norm_eps=float(val("bert.attention.layer_norm_epsilon")), hidden_act="gelu_new",  # ggml_gelu is the tanh form
                     hidden_dropout_prob=0.0, attention_probs_dropout_prob=0.0)
    m = BertModel(cfg, add_pooling_layer=False).to(torch.float32).eval()
    sd = {"embeddings.word_embeddings.weight": t["token_embd.weight"], "embeddings.position_embeddings.weight": t["position_embd.weight"],
          "embeddings.token_type_embeddings.weight": t["token_types.weight"], "embeddings.LayerNorm.weight": t["token_embd_norm.weight"],
          "embeddings.LayerNorm.bias": t["token_embd_norm.bias"]}
    names = {"attn_q": "attention.self.query", "attn_k": "attention.self.key", "attn_v": "attention.self.value",
             "attn_output": "attention.output.dense", "attn_output_norm": "attention.output.LayerNorm", "ffn_up": "intermediate.dense",
             "ffn_down": "output.dense", "layer_output_norm": "output.LayerNorm"}
    for l in range(n_layer):
        for g, h in names.items():
            for suffix in ("weight", "bias"):
                sd[f"encoder.layer.{l}.{h}.{suffix}"] = t[f"blk.{l}.{g}.{suffix}"]
    missing, unexpected = m.load_state_dict({k: torch.from_numpy(np.ascontiguousarray(v)) for k, v in sd.items()}, strict=False)
    assert not unexpected and all("position_ids" in k for k in missing), (missing, unexpected)
    with torch.no_grad():
        out = m(input_ids=torch.tensor(np.asarray(ids, dtype=np.int64))[None]).last_hidden_state[0]
    return out.numpy().astype(np.float32)


def measure() -> dict:
    import make_bert_gguf as mb
    import make_gguf as mg
    from oracle import bert_ref as B
    from oracle import oracle as O

    res = {}
    rng = np.random.default_rng(5)
    with tempfile.TemporaryDirectory() as td:
        for arch, preset in (("mini", "f16"), ("mini", "q4_k_m"), ("mini", "q8_0"), ("mini-qwen2", "f16"), ("mini-qwen2", "q4_k_m"),
                             ("mini-phi3", "f16"), ("mini-phi3", "q4_k_m")):
            path = os.path.join(td, f"{arch}_{preset}.gguf")
            try:
                mg.write_gguf(path, mg.ARCHS[arch], preset, seed=7, quiet=True)
            except Exception as ex:  # preset not defined for this arch
                res[f"llama_{arch}_{preset}".replace("llama_mini-qwen2", "qwen2_mini").replace("llama_mini-phi3", "phi3_mini")] = {"skipped": str(ex)}
                continue
            ids = rng.integers(0, 4000, 32)
            want = hf_llama_logits(path, ids)
            got = O.OracleModel(path, n_ctx=64).eval(ids, logits_all=True)
            sig = float(want.std())
            err = np.abs(got - want)
            res[f"llama_{arch}_{preset}".replace("llama_mini-qwen2", "qwen2_mini").replace("llama_mini-phi3", "phi3_mini")] = {
                "n_tokens": int(len(ids)), "sigma_logits": sig, "max_err_sigma": float(err.max() / sig),
                "median_err_sigma": float(np.median(err) / sig), "rel_to_max": float(err.max() / np.abs(want).max()),
                "argmax_agree": int((got.argmax(-1) == want.argmax(-1)).sum())}
        # RoPE scaling: linear, YaRN, and the per-dimension divisors of Llama-3.1 (rope_freqs.weight) against transformers'
        # own implementations of the three (modeling_rope_utils: "linear", "yarn", "llama3"); 96 positions, so that the
        # scaled rotations differ from the unscaled ones by more than a sigma of the logits
        for arch, rs in (("micro-linear", {"rope_type": "linear", "factor": mg.ARCHS["micro-linear"].rope_factor}),
                         ("micro-yarn", {"rope_type": "yarn", "factor": mg.ARCHS["micro-yarn"].rope_factor,
                                         "original_max_position_embeddings": mg.ARCHS["micro-yarn"].rope_orig_ctx}),
                         ("micro-l31", {"rope_type": "llama3", "factor": 8.0, "low_freq_factor": 1.0, "high_freq_factor": 4.0,
                                        "original_max_position_embeddings": mg.ARCHS["micro-l31"].rope_orig_ctx or 64})):
            path = os.path.join(td, f"{arch}_f16.gguf")
            mg.write_gguf(path, mg.ARCHS[arch], "f16", seed=7, quiet=True)
            ids = rng.integers(0, 4000, 96)
            want = hf_llama_logits(path, ids, rope_scaling=rs)
            plain = hf_llama_logits(path, ids)
            got = O.OracleModel(path, n_ctx=128).eval(ids, logits_all=True)
            sig = float(want.std())
            err = np.abs(got - want)
            res["rope_" + arch.split("-")[1] + "_f16"] = {
                "n_tokens": int(len(ids)), "sigma_logits": sig, "max_err_sigma": float(err.max() / sig),
                "median_err_sigma": float(np.median(err) / sig), "rel_to_max": float(err.max() / np.abs(want).max()),
                "argmax_agree": int((got.argmax(-1) == want.argmax(-1)).sum()),
                "unscaled_max_err_sigma": float(np.abs(got - plain).max() / sig)}
        bpath = os.path.join(td, "bert_mini.gguf")
        mb.write_bert_gguf(bpath, mb.BERT_ARCHS["bert-mini"], seed=3)
        ids = np.concatenate([[2], rng.integers(5, 2048, 30), [3]])
        want = hf_bert_hidden(bpath, ids)
        got = B.BertRef(bpath).embed_tokens(ids, return_hidden=True)
        err = np.abs(got - want)
        res["bert_mini_f16"] = {"n_tokens": int(len(ids)), "sigma_hidden": float(want.std()), "max_err_sigma": float(err.max() / want.std()),
                                "median_err_sigma": float(np.median(err) / want.std())}
    return res


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=os.path.join(ROOT, "tests", "golden", "transformers_crosscheck.json"))
    a = ap.parse_args()
    r = measure()
    print(json.dumps(r, indent=1))
    with open(a.out, "w") as f:
        json.dump(r, f, indent=1)
