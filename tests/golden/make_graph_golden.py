"""Generates tests/golden/graph_golden.npz: fp32 logits of HF transformers' LlamaForCausalLM / Qwen2ForCausalLM
on the small synthetic GGUFs (weights dequantised by gguf-py, loaded through
transformers.modeling_gguf_pytorch_utils.load_gguf_checkpoint) for fixed prompts.  This is tier T2 of the oracle
(SURVEY 8c): it pins the *graph* (RMSNorm, RoPE pairing and Q/K permutation, GQA attention, SwiGLU, residuals,
tied embeddings, QKV bias, llama-3 rope factors) independently of anything written in this repo.
Run from the repo root:  python tests/golden/make_graph_golden.py      (needs torch + transformers, CPU only)"""
import hashlib
import importlib
import os
import sys
import tempfile

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sg = importlib.import_module("llms-on-kubernetes_b200.synth_gguf")

from transformers import LlamaConfig, LlamaForCausalLM, Qwen2Config, Qwen2ForCausalLM  # noqa: E402
from transformers.modeling_gguf_pytorch_utils import load_gguf_checkpoint  # noqa: E402

SEED = 7
MODELS = ["tiny-llama-q4_k_m", "tiny-llama3-q4_k_m", "tiny-llama-q5_k_m", "tiny-qwen2-q8_0"]
PROMPTS = [[1, 17, 300, 5, 9, 77, 402, 31], list(range(3, 67))]


def hf_model(path, spec):
    common = dict(vocab_size=spec.vocab, hidden_size=spec.d_model, intermediate_size=spec.d_ff,
                  num_hidden_layers=spec.n_layer, num_attention_heads=spec.n_head, num_key_value_heads=spec.n_kv,
                  rms_norm_eps=spec.eps, max_position_embeddings=spec.ctx, tie_word_embeddings=spec.tied,
                  attention_dropout=0.0)
    if spec.arch == "llama":
        rope = {"rope_type": "llama3", "factor": 8.0, "low_freq_factor": 1.0, "high_freq_factor": 4.0,
                "original_max_position_embeddings": 8192, "rope_theta": spec.rope_base} if spec.rope_freqs else \
               {"rope_type": "default", "rope_theta": spec.rope_base}
        cfg = LlamaConfig(head_dim=spec.head_dim, attention_bias=False, mlp_bias=False, rope_parameters=rope, **common)
        model = LlamaForCausalLM(cfg)
    else:
        cfg = Qwen2Config(rope_parameters={"rope_type": "default", "rope_theta": spec.rope_base}, **common)
        model = Qwen2ForCausalLM(cfg)
    model = model.to(torch.float32).eval()
    parsed = load_gguf_checkpoint(path, return_tensors=True, model_to_load=model)
    missing, unexpected = model.load_state_dict(parsed["tensors"], strict=False)
    missing = [m for m in missing if not (spec.tied and m == "lm_head.weight")]
    assert not missing and not unexpected, (missing, unexpected)
    if spec.tied:
        model.tie_weights()
    return model


out = {"generator": np.array(f"torch {torch.__version__}, transformers {importlib.import_module('transformers').__version__}, seed {SEED}")}
with tempfile.TemporaryDirectory() as d:
    for name in MODELS:
        spec = sg.SPECS[name]
        path = os.path.join(d, name + ".gguf")
        sg.write_model(path, name, seed=SEED, norm_jitter=0.1)
        out[f"{name}/file_sha256"] = np.array(hashlib.sha256(open(path, "rb").read()).hexdigest())
        model = hf_model(path, spec)
        for pi, prompt in enumerate(PROMPTS):
            ids = [t % spec.vocab for t in prompt]
            with torch.no_grad():
                logits = model(torch.tensor([ids])).logits[0, -1].to(torch.float32).numpy()
            out[f"{name}/prompt{pi}"] = np.array(ids, np.int32)
            out[f"{name}/logits{pi}"] = logits
            print(name, pi, logits[:4], float(np.abs(logits).max()))
np.savez_compressed(os.path.join(os.path.dirname(os.path.abspath(__file__)), "graph_golden.npz"), **out)
print("wrote graph_golden.npz")
