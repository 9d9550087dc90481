"""tiny-qwen2-q8_0, 33-token prompt: logits of the last position through the default prompt path (tcgen05 pass from 24
tokens) and through the per-token kernels (B2_NO_TC_PREFILL=1), each against the oracle.  Run once per setting:
  python tools/diag_prompt_path.py            B2_NO_TC_PREFILL=1 python tools/diag_prompt_path.py"""
import importlib, os, sys, tempfile
import numpy as np
sys.path.insert(0, ".")
b2 = importlib.import_module("llms-on-kubernetes_b200")
synth = importlib.import_module("llms-on-kubernetes_b200.synth_gguf")
from oracle import ref_model as rm
path = os.path.join(tempfile.mkdtemp(), "m.gguf")
synth.write_model(path, "tiny-qwen2-q8_0", seed=7, norm_jitter=0.1)
m = b2.Model(path); ctx = b2.Context(m, 256)
ref = rm.RefGGUFModel(path, n_ctx=256)
prompt = np.random.default_rng(1).integers(3, m.info.vocab, size=33).tolist()
tok, lg = ctx.decode(prompt, 0)
for p, t in enumerate(prompt):
    rt, rl = ref.decode(t, p, want_logits=p == len(prompt) - 1)
print("B2_NO_TC_PREFILL =", os.environ.get("B2_NO_TC_PREFILL"), "max |logit diff| / range =", float(np.abs(lg - rl).max() / (rl.max() - rl.min())))
