"""Diagnostic: b2_ctx_decode's 32-column chunked prompt path past 256 positions against the oracle."""
import importlib, os, sys, tempfile
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
os.environ["B2_NO_TC_PREFILL"] = "1"
b2 = importlib.import_module("llms-on-kubernetes_b200")
sg = importlib.import_module("llms-on-kubernetes_b200.synth_gguf")
from oracle import ref_model as rm
d = tempfile.mkdtemp()
name = "tiny-llama-q4_k_m"
path = os.path.join(d, name + ".gguf")
sg.write_model(path, name, seed=1)
m = b2.Model(path)
toks = np.random.default_rng(1).integers(3, m.info.vocab, size=400).tolist()
ref = rm.RefGGUFModel(path, n_ctx=512)
want = {}
for p, t in enumerate(toks):
    _, lr = ref.decode(t, p, want_logits=True)
    want[p + 1] = lr
for graph in ("0", "1"):
    os.environ["B2_NO_GRAPH"] = graph
    for n in (200, 250, 256, 257, 260, 288, 289, 333, 400):
        ctx = b2.Context(m, 512)
        _, lg = ctx.decode(toks[:n], 0)
        lr = want[n]
        print(f"no_graph={graph} n={n}: logits err {np.abs(lg-lr).max()/np.abs(lr).max():.4f}", flush=True)
        ctx.free()
