"""Diagnostic: error of the prompt path (and of the decode steps after it) against the oracle, per tiny model."""
import importlib, os, sys, tempfile
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
b2 = importlib.import_module("llms-on-kubernetes_b200")
sg = importlib.import_module("llms-on-kubernetes_b200.synth_gguf")
from oracle import ref_model as rm

def kl(p, q):
    p = p.astype(np.float64); q = q.astype(np.float64); p -= p.max(); q -= q.max()
    lp = p - np.log(np.exp(p).sum()); lq = q - np.log(np.exp(q).sum())
    return float((np.exp(lp) * (lp - lq)).sum())

d = tempfile.mkdtemp()
for name in ["tiny-llama-q4_k_m", "tiny-llama3-q4_k_m", "tiny-llama-q5_k_m", "tiny-qwen2-q8_0"]:
    path = os.path.join(d, name + ".gguf")
    sg.write_model(path, name, seed=1)
    for n in (48, 65, 200, 333):
        ref = rm.RefGGUFModel(path, n_ctx=512); m = b2.Model(path); ctx = b2.Context(m, 512)
        toks = np.random.default_rng(n).integers(3, m.info.vocab, size=n).tolist()
        for p, t in enumerate(toks):
            tok, lr = ref.decode(t, p, want_logits=p == n - 1)
        gt, lg = ctx.prefill(toks)
        out = [f"{name} n={n}: prefill err {np.abs(lg-lr).max()/np.abs(lr).max():.4f} kl {kl(lr,lg):.2e}"]
        # int8 path for the same prompt (32-column chunks) for comparison
        os.environ["B2_NO_TC_PREFILL"] = "1"
        ctx2 = b2.Context(m, 512); _, l8 = ctx2.decode(toks, 0); ctx2.free()
        del os.environ["B2_NO_TC_PREFILL"]
        out.append(f"int8 path err {np.abs(l8-lr).max()/np.abs(lr).max():.4f} kl {kl(lr,l8):.2e}")
        nxt = tok; worst = 0; wk = 0
        for i in range(8):
            tr, lr = ref.decode(nxt, n + i); tg, lg = ctx.decode([nxt], n + i)
            worst = max(worst, np.abs(lg-lr).max()/np.abs(lr).max()); wk = max(wk, kl(lr, lg)); nxt = tr
        out.append(f"decode after: err {worst:.4f} kl {wk:.2e}")
        print(" | ".join(out), flush=True)
        ctx.free(); m.free(); ref.close()
