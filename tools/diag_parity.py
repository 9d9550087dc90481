"""Logit error of the decode path against the oracle on a full-shape synthetic model, position by position.
usage: python tools/diag_parity.py <spec> [n_positions]   (environment selects the kernel: B2_STEP_V=1, B2_NO_STEP=1)"""
import importlib, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
b2 = importlib.import_module("llms-on-kubernetes_b200")
synth = importlib.import_module("llms-on-kubernetes_b200.synth_gguf")
from oracle import parity as par, ref_model as rm

name = sys.argv[1]
n = int(sys.argv[2]) if len(sys.argv) > 2 else 12
path = os.path.join(os.environ.get("B2_BENCH_DIR", "/tmp"), f"b2_synth_{name}_s0.gguf")
if not os.path.exists(path):
    synth.write_model(path, name, seed=0)
rm.lib().ref_set_num_threads(len(os.sched_getaffinity(0)))
ref = rm.RefGGUFModel(path, n_ctx=n + 8)
m = b2.Model(path)
ctx = b2.Context(m, 64 if n < 56 else n + 8)
prompt = np.random.default_rng(5).integers(3, m.info.vocab, size=n).tolist()
for pos, t in enumerate(prompt):
    _, want = ref.decode(t, pos, want_logits=True)
    _, got = ctx.decode([t], pos, want_logits=True)
    st = par.logit_stats(want, got)
    hid_w = np.ctypeslib.as_array(rm.lib().ref_ctx_hidden(ref.ctx), shape=(m.info.d_model,)).copy()
    hid_g = ctx.hidden()[: m.info.d_model]
    herr = float(np.abs(hid_w - hid_g).max() / (np.abs(hid_w).max() or 1.0))
    print(f"pos {pos:3d}: logits err/range {st['err']:.5f}  kl {st['kl']:.2e}  oracle margin {st['margin']:.4f}  range {st['range']:.3f}  hidden err/max {herr:.2e}  launches/step {ctx.launches_per_step}")
