"""A small run of every GPU path for compute-sanitizer (memcheck / racecheck / synccheck): the persistent decode kernel,
one batched decode pass on the int8 kernel, one prompt pass on the tcgen05 path, the paged KV pool -- on the 3-layer test
model, a few positions each (the tools slow kernels down by orders of magnitude).
  compute-sanitizer --tool memcheck python tools/sanitize_target.py"""
import importlib, os, sys, tempfile
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
b2 = importlib.import_module("llms-on-kubernetes_b200")
synth = importlib.import_module("llms-on-kubernetes_b200.synth_gguf")

which = sys.argv[1] if len(sys.argv) > 1 else "all"
with tempfile.TemporaryDirectory() as d:
    path = os.path.join(d, "m.gguf")
    synth.write_model(path, "tiny-llama-q4_k_m", seed=3, norm_jitter=0.1)
    m = b2.Model(path)
    rng = np.random.default_rng(0)
    if which in ("all", "step"):
        ctx = b2.Context(m, 128)
        ids, _ = ctx.greedy(rng.integers(3, m.info.vocab, size=6).tolist(), 4)
        print("step kernel:", ids, "launches/step", ctx.launches_per_step)
        ctx.free()
    if which in ("all", "batch"):
        ctx = b2.Context(m, 128, n_slots=8)
        toks, _ = ctx.decode_batch(rng.integers(3, m.info.vocab, size=8).tolist(), [0] * 8, slots=list(range(8)))
        toks, _ = ctx.decode_batch(toks, [1] * 8, slots=list(range(8)))
        print("batched decode:", toks)
        ctx.free()
    if which in ("all", "prefill"):
        ctx = b2.Context(m, 128)
        nxt, _ = ctx.prefill(rng.integers(3, m.info.vocab, size=48).tolist())
        print("prompt pass:", nxt)
        ctx.free()
    if which in ("all", "paged"):
        os.environ["B2_KV_POOL_TOKENS"] = str(6 * 256)
        ctx = b2.Context(m, 512, n_slots=4)
        nxt, _ = ctx.prefill(rng.integers(3, m.info.vocab, size=40).tolist(), slot=2)
        toks, _ = ctx.decode_batch([nxt], [40], slots=[2])
        print("paged KV:", toks, ctx.kv_pages())
        ctx.free()
    m.free()
print("done")
