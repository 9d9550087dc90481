"""Run one b2_gemv shape a few times (for ncu).  usage: prof_one.py TYPE N K [reps]"""
import importlib, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
b2 = importlib.import_module("llms-on-kubernetes_b200")
sg = importlib.import_module("llms-on-kubernetes_b200.synth_gguf")
qt = {"Q4_K": sg.Q4_K, "Q5_K": sg.Q5_K, "Q6_K": sg.Q6_K, "Q8_0": sg.Q8_0}[sys.argv[1]]
N, K = int(sys.argv[2]), int(sys.argv[3])
reps = int(sys.argv[4]) if len(sys.argv) > 4 else 3
rng = np.random.default_rng(0)
be, bb = sg.BLOCK[qt]
W = sg.random_blocks(qt, N * (K // be), rng).reshape(-1)
dW = b2.DevBuf.from_numpy(W)
dx, dy = b2.DevBuf.from_numpy(rng.standard_normal(K).astype(np.float32)), b2.DevBuf(4 * N)
for _ in range(reps):
    b2.check(b2.lib.b2_gemv(qt, dW.ptr, dx.ptr, dy.ptr, N, K, 1, None))
b2.check(b2.lib.b2_stream_sync(None))
print("ok")
