import importlib, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
b2 = importlib.import_module("llms-on-kubernetes_b200")
sg = importlib.import_module("llms-on-kubernetes_b200.synth_gguf")
qt = {"Q8_0": sg.Q8_0, "Q4_K": sg.Q4_K}[sys.argv[1]]
N, K = int(sys.argv[2]), int(sys.argv[3])
rng = np.random.default_rng(0)
W = sg.random_blocks(qt, N * (K // sg.BLOCK[qt][0]), rng).reshape(N, -1)
x = rng.standard_normal(K).astype(np.float32)
print(b2.gemv(qt, W, x, N, K)[:4])
