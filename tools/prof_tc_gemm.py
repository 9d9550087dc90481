"""One shape of b2_tc_gemm a few times, for `ncu -k regex:tc_gemm_kernel`: the Llama-3.1-8B gate|up matrix (28672 x 4096,
Q4_K) against 2048 prompt rows.  Usage: python tools/prof_tc_gemm.py [T]"""
import importlib, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
b2 = importlib.import_module("llms-on-kubernetes_b200")
sg = importlib.import_module("llms-on-kubernetes_b200.synth_gguf")
T = int(sys.argv[1]) if len(sys.argv) > 1 else 2048
N, K, qt = 28672, 4096, sg.Q4_K
rng = np.random.default_rng(0)
W = sg.random_blocks(qt, N * (K // sg.BLOCK[qt][0]), rng).reshape(-1)
X = rng.standard_normal((T, K)).astype(np.float32)
dW, dx, dy = b2.DevBuf.from_numpy(W), b2.DevBuf.from_numpy(X), b2.DevBuf(4 * N * T)
sc = b2.DevBuf(b2.lib.b2_tc_gemm_scratch_bytes(N, K, T))
for i in range(4):
    b2.check(b2.lib.b2_tc_gemm(qt, dW.ptr, dx.ptr, dy.ptr, N, K, T, 0, sc.ptr, None))
b2.check(b2.lib.b2_stream_sync(None))
print("ok", 2.0 * N * K * T / 1e12, "TFLOP per call")
