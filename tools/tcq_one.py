"""One b2_tcq_gemm call on the gate|up shape (for B2_TCQ_TRACE=1 / B2_TCQ_DBG experiments)."""
import sys, importlib
import numpy as np
sys.path.insert(0, ".")
b2 = importlib.import_module("llms-on-kubernetes_b200")
N, K, T = 28672, 4096, int(sys.argv[1]) if len(sys.argv) > 1 else 32
rng = np.random.default_rng(0)
W = rng.integers(0, 256, size=(N, K // 256 * 144), dtype=np.uint8)
W[:, 0::144] = 0; W[:, 1::144] = 0x2c; W[:, 2::144] = 0; W[:, 3::144] = 0x2c
x = rng.standard_normal((T, K)).astype(np.float32)
for _ in range(2):
    y = b2.tcq_gemm(12, W, x, N, K)
print(float(np.abs(y).max()))
