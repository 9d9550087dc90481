"""Run b2_tcq_gemm (and b2_tc_gemm for comparison) on the Llama-3.1-8B matrix shapes; meant to run under
`ncu --metrics gpu__time_duration.sum` so that the launch list gives the kernels' durations:
  python tools/tcq_shapes.py [T ...]"""
import sys
import importlib
import numpy as np

sys.path.insert(0, ".")
b2 = importlib.import_module("llms-on-kubernetes_b200")
Q4_K, Q6_K = 12, 14
BB = {Q4_K: 144, Q6_K: 210}
SHAPES = [("qkv", 6144, 4096, Q4_K), ("o", 4096, 4096, Q4_K), ("gate_up", 28672, 4096, Q4_K), ("down_q4", 4096, 14336, Q4_K),
          ("down_q6", 4096, 14336, Q6_K), ("lm_head/4", 32064, 4096, Q6_K)]
rng = np.random.default_rng(0)
for T in [int(a) for a in sys.argv[1:]] or [8, 32]:
    for name, N, K, qt in SHAPES:
        W = rng.integers(0, 256, size=(N, K // 256 * BB[qt]), dtype=np.uint8)
        W[:, (0 if qt == Q4_K else 208)::BB[qt]] = 0          # keep the fp16 scales finite: low byte 0 ...
        W[:, (1 if qt == Q4_K else 209)::BB[qt]] = 0x2c        # ... high byte 0x2c -> 0.0625
        if qt == Q4_K:
            W[:, 2::144] = 0; W[:, 3::144] = 0x2c
        x = rng.standard_normal((T, K)).astype(np.float32)
        y = b2.tcq_gemm(qt, W, x, N, K)
        print(name, T, N, K, float(np.abs(y).max()), flush=True)
