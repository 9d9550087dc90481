"""Stand-alone timing of b2_tc_gemm (tcgen05 prompt-phase mat-mul) on the Llama-3.1-8B layer shapes.
Usage: python tools/bench_tc_gemm.py [reps]   -- wall clock over back-to-back calls; each call dequantises the matrix
to bf16, rounds the activations to bf16 and runs the tensor-core kernel (run under `ncu --metrics
gpu__time_duration.sum` for the split between the three kernels)."""
import importlib, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
b2 = importlib.import_module("llms-on-kubernetes_b200")
sg = importlib.import_module("llms-on-kubernetes_b200.synth_gguf")
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 10
shapes = [("qkv", 6144, 4096), ("gate_up", 28672, 4096), ("down", 4096, 14336)]
rng = np.random.default_rng(0)
qt = sg.Q4_K
for sname, N, K in shapes:
    be, bb = sg.BLOCK[qt]
    W = sg.random_blocks(qt, N * (K // be), rng).reshape(-1)
    dW = b2.DevBuf.from_numpy(W)
    for T in (128, 256, 512, 1024, 2048):
        X = rng.standard_normal((T, K)).astype(np.float32)
        dx, dy = b2.DevBuf.from_numpy(X), b2.DevBuf(4 * N * T)
        sc = b2.DevBuf(b2.lib.b2_tc_gemm_scratch_bytes(N, K, T))
        for i in range(2):
            b2.check(b2.lib.b2_tc_gemm(qt, dW.ptr, dx.ptr, dy.ptr, N, K, T, 0, sc.ptr, None))
        b2.check(b2.lib.b2_stream_sync(None))
        t0 = time.perf_counter()
        for i in range(reps):
            b2.check(b2.lib.b2_tc_gemm(qt, dW.ptr, dx.ptr, dy.ptr, N, K, T, 0, sc.ptr, None))
        b2.check(b2.lib.b2_stream_sync(None))
        dt = (time.perf_counter() - t0) / reps
        print(f"Q4_K {sname:8s} N={N:6d} K={K:5d} T={T:4d} {dt*1e6:9.1f} us  {2.0*N*K*T/dt/1e12:7.1f} TFLOP/s (whole call)", flush=True)
        for d in (dx, dy, sc): d.free()
    dW.free()
