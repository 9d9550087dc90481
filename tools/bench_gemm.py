"""Stand-alone timing of b2_gemm (batched tensor-core mat-mul) on the Llama-3.1-8B layer shapes.
Usage: python tools/bench_gemm.py [reps]   -- prints weight GB/s per shape and column count (wall clock over many
back-to-back launches on distinct weight copies so the working set exceeds L2; includes the activation quantiser)."""
import importlib, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
b2 = importlib.import_module("llms-on-kubernetes_b200")
sg = importlib.import_module("llms-on-kubernetes_b200.synth_gguf")
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 20
shapes = [("qkv", 6144, 4096), ("gate_up", 28672, 4096), ("down", 4096, 14336), ("lm_head", 128256, 4096)]
rng = np.random.default_rng(0)
for qt, name in ((sg.Q4_K, "Q4_K"), (sg.Q6_K, "Q6_K")):
    for sname, N, K in shapes:
        be, bb = sg.BLOCK[qt]
        row = K // be * bb
        n_copies = max(2, min(16, int(400e6 // (N * row)) + 1))     # > 126 MB L2 in total
        W = sg.random_blocks(qt, N * (K // be), rng).reshape(-1)
        dW = [b2.DevBuf.from_numpy(W) for _ in range(n_copies)]
        for B in (1, 8, 16, 32):
            X = rng.standard_normal((B, K)).astype(np.float32)
            dx, dy = b2.DevBuf.from_numpy(X), b2.DevBuf(4 * N * B)
            sc = b2.DevBuf(b2.lib.b2_gemm_scratch_bytes(K, B))
            for i in range(3):
                b2.check(b2.lib.b2_gemm(qt, dW[i % n_copies].ptr, dx.ptr, dy.ptr, N, K, B, sc.ptr, None))
            b2.check(b2.lib.b2_stream_sync(None))
            t0 = time.perf_counter()
            for i in range(reps):
                b2.check(b2.lib.b2_gemm(qt, dW[i % n_copies].ptr, dx.ptr, dy.ptr, N, K, B, sc.ptr, None))
            b2.check(b2.lib.b2_stream_sync(None))
            dt = (time.perf_counter() - t0) / reps
            print(f"{name} {sname:8s} N={N:6d} K={K:5d} B={B:2d} {N*row/1e6:8.1f} MB  {dt*1e6:8.1f} us  {N*row/dt/1e9:7.1f} GB/s "
                  f"{2.0*N*K*B/dt/1e12:6.2f} TOP/s", flush=True)
            for d in (dx, dy, sc): d.free()
        for d in dW: d.free()
