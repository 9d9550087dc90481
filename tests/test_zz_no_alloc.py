"""b2_attn_decode_static -- decode attention for callers that are not allowed to allocate GPU memory.

Who needs it: a host that calls attention from inside a CUDA-graph capture or a tight serving loop (where
cudaMalloc would stall the whole device), a plugin that was given a fixed memory budget, or an embedding
application whose own allocator has claimed all of HBM.  The entry point keeps its split-merge scratch in a
fixed array compiled into libb2.so instead of taking a scratch pointer, so it has hard size limits.

This file is named test_zz_* so pytest collects it last: a run that stops at the first failure has then
already executed every other test.  The test calls the entry point through the C ABI twice in a row (the
array keeps the previous call's contents) and compares each result with b2_attn_decode on caller-provided
scratch -- same kernel, so the outputs must be identical bit for bit; it also checks that requests beyond
the built-in limits are refused with a clear error instead of overrunning the array.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _run_pair(b2, rng, n_head, n_kv, hd, n_ctx, T):
    q = rng.standard_normal((n_head, hd)).astype(np.float32)
    K = (rng.standard_normal((n_kv, n_ctx, hd)) * 0.5).astype(np.float16)
    V = rng.standard_normal((n_kv, n_ctx, hd)).astype(np.float16)
    dq, dk, dv = b2.DevBuf.from_numpy(q), b2.DevBuf.from_numpy(K), b2.DevBuf.from_numpy(V)
    d_ref, d_new = b2.DevBuf(q.nbytes, zero=True), b2.DevBuf(q.nbytes, zero=True)
    scratch = b2.DevBuf(b2.lib.b2_attn_scratch_bytes(n_head, hd, T), zero=True)
    b2.check(b2.lib.b2_attn_decode(dq.ptr, dk.ptr, dv.ptr, d_ref.ptr, n_head, n_kv, hd, n_ctx, T, scratch.ptr, None))
    b2.check(b2.lib.b2_attn_decode_static(dq.ptr, dk.ptr, dv.ptr, d_new.ptr, n_head, n_kv, hd, n_ctx, T, None))
    b2.check(b2.lib.b2_stream_sync(None))
    want, got = d_ref.to_numpy(np.float32), d_new.to_numpy(np.float32)
    assert np.isfinite(want).all() and np.abs(want).max() > 0
    # identical kernel and inputs: the tolerance of the attention tests (2e-4 of the output range) is met with 0
    assert np.abs(got - want).max() <= 2e-4 * max(1.0, np.abs(want).max())
    assert np.array_equal(got, want)


def test_static_scratch_matches_caller_scratch_twice_in_a_row(b2):
    rng = np.random.default_rng(2026)
    # two consecutive calls on the shared array: first a multi-split case (uses the partial records and the
    # arrival counters), then a different shape that reinterprets the same bytes
    _run_pair(b2, rng, 32, 8, 128, 4096, 3000)
    _run_pair(b2, rng, 32, 4, 64, 2048, 2048)
    # and once more with a single split (no scratch traffic at all)
    _run_pair(b2, rng, 12, 2, 128, 512, 77)


def test_static_scratch_refuses_oversize_requests(b2):
    d = b2.DevBuf(1 << 16)
    for n_head, n_kv, hd, n_ctx, T in [(32, 8, 128, 8192, 4097), (128, 8, 128, 1024, 16), (32, 8, 256, 1024, 16)]:
        rc = b2.lib.b2_attn_decode_static(d.ptr, d.ptr, d.ptr, d.ptr, n_head, n_kv, hd, n_ctx, T, None)
        assert rc == -6, (n_head, hd, T, rc)
        assert b"built-in scratch" in b2.lib.b2_last_error()


def test_static_scratch_refuses_zero_and_negative_sizes(b2):
    """Sizes below the valid range must be turned away before the built-in array is addressed at all: with a negative
    head count the offset of the arrival counters would wrap around and the clearing memset would land in front of
    the array, in some other device global of the library."""
    d = b2.DevBuf(1 << 16)
    for n_head, n_kv, hd, n_ctx, T in [(-1, 1, 128, 1024, 16), (0, 1, 128, 1024, 16), (32, 0, 128, 1024, 16), (32, -8, 128, 1024, 16),
                                        (32, 5, 128, 1024, 16), (32, 8, -128, 1024, 16), (32, 8, 0, 1024, 16), (32, 8, 96, 1024, 16),
                                        (32, 8, 128, 1024, 0), (32, 8, 128, 1024, -3), (32, 8, 128, 0, 16), (32, 8, 128, 64, 65)]:
        rc = b2.lib.b2_attn_decode_static(d.ptr, d.ptr, d.ptr, d.ptr, n_head, n_kv, hd, n_ctx, T, None)
        assert rc == -1, (n_head, n_kv, hd, n_ctx, T, rc)
        assert b"b2_attn_decode_static" in b2.lib.b2_last_error()
    b2.check(b2.lib.b2_stream_sync(None))

