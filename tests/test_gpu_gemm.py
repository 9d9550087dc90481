"""Batched mat-mul on the tensor cores (b2_gemm) against the oracle's mat-vec, column by column.
The integer sub-block sums are the same as the mat-vec's (and ggml's); only the fp32 accumulation order
across super-blocks differs, so the tolerance is a few ulp of the largest partial sum."""
import numpy as np
import pytest

from oracle import quants_np as qn
from oracle import ref_model as rm

pytestmark = pytest.mark.gpu
TYPES = [qn.Q4_K, qn.Q5_K, qn.Q6_K, qn.Q8_0]


def _case(b2, synth, qt, N, K, B, seed, scale=1.0):
    rng = np.random.default_rng(seed)
    W = synth.random_blocks(qt, N * (K // qn.BLOCK[qt][0]), rng).reshape(N, -1)
    X = (rng.standard_normal((B, K)) * scale).astype(np.float32)
    got = b2.gemm(qt, W, X, N, K)
    assert got.shape == (B, N)
    absW = np.abs(qn.dequantize(W, qt)).astype(np.float64)
    for b in range(B):
        want = rm.gemv(qt, W, X[b], N, K)
        tol = 4e-6 * absW.dot(np.abs(X[b]).astype(np.float64)).max()
        err = np.abs(got[b] - want).max()
        assert err <= tol, (b, err, tol)
    return got


@pytest.mark.parametrize("qt", TYPES)
@pytest.mark.parametrize("N,K,B", [(16, 256, 8), (2, 256, 1), (64, 2048, 8), (33, 4096, 5), (128, 4096, 16), (48, 6144, 12),
                                   (32, 14336, 8), (40, 4096, 32), (18, 2304, 27)])
def test_gemm_matches_oracle_per_column(b2, synth, qt, N, K, B):
    _case(b2, synth, qt, N, K, B, seed=N * 7 + K + B + qt)


@pytest.mark.parametrize("qt", TYPES)
def test_gemm_more_tiles_than_ctas(b2, synth, qt):
    _case(b2, synth, qt, 8192, 1024, 8, seed=qt)


@pytest.mark.parametrize("qt", TYPES)
def test_gemm_agrees_with_gemv_kernel(b2, synth, qt):
    """Same integer arithmetic as the single-column kernel: results agree to fp32 summation order."""
    rng = np.random.default_rng(11 + qt)
    N, K, B = 256, 4096, 8
    W = synth.random_blocks(qt, N * (K // qn.BLOCK[qt][0]), rng).reshape(N, -1)
    X = rng.standard_normal((B, K)).astype(np.float32)
    Y = b2.gemm(qt, W, X, N, K)
    for b in range(B):
        y = b2.gemv(qt, W, X[b], N, K)
        assert np.allclose(Y[b], y, rtol=0, atol=4e-6 * np.abs(y).max() + 1e-6)


@pytest.mark.parametrize("N,K,B", [(24, 1536, 8), (16, 8960, 20), (10, 96, 3)])
def test_gemm_q8_0_ragged_k(b2, synth, N, K, B):
    """Q8_0 rows need K to be a multiple of 32 only (Qwen2.5: 1536, 8960): partial last chunk, partial warp."""
    _case(b2, synth, qn.Q8_0, N, K, B, seed=N + K + B)


def test_gemm_zero_columns_and_tiny_values(b2, synth):
    rng = np.random.default_rng(5)
    N, K = 32, 1024
    W = synth.random_blocks(qn.Q4_K, N * (K // 256), rng).reshape(N, -1)
    X = np.zeros((8, K), np.float32)
    X[3] = rng.standard_normal(K).astype(np.float32) * 1e-20
    Y = b2.gemm(qn.Q4_K, W, X, N, K)
    assert not Y[[0, 1, 2, 4, 5, 6, 7]].any()
    assert np.allclose(Y[3], rm.gemv(qn.Q4_K, W, X[3], N, K), rtol=1e-5, atol=0)


def test_gemm_rejects_unsupported(b2):
    d = b2.DevBuf(1 << 16)
    assert b2.lib.b2_gemm(b2.F16, d.ptr, d.ptr, d.ptr, 16, 256, 8, d.ptr, None) < 0 and b2.lib.b2_gemm(99, d.ptr, d.ptr, d.ptr, 16, 256, 8, d.ptr, None) < 0
    assert b2.lib.b2_gemm(b2.Q4_K, d.ptr, d.ptr, d.ptr, 16, 256, 33, d.ptr, None) < 0
    assert b2.lib.b2_gemm(b2.Q4_K, d.ptr, d.ptr, d.ptr, 16, 100, 8, d.ptr, None) < 0
    assert b2.lib.b2_gemm_scratch_bytes(4096, 8) == max(2 * (8 * 2080 + 8 * 8 * 20), 4 * (8 * 1056 + 32 * 8 * 4))
