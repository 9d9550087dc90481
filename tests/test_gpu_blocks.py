"""CUDA dequantise + mat-vec through the C ABI against the oracle.  Bit-exact for dequantised weights;
the mat-vec is integer inside a block and fp32 across blocks, so it is compared with a stated tolerance
on the summation order only."""
import hashlib
import os

import numpy as np
import pytest

from oracle import quants_np as qn
from oracle import ref_model as rm

pytestmark = pytest.mark.gpu
TYPES = [qn.Q8_0, qn.Q4_K, qn.Q5_K, qn.Q6_K]
GOLDEN = os.path.join(os.path.dirname(__file__), "golden")


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32).reshape(-1)


@pytest.mark.parametrize("qt", TYPES)
def test_dequant_bit_exact_edge_cases(b2, synth, qt):
    rng = np.random.default_rng(200 + qt)
    blocks = synth.random_blocks(qt, 4096, rng, edge_cases=True)
    got = b2.dequantize(blocks, qt)
    assert np.array_equal(bits(got), bits(qn.dequantize(blocks, qt)))


@pytest.mark.parametrize("qt", TYPES)
def test_dequant_golden_fixture(b2, synth, qt):
    g = np.load(os.path.join(GOLDEN, "blocks_golden.npz"))
    name = synth.TYPE_NAME[qt]
    got = b2.dequantize(g[f"{name}_blocks"], qt)
    assert hashlib.sha256(got.tobytes()).hexdigest() == str(g[f"{name}_out_sha256"])


@pytest.mark.parametrize("qt", TYPES)
def test_dequant_uniform_random_bytes(b2, qt):
    rng = np.random.default_rng(qt + 300)
    bb = qn.BLOCK[qt][1]
    raw = rng.integers(0, 256, size=(8192, bb), dtype=np.uint8)
    for off in ({qn.Q8_0: [0], qn.Q4_K: [0, 2], qn.Q5_K: [0, 2], qn.Q6_K: [208]}[qt]):
        h = raw[:, off:off + 2].copy().view(np.uint16)
        h[(h & 0x7C00) == 0x7C00] &= 0x3FFF
        raw[:, off:off + 2] = h.view(np.uint8)
    assert np.array_equal(bits(b2.dequantize(raw, qt)), bits(qn.dequantize(raw, qt)))


@pytest.mark.parametrize("qt", TYPES)
def test_dequant_million_blocks_vs_c_oracle(b2, synth, qt):
    """>= 1e6 blocks (SURVEY section 7 minimum slice), compared through a digest of the full output."""
    rng = np.random.default_rng(qt + 400)
    n = 1 << 20
    blocks = synth.random_blocks(qt, n, rng, edge_cases=True)
    got = b2.dequantize(blocks, qt)
    want = rm.dequantize(blocks, qt, got.size)
    assert hashlib.sha256(got.tobytes()).digest() == hashlib.sha256(want.tobytes()).digest()


def test_dequant_f16_bf16_f32(b2):
    rng = np.random.default_rng(9)
    x = rng.standard_normal(4096).astype(np.float32)
    assert np.array_equal(b2.dequantize(x.view(np.uint8), b2.F32), x)
    h = x.astype(np.float16)
    assert np.array_equal(b2.dequantize(h.view(np.uint8), b2.F16), h.astype(np.float32))
    bf = (x.view(np.uint32) >> 16).astype(np.uint16)
    assert np.array_equal(bits(b2.dequantize(bf.view(np.uint8), b2.BF16)), bf.astype(np.uint32) << 16)


def test_dequant_empty_and_ragged(b2):
    assert b2.dequantize(np.zeros(0, np.uint8), b2.Q4_K).size == 0
    with pytest.raises(b2.B2Error):
        b2.dequantize(np.zeros(143, np.uint8), b2.Q4_K)
    with pytest.raises(b2.B2Error):
        b2.dequantize(np.zeros(144, np.uint8), 99)


def _gemv_case(b2, synth, qt, N, K, seed, scale=1.0):
    rng = np.random.default_rng(seed)
    be = qn.BLOCK[qt][0]
    W = synth.random_blocks(qt, N * (K // be), rng).reshape(N, -1)
    x = (rng.standard_normal(K) * scale).astype(np.float32)
    got = b2.gemv(qt, W, x, N, K)
    want = rm.gemv(qt, W, x, N, K)
    # identical integer block sums; only the fp32 accumulation order differs (8-lane partial sums on the
    # CPU, 32-lane tree on the GPU): a few ulp of the largest partial
    tol = 4e-6 * np.abs(qn.dequantize(W, qt)).astype(np.float64).dot(np.abs(x).astype(np.float64)).max()
    assert np.abs(got - want).max() <= tol, (np.abs(got - want).max(), tol)
    return got, want


@pytest.mark.parametrize("qt", TYPES)
@pytest.mark.parametrize("N,K", [(2, 256), (64, 512), (7, 1024), (33, 2048), (128, 4096), (10, 6144), (16, 14336)])
def test_gemv_matches_oracle(b2, synth, qt, N, K):
    _gemv_case(b2, synth, qt, N, K, seed=N * 131 + K + qt)


@pytest.mark.parametrize("qt", TYPES)
def test_gemv_more_rows_than_one_wave(b2, synth, qt):
    _gemv_case(b2, synth, qt, 6144, 1024, seed=qt)


@pytest.mark.parametrize("qt", TYPES)
def test_gemv_zero_and_tiny_activations(b2, synth, qt):
    rng = np.random.default_rng(1)
    N, K = 8, 1024
    W = synth.random_blocks(qt, N * (K // qn.BLOCK[qt][0]), rng).reshape(N, -1)
    assert not b2.gemv(qt, W, np.zeros(K, np.float32), N, K).any()
    _gemv_case(b2, synth, qt, N, K, seed=3, scale=1e-20)
    x = np.zeros(K, np.float32); x[5] = -2.5; x[700] = 2.5       # tie between -max and +max in a block
    assert np.allclose(b2.gemv(qt, W, x, N, K), rm.gemv(qt, W, x, N, K), rtol=1e-5, atol=1e-7)


def test_gemv_linearity_property(b2, synth):
    """Size-independent property at a full-size shape: W.(2x) == 2 (W.x) exactly (power-of-two scaling commutes
    with the activation quantiser), for the lm_head shape of Llama-3.1-8B rows sampled down."""
    rng = np.random.default_rng(2)
    N, K = 4096, 4096
    W = synth.random_blocks(qn.Q6_K, N * (K // 256), rng).reshape(N, -1)
    x = rng.standard_normal(K).astype(np.float32)
    y1, y2 = b2.gemv(qn.Q6_K, W, x, N, K), b2.gemv(qn.Q6_K, W, 2 * x, N, K)
    assert np.array_equal(2 * y1, y2)


def test_gemv_batch_columns(b2, synth):
    rng = np.random.default_rng(4)
    N, K, B = 32, 1024, 3
    W = synth.random_blocks(qn.Q4_K, N * (K // 256), rng).reshape(N, -1)
    X = rng.standard_normal((B, K)).astype(np.float32)
    Y = b2.gemv(qn.Q4_K, W, X, N, K)
    for b in range(B):
        assert np.array_equal(Y[b], b2.gemv(qn.Q4_K, W, X[b], N, K))


def test_gemv_rejects_bad_shapes(b2):
    d = b2.DevBuf(4096)
    for args in [(b2.Q4_K, 8, 100), (b2.F16, 8, 256), (b2.Q8_0, 8, 48), (b2.Q4_K, 0, 256)]:
        assert b2.lib.b2_gemv(args[0], d.ptr, d.ptr, d.ptr, args[1], args[2], 1, None) < 0
        assert b2.lib.b2_last_error()
