"""b2_tcq_gemm: the decode-batch mat-mul whose producer dequantises the raw GGUF blocks (tcq_gemm.cu).

Two checks against the oracle:
  * one-hot activations: row t of X is e_k, so Y[t][n] is the ONE fp16 value the kernel's dequantising warps wrote for
    W[n][k] -- compared bit for bit with the oracle's dequantised block rounded to fp16, for every position of a block
    and every row of a tile (layout of the swizzled tile, nibble / 2-bit planes, scale unpacking, the stream-K slices);
  * random activations: the fp16 reference product (oracle-dequantised W and oracle-quantised X, both rounded to fp16,
    summed in float64) within the fp32 summation-order bound, and the reference's integer mat-vec within the fp16 rounding.
"""
import numpy as np
import pytest

from oracle import quants_np as qn
from oracle import ref_model as rm

pytestmark = pytest.mark.gpu

TYPES = [qn.Q4_K, qn.Q6_K]


def f16_round(a: np.ndarray) -> np.ndarray:
    return np.clip(np.asarray(a, np.float32), -65504, 65504).astype(np.float16).astype(np.float32)


def blocks(synth, qt, N, K, seed):
    rng = np.random.default_rng(seed)
    packed = synth.random_blocks(qt, N * (K // 256), rng).reshape(N, -1)
    return packed, qn.dequantize(packed, qt).reshape(N, K).astype(np.float32)


@pytest.mark.parametrize("qt", TYPES)
@pytest.mark.parametrize("N,K", [(128, 256), (384, 1024), (200, 512), (4096, 4096)])
def test_one_hot_columns_return_the_dequantised_weights_exactly(b2, synth, qt, N, K, monkeypatch):
    monkeypatch.setenv("B2_TCQ_EXACT", "1")       # the exact conversion (the default converts in fp16 arithmetic, next test)
    if qt == qn.Q6_K:
        K = -(-K // 2048) * 2048          # rows of 210-byte blocks: a 16-byte row stride needs K % 2048 == 0
    packed, wd = blocks(synth, qt, N, K, seed=N + K + qt)
    want = f16_round(wd)
    rng = np.random.default_rng(5)
    # every position inside a 256-block at least once, blocks chosen at random
    for base in (0, 128):
        ks = np.array([256 * int(rng.integers(0, K // 256)) + base + i for i in range(128)])
        x = np.zeros((128, K), np.float32)
        x[np.arange(128), ks] = 1.0
        got = b2.tcq_gemm(qt, packed, x, N, K)
        assert np.array_equal(got, want[:, ks].T), (qt, N, K, base, float(np.abs(got - want[:, ks].T).max()))


@pytest.mark.parametrize("qt", TYPES)
@pytest.mark.parametrize("N,K", [(256, 512), (4096, 4096)])
def test_fast_conversion_stays_within_two_fp16_ulps(b2, synth, qt, N, K):
    """The default conversion rounds scale and minimum to fp16 before its fp16 multiply-add (and wants the activations' K
    order permuted, which b2_tcq_gemm does): every weight it produces is within 2 ulp of the exactly rounded fp16 value."""
    if qt == qn.Q6_K:
        K = -(-K // 2048) * 2048
    packed, wd = blocks(synth, qt, N, K, seed=N + K + qt + 1)
    want = f16_round(wd)
    rng = np.random.default_rng(6)
    for base in (0, 128):
        ks = np.array([256 * int(rng.integers(0, K // 256)) + base + i for i in range(128)])
        x = np.zeros((128, K), np.float32)
        x[np.arange(128), ks] = 1.0
        got = b2.tcq_gemm(qt, packed, x, N, K)
        w = want[:, ks].T
        # the error of a weight is a few fp16 roundings of its two terms (scale x code, minimum), which can be much larger
        # than the weight itself where they cancel: bound it by the largest magnitude of the weight's 256-block
        blk_max = np.abs(wd.reshape(N, K // 256, 256)).max(axis=2)            # [N][blocks]
        bound = blk_max[:, ks // 256].T                                        # [128][N]
        err = np.abs(got - w)
        assert (err <= 2.0 ** -9 * bound + 1e-12).all(), float((err / (bound + 1e-30)).max())
        assert (err <= 2.0 ** -11 * bound + 1e-12).mean() > 0.5               # and most weights are as good as the exact path


@pytest.mark.parametrize("qt", TYPES)
@pytest.mark.parametrize("N,K,T", [(128, 256, 2), (256, 512, 8), (384, 1024, 33), (1024, 2048, 64), (200, 256, 17), (4096, 4096, 32),
                                   (512, 14336, 16), (6144, 4096, 128)])
def test_tcq_gemm_matches_the_f16_reference(b2, synth, qt, N, K, T):
    if qt == qn.Q6_K:
        K = -(-K // 2048) * 2048
    packed, wd = blocks(synth, qt, N, K, seed=N + K + T + qt)
    rng = np.random.default_rng(T)
    x = rng.standard_normal((T, K)).astype(np.float32)
    x[:, ::97] *= 8.0
    got = b2.tcq_gemm(qt, packed, x, N, K)
    xq = np.empty_like(x)
    for t in range(T):
        d, q, _ = qn.quantize_q8_K(x[t])
        xq[t] = (d[:, None] * q.reshape(-1, 256).astype(np.float32)).reshape(-1)
    wb, xb = f16_round(wd).astype(np.float64), f16_round(xq).astype(np.float64)
    want = xb @ wb.T
    mag = np.abs(xb) @ np.abs(wb).T + 1e-30
    # (the default fp16 conversion adds up to ~2^-10 relative per weight on top of the fp32 summation order)
    assert np.abs(got - want).max() <= 1e-3 * mag.max(), float(np.abs(got - want).max() / mag.max())
    assert (np.abs(got - want) <= 1e-3 * mag + 1e-6).all()
    for t in range(0, T, max(1, T // 3)):
        ref = rm.gemv(qt, packed, x[t], N, K)
        assert np.abs(got[t] - ref).max() <= 1.5e-3 * mag[t].max()
    # the residual form adds into y
    y0 = rng.standard_normal((T, N)).astype(np.float32)
    got2 = b2.tcq_gemm(qt, packed, x, N, K, y0=y0)
    assert np.abs(got2 - (y0 + got)).max() <= 1e-5 * (np.abs(y0).max() + np.abs(got).max())


def test_shapes_outside_the_fused_path_are_refused(b2, synth):
    packed, _ = blocks(synth, qn.Q4_K, 128, 256, seed=1)
    x = np.zeros((129, 256), np.float32)
    with pytest.raises(b2.B2Error) as e:
        b2.tcq_gemm(qn.Q4_K, packed, x, 128, 256)
    assert e.value.code == -4
    q8 = synth.random_blocks(qn.Q8_0, 128 * 8, np.random.default_rng(0)).reshape(128, -1)
    with pytest.raises(b2.B2Error):
        b2.tcq_gemm(qn.Q8_0, q8, np.zeros((4, 256), np.float32), 128, 256)
    q6, _ = blocks(synth, qn.Q6_K, 128, 512, seed=2)           # 420-byte rows: no tensor map
    with pytest.raises(b2.B2Error) as e:
        b2.tcq_gemm(qn.Q6_K, q6, np.zeros((4, 512), np.float32), 128, 512)
    assert e.value.code == -4
