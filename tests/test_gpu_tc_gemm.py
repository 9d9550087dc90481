"""b2_tc_gemm (tcgen05 prompt-phase mat-mul) against the oracle's dequantised weights and activation quantisers.

The kernel multiplies fp16 copies of the dequantised matrix and of the activations with fp32 accumulation; the
activations first go through the reference's own quantisation for that weight type (Q8_K blocks for K-quants, Q8_0
blocks for Q8_0 weights) and d * q is rounded to fp16.  So the exact reference is: oracle-dequantised W rounded to
fp16, oracle-quantised X (d * q) rounded to fp16, products summed in float64; against that only the fp32 summation
order differs (2e-5 of sum |w||x|).  Against the reference's integer mat-vec (the oracle's gemv, what llama.cpp
computes) the fp16 roundings of W and d * q show: 2^-11 relative each, a few 1e-4 of sum |w||x| at most."""
import numpy as np
import pytest

from oracle import quants_np as qn
from oracle import ref_model as rm

pytestmark = pytest.mark.gpu

TYPES = [qn.Q4_K, qn.Q5_K, qn.Q6_K, qn.Q8_0]


def f16_round(a: np.ndarray) -> np.ndarray:
    return np.clip(np.asarray(a, np.float32), -65504, 65504).astype(np.float16).astype(np.float32)


def quantised_acts(qt, x: np.ndarray) -> np.ndarray:
    """d * q of every row, as the reference's mat-mul with weights of type qt would quantise it."""
    out = np.empty_like(x)
    for t in range(x.shape[0]):
        if qt == qn.Q8_0:
            blk = qn.quantize_q8_0(x[t]).reshape(-1, 34)
            d = blk[:, :2].copy().view(np.float16).astype(np.float32)
            out[t] = (d * blk[:, 2:].view(np.int8).astype(np.float32)).reshape(-1)
        else:
            d, q, _ = qn.quantize_q8_K(x[t])
            out[t] = (d[:, None] * q.reshape(-1, 256).astype(np.float32)).reshape(-1)
    return out


def make(synth, qt, N, K, T, seed):
    rng = np.random.default_rng(seed)
    packed = synth.random_blocks(qt, N * (K // qn.BLOCK[qt][0]), rng).reshape(N, -1)
    wd = qn.dequantize(packed, qt).reshape(N, K).astype(np.float32)
    x = rng.standard_normal((T, K)).astype(np.float32)
    x[:, ::97] *= 8.0
    return packed, wd, x


@pytest.mark.parametrize("qt", TYPES)
@pytest.mark.parametrize("N,K,T", [(128, 256, 128), (256, 512, 40), (384, 1024, 300), (1024, 2048, 512), (200, 256, 1), (4096, 4096, 257)])
def test_tc_gemm_matches_f16_reference(b2, synth, qt, N, K, T):
    packed, wd, x = make(synth, qt, N, K, T, seed=N + K + T + qt)
    got = b2.tc_gemm(qt, packed, x, N, K)
    xq = quantised_acts(qt, x)
    wb, xb = f16_round(wd).astype(np.float64), f16_round(xq).astype(np.float64)
    want = xb @ wb.T
    mag = np.abs(xb) @ np.abs(wb).T + 1e-30
    assert np.abs(got - want).max() <= 2e-5 * mag.max(), float(np.abs(got - want).max() / mag.max())
    assert (np.abs(got - want) <= 2e-5 * mag + 1e-6).all()
    # against the reference's own arithmetic for this row (integer sub-block dots with fp32 scales)
    for t in range(0, T, max(1, T // 4)):
        ref = rm.gemv(qt, packed, x[t], N, K)
        assert (np.abs(got[t] - ref) <= 5e-4 * mag[t]).all(), float((np.abs(got[t] - ref) / mag[t]).max())


def test_tc_gemm_accumulates_into_the_output(b2, synth):
    qt, N, K, T = qn.Q4_K, 512, 512, 200
    packed, wd, x = make(synth, qt, N, K, T, seed=3)
    y0 = np.random.default_rng(4).standard_normal((T, N)).astype(np.float32)
    got = b2.tc_gemm(qt, packed, x, N, K, y0=y0)
    want = y0 + f16_round(quantised_acts(qt, x)).astype(np.float64) @ f16_round(wd).astype(np.float64).T
    assert np.abs(got - want).max() <= 1e-4 * np.abs(want).max()


def test_tc_gemm_rejects_bad_shapes(b2, synth):
    packed, wd, x = make(synth, qn.Q8_0, 128, 96, 4, seed=1)
    with pytest.raises(b2.B2Error):
        b2.tc_gemm(qn.Q8_0, packed, x, 128, 96)   # K not a multiple of 64
