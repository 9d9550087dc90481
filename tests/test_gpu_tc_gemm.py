"""b2_tc_gemm (tcgen05 prompt-phase mat-mul) against the oracle's dequantised weights.

The kernel multiplies bf16 copies of the dequantised matrix and of the activations with fp32 accumulation, so the
exact reference is: oracle-dequantised W rounded to bf16, X rounded to bf16, products summed in float64.  Against
that the only difference is fp32 summation order (tolerance 2e-5 of the row's absolute dot product).  Against the
un-rounded fp32 mat-mul the bf16 rounding shows: each factor carries 2^-9 relative error, random in sign, so the
result stays within a few 1e-3 of sum |w||x| -- the second assertion states that bound."""
import numpy as np
import pytest

from oracle import quants_np as qn

pytestmark = pytest.mark.gpu

TYPES = [qn.Q4_K, qn.Q5_K, qn.Q6_K, qn.Q8_0]


def bf16_round(a: np.ndarray) -> np.ndarray:
    u = np.ascontiguousarray(a, np.float32).view(np.uint32).astype(np.uint64)
    u = (u + 0x7FFF + ((u >> 16) & 1)) & 0xFFFF0000
    return u.astype(np.uint32).view(np.float32).reshape(a.shape)


def make(synth, qt, N, K, T, seed):
    rng = np.random.default_rng(seed)
    packed = synth.random_blocks(qt, N * (K // qn.BLOCK[qt][0]), rng).reshape(N, -1)
    wd = qn.dequantize(packed, qt).reshape(N, K).astype(np.float32)
    x = rng.standard_normal((T, K)).astype(np.float32)
    x[:, ::97] *= 8.0
    return packed, wd, x


@pytest.mark.parametrize("qt", TYPES)
@pytest.mark.parametrize("N,K,T", [(128, 256, 128), (256, 512, 40), (384, 1024, 300), (1024, 2048, 512), (200, 256, 1), (4096, 4096, 257)])
def test_tc_gemm_matches_bf16_reference(b2, synth, qt, N, K, T):
    packed, wd, x = make(synth, qt, N, K, T, seed=N + K + T + qt)
    got = b2.tc_gemm(qt, packed, x, N, K)
    wb, xb = bf16_round(wd).astype(np.float64), bf16_round(x).astype(np.float64)
    want = xb @ wb.T
    mag = np.abs(xb) @ np.abs(wb).T + 1e-30
    assert np.abs(got - want).max() <= 2e-5 * mag.max(), float(np.abs(got - want).max() / mag.max())
    assert (np.abs(got - want) <= 2e-5 * mag + 1e-6).all()
    full = x.astype(np.float64) @ wd.astype(np.float64).T
    assert (np.abs(got - full) <= 4e-3 * mag).all()


def test_tc_gemm_accumulates_into_the_output(b2, synth):
    qt, N, K, T = qn.Q4_K, 512, 512, 200
    packed, wd, x = make(synth, qt, N, K, T, seed=3)
    y0 = np.random.default_rng(4).standard_normal((T, N)).astype(np.float32)
    got = b2.tc_gemm(qt, packed, x, N, K, y0=y0)
    want = y0 + bf16_round(x).astype(np.float64) @ bf16_round(wd).astype(np.float64).T
    assert np.abs(got - want).max() <= 1e-4 * np.abs(want).max()


def test_tc_gemm_rejects_bad_shapes(b2, synth):
    packed, wd, x = make(synth, qn.Q8_0, 128, 96, 4, seed=1)
    with pytest.raises(b2.B2Error):
        b2.tc_gemm(qn.Q8_0, packed, x, 128, 96)   # K not a multiple of 64
