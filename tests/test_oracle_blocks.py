"""The oracle's block arithmetic against llama.cpp's own Python package (gguf-py) and the committed
golden fixtures -- this is what pins the oracle (SURVEY 8c tier T1).  CPU only."""
import hashlib
import os

import gguf
import numpy as np
import pytest

from oracle import quants_np as qn
from oracle import ref_model as rm

GG = {qn.Q8_0: gguf.GGMLQuantizationType.Q8_0, qn.Q4_K: gguf.GGMLQuantizationType.Q4_K,
      qn.Q5_K: gguf.GGMLQuantizationType.Q5_K, qn.Q6_K: gguf.GGMLQuantizationType.Q6_K}
GOLDEN = os.path.join(os.path.dirname(__file__), "golden")


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


@pytest.mark.parametrize("qt", sorted(GG))
def test_dequant_numpy_and_c_match_gguf_py(synth, qt):
    rng = np.random.default_rng(100 + qt)
    blocks = synth.random_blocks(qt, 2048, rng, edge_cases=True)
    want = gguf.dequantize(blocks, GG[qt])
    assert np.array_equal(bits(qn.dequantize(blocks, qt)), bits(want))
    assert np.array_equal(bits(rm.dequantize(blocks, qt, want.size)), bits(want).reshape(-1))


@pytest.mark.parametrize("qt", sorted(GG))
def test_dequant_fully_random_bytes(qt):
    """Any byte pattern whose fp16 fields are finite is a valid block: uniform bytes, NaN/Inf scales patched."""
    rng = np.random.default_rng(qt)
    bb = qn.BLOCK[qt][1]
    raw = rng.integers(0, 256, size=(4096, bb), dtype=np.uint8)
    for off in ({qn.Q8_0: [0], qn.Q4_K: [0, 2], qn.Q5_K: [0, 2], qn.Q6_K: [208]}[qt]):
        h = raw[:, off:off + 2].copy().view(np.uint16)
        h[(h & 0x7C00) == 0x7C00] &= 0x3FFF  # clear Inf/NaN exponents
        raw[:, off:off + 2] = h.view(np.uint8)
    want = gguf.dequantize(raw, GG[qt])
    assert np.array_equal(bits(qn.dequantize(raw, qt)), bits(want))
    assert np.array_equal(bits(rm.dequantize(raw, qt, want.size)), bits(want).reshape(-1))


@pytest.mark.parametrize("qt", sorted(GG))
def test_golden_fixture(synth, qt):
    """Committed vectors (tests/golden/make_block_golden.py): blocks regenerated from the seed must hash to the
    recorded input digest, and the oracle must reproduce the recorded gguf-py output digest."""
    g = np.load(os.path.join(GOLDEN, "blocks_golden.npz"))
    name = synth.TYPE_NAME[qt]
    blocks = g[f"{name}_blocks"]
    assert hashlib.sha256(blocks.tobytes()).hexdigest() == str(g[f"{name}_in_sha256"])
    out = qn.dequantize(blocks, qt)
    assert hashlib.sha256(out.tobytes()).hexdigest() == str(g[f"{name}_out_sha256"])
    assert np.array_equal(bits(out[:4]), bits(g[f"{name}_first_rows"]))
    assert hashlib.sha256(rm.dequantize(blocks, qt, out.size).tobytes()).hexdigest() == str(g[f"{name}_out_sha256"])


def test_q8_0_quantiser_matches_gguf_py():
    rng = np.random.default_rng(5)
    x = (rng.standard_normal(32 * 513) * rng.uniform(1e-3, 30.0)).astype(np.float32)
    x[:32] = 0.0                      # all-zero block: d = 0
    x[32:64] = 127.0                  # exact multiples
    x[64] = 0.5 * (np.abs(x[64:96]).max() / 127)  # a .5 tie -> rounds away from zero
    want = gguf.quantize(x, gguf.GGMLQuantizationType.Q8_0)
    assert np.array_equal(qn.quantize_q8_0(x), want)
    assert np.array_equal(rm.quantize_q8_0(x), want)


def test_q8_K_numpy_matches_c():
    rng = np.random.default_rng(6)
    x = rng.standard_normal(256 * 40).astype(np.float32)
    x[:256] = 0.0
    x[256:512] = -3.0                 # every |x| equal: first element is the max, negative
    x[512] = 5.0; x[513] = -5.0       # tie between +max and -max: the first wins
    d, qs, bs = qn.quantize_q8_K(x)
    raw = rm.quantize_q8_K(x).reshape(-1, 292)
    assert np.array_equal(raw[:, 0:4].copy().view(np.float32).reshape(-1), d)
    assert np.array_equal(raw[:, 4:260].copy().view(np.int8).reshape(-1), qs)
    assert np.array_equal(raw[:, 260:].copy().view(np.int16).reshape(-1), bs)
    assert qs.max() <= 127 and qs.min() >= -127
    assert d[0] == 0 and not qs[:256].any()


@pytest.mark.parametrize("qt", [qn.Q4_K, qn.Q5_K, qn.Q6_K])
def test_c_vec_dot_matches_exact_integer_pieces(synth, qt):
    """ggml_ref.c's dot products equal the exact per-block integer sums evaluated in float64 (to fp32 rounding)."""
    rng = np.random.default_rng(70 + qt)
    N, K = 16, 2048
    W = synth.random_blocks(qt, N * K // 256, rng).reshape(N, -1)
    x = rng.standard_normal(K).astype(np.float32)
    y = rm.gemv(qt, W, x, N, K)
    d8, q8, _ = qn.quantize_q8_K(x)
    exact = np.array([qn.int_dot_k(W[r], qt, d8, q8).sum() for r in range(N)])
    assert np.allclose(y, exact, rtol=2e-6, atol=2e-6 * np.abs(exact).max())


@pytest.mark.parametrize("qt", sorted(GG))
def test_gemv_close_to_fp32_dequant_product(synth, qt):
    """Activation quantisation noise is small: int8-activation result within ~1% of dequant(W) @ x."""
    rng = np.random.default_rng(80 + qt)
    N, K = 64, 1024
    be = qn.BLOCK[qt][0]
    W = synth.random_blocks(qt, N * K // be, rng).reshape(N, -1)
    x = rng.standard_normal(K).astype(np.float32)
    y = rm.gemv(qt, W, x, N, K)
    yf = qn.dequantize(W, qt).astype(np.float64) @ x.astype(np.float64)
    assert np.abs(y - yf).max() < 0.02 * np.abs(yf).max()
