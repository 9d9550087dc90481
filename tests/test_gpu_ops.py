"""Stand-alone operators through the C ABI against the oracle: RMSNorm, RoPE, decode attention, argmax."""
import numpy as np
import pytest

from oracle import ref_model as rm

pytestmark = pytest.mark.gpu


def test_rmsnorm(b2):
    rng = np.random.default_rng(0)
    for n, rows, eps in [(256, 1, 1e-5), (4096, 3, 1e-5), (1536, 2, 1e-6), (8192, 1, 1e-5)]:
        x = (rng.standard_normal((rows, n)) * 3).astype(np.float32)
        w = rng.uniform(0.5, 1.5, n).astype(np.float32)
        dx, dw, dy = b2.DevBuf.from_numpy(x), b2.DevBuf.from_numpy(w), b2.DevBuf(x.nbytes)
        b2.check(b2.lib.b2_rmsnorm(dx.ptr, dw.ptr, dy.ptr, n, rows, eps, None))
        got = dy.to_numpy(np.float32).reshape(rows, n)
        want = np.stack([rm.rmsnorm(x[r], w, eps) for r in range(rows)])
        # double-accumulated sum of squares on both sides: at most 1 ulp apart through sqrt/div
        assert np.allclose(got, want, rtol=3e-7, atol=0)


@pytest.mark.parametrize("mode", [0, 2])
@pytest.mark.parametrize("with_factors", [False, True])
def test_rope(b2, mode, with_factors):
    rng = np.random.default_rng(1)
    n_heads, hd = 8, 128
    x = rng.standard_normal((n_heads, hd)).astype(np.float32)
    ff = rng.uniform(1.0, 8.0, hd // 2).astype(np.float32) if with_factors else None
    for pos, base in [(0, 1e4), (1, 1e4), (127, 5e5), (2047, 1e6)]:
        dx = b2.DevBuf.from_numpy(x)
        dff = b2.DevBuf.from_numpy(ff) if with_factors else None
        b2.check(b2.lib.b2_rope(dx.ptr, n_heads, hd, hd, mode, pos, base, dff.ptr if dff else None, None))
        got = dx.to_numpy(np.float32).reshape(n_heads, hd)
        want = rm.rope(x, n_heads, hd, hd, mode, rm.rope_cache(pos, hd, base, ff))
        # device cosf/sinf are within 2 ulp of libm's; angles up to 2047 rad
        assert np.allclose(got, want, rtol=0, atol=3e-6 * np.abs(x).max())


def _attn_ref(q, K, V, n_head, n_kv, hd, T):
    """oracle/ggml_ref.c ref_attention restated in numpy (q, probabilities rounded to fp16; fp32 accumulate)."""
    g = n_head // n_kv
    out = np.zeros((n_head, hd), np.float32)
    qh = q.astype(np.float16).astype(np.float32)
    for h in range(n_head):
        k = K[h // g, :T].astype(np.float32)
        v = V[h // g, :T].astype(np.float32)
        s = (k @ qh[h]).astype(np.float32) * np.float32(1.0 / np.sqrt(np.float32(hd)))
        e = np.exp(s - s.max()).astype(np.float32)
        p = (e * np.float32(1.0 / e.astype(np.float64).sum())).astype(np.float32)
        out[h] = p.astype(np.float16).astype(np.float32) @ v
    return out


@pytest.mark.parametrize("n_head,n_kv,hd", [(32, 8, 128), (32, 4, 64), (12, 2, 128), (8, 8, 64), (64, 8, 128)])
@pytest.mark.parametrize("T", [1, 2, 127, 128, 1000, 1024])
def test_attention_single_split(b2, n_head, n_kv, hd, T):
    rng = np.random.default_rng(T + n_head)
    n_ctx = 1024
    q = rng.standard_normal((n_head, hd)).astype(np.float32)
    K = rng.standard_normal((n_kv, n_ctx, hd)).astype(np.float16)
    V = rng.standard_normal((n_kv, n_ctx, hd)).astype(np.float16)
    dq, dk, dv = b2.DevBuf.from_numpy(q), b2.DevBuf.from_numpy(K), b2.DevBuf.from_numpy(V)
    do = b2.DevBuf(q.nbytes)
    sc = b2.DevBuf(b2.lib.b2_attn_scratch_bytes(n_head, hd, T), zero=True)
    b2.check(b2.lib.b2_attn_decode(dq.ptr, dk.ptr, dv.ptr, do.ptr, n_head, n_kv, hd, n_ctx, T, sc.ptr, None))
    got = do.to_numpy(np.float32).reshape(n_head, hd)
    want = _attn_ref(q, K, V, n_head, n_kv, hd, T)
    # same fp16 roundings; fp32 summation order differs, and a probability that lands on an fp16 rounding
    # boundary may flip by one fp16 ulp (2^-11 relative) times |v| <= ~4
    assert np.abs(got - want).max() < 2e-4 * max(1.0, np.abs(want).max())


@pytest.mark.parametrize("T", [1025, 3000, 8192])
def test_attention_multi_split_twice(b2, T):
    """Several sequence splits merged by the last-arriving CTA; run twice on the same scratch to check the
    arrival counters reset themselves."""
    rng = np.random.default_rng(T)
    n_head, n_kv, hd, n_ctx = 32, 8, 128, 8192
    q = rng.standard_normal((n_head, hd)).astype(np.float32)
    K = (rng.standard_normal((n_kv, n_ctx, hd)) * 0.5).astype(np.float16)
    V = rng.standard_normal((n_kv, n_ctx, hd)).astype(np.float16)
    dq, dk, dv = b2.DevBuf.from_numpy(q), b2.DevBuf.from_numpy(K), b2.DevBuf.from_numpy(V)
    do = b2.DevBuf(q.nbytes)
    sc = b2.DevBuf(b2.lib.b2_attn_scratch_bytes(n_head, hd, T), zero=True)
    want = _attn_ref(q, K, V, n_head, n_kv, hd, T)
    for _ in range(2):
        b2.check(b2.lib.b2_dev_memset(do.ptr, 0, q.nbytes, None))
        b2.check(b2.lib.b2_attn_decode(dq.ptr, dk.ptr, dv.ptr, do.ptr, n_head, n_kv, hd, n_ctx, T, sc.ptr, None))
        got = do.to_numpy(np.float32).reshape(n_head, hd)
        # probabilities are rounded to fp16 before the split merge: up to 2^-11 relative on each
        assert np.abs(got - want).max() < 2e-3 * np.abs(want).max()


def test_attention_rejects_bad_config(b2):
    d = b2.DevBuf(1 << 16)
    assert b2.lib.b2_attn_decode(d.ptr, d.ptr, d.ptr, d.ptr, 32, 8, 96, 128, 4, d.ptr, None) < 0
    assert b2.lib.b2_attn_decode(d.ptr, d.ptr, d.ptr, d.ptr, 32, 5, 128, 128, 4, d.ptr, None) < 0
    assert b2.lib.b2_attn_decode(d.ptr, d.ptr, d.ptr, d.ptr, 32, 8, 128, 128, 129, d.ptr, None) < 0


@pytest.mark.parametrize("n", [1, 5, 32000, 128256, 151936])
def test_argmax(b2, n):
    rng = np.random.default_rng(n)
    x = rng.standard_normal(n).astype(np.float32)
    if n > 100:
        x[77] = x.max() + 1; x[n - 3] = x[77]      # duplicate maximum: the lower index wins
    dx, dt = b2.DevBuf.from_numpy(x), b2.DevBuf(4)
    sc = b2.DevBuf(b2.lib.b2_argmax_scratch_bytes(n), zero=True)
    for _ in range(2):
        b2.check(b2.lib.b2_sample_greedy(dx.ptr, n, dt.ptr, sc.ptr, None))
        assert dt.to_numpy(np.int32)[0] == int(np.argmax(x))
