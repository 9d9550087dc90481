"""Device-side sampling (the non-greedy half): llama-server's defaults when a request carries no
sampling fields -- temperature 0.8, top-k 40, top-p 0.95, min-p 0.05 -- are what the README's curl
(ramalama-models/README.md:278-285) and OpenWebUI get.  The filter chain is restated here in numpy (float64) from the
server's host sampler (csrc/server/engine.h: Engine::sample, itself llama.cpp's default sampler order); the device
sampler must (1) keep exactly the candidate set that restatement keeps, (2) draw from its distribution (chi-square on
20 000 draws), (3) repeat itself for a (seed, draw index) pair and change with either, (4) fall back to the first index
of the maximum when greedy, (5) serve every decode entry point of a context slot that has a sampler."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def target_distribution(logits, temperature=0.8, top_k=40, top_p=0.95, min_p=0.05, k_max=1024):
    """{token id: probability} after top-k -> softmax(T) -> min-p -> top-p, as the server applies them."""
    V = logits.size
    k = k_max if top_k <= 0 or top_k > k_max else top_k
    k = min(k, V)
    order = np.lexsort((np.arange(V), -logits.astype(np.float64)))[:k]      # value descending, id ascending
    v = logits[order].astype(np.float64)
    pr = np.exp((v - v[0]) / temperature)
    pr /= pr.sum()
    keep = k
    if min_p > 0:
        keep = max(1, int(np.argmin(pr >= min_p * pr[0])) if not (pr >= min_p * pr[0]).all() else k)
    if top_p < 1.0:
        acc = np.cumsum(pr[:keep])
        n = int(np.searchsorted(acc, top_p, side="left")) + 1
        keep = max(1, min(keep, n))
    p = pr[:keep] / pr[:keep].sum()
    return dict(zip(order[:keep].tolist(), p.tolist()))


@pytest.mark.parametrize("V,params", [(32000, {}), (128256, {}), (1024, dict(top_k=0, top_p=1.0, min_p=0.0, temperature=1.0)),
                                       (151936, dict(top_k=8, temperature=1.5, top_p=0.9, min_p=0.0)), (5000, dict(top_k=100, top_p=0.5, min_p=0.2))])
def test_draws_follow_the_server_s_distribution(b2, V, params):
    rng = np.random.default_rng(V)
    logits = (rng.standard_normal(V) * 2.0).astype(np.float32)
    logits[rng.integers(0, V, size=5)] += 6.0                      # a few clear favourites, as trained models have
    want = target_distribution(logits, **params)
    n = 20000
    got = b2.sample(logits, n, seed=1234, **params)
    ids, counts = np.unique(got, return_counts=True)
    assert set(ids.tolist()) <= set(want), "a token outside the kept set was drawn"
    # chi-square over the cells with an expected count of at least 5 (the rest pooled)
    exp = np.array([want[i] * n for i in want]); obs = np.array([counts[ids == i][0] if i in ids else 0 for i in want])
    big = exp >= 5
    e = np.append(exp[big], exp[~big].sum()); o = np.append(obs[big], obs[~big].sum())
    e, o = e[e > 0], o[e > 0]
    chi2 = float(((o - e) ** 2 / e).sum()); dof = max(1, len(e) - 1)
    assert chi2 < dof + 6 * np.sqrt(2 * dof) + 10, (chi2, dof)     # far tail of chi-square(dof): a wrong distribution is off by hundreds


def test_same_seed_and_index_same_token(b2):
    rng = np.random.default_rng(1)
    logits = rng.standard_normal(50000).astype(np.float32)
    a = b2.sample(logits, 512, seed=7)
    assert np.array_equal(a, b2.sample(logits, 512, seed=7))
    assert np.array_equal(a[100:], b2.sample(logits, 412, seed=7, first_draw=100))        # the draw index, not the row, decides
    assert not np.array_equal(a, b2.sample(logits, 512, seed=8))


def test_greedy_settings_return_the_first_maximum(b2):
    rng = np.random.default_rng(2)
    logits = rng.standard_normal(40000).astype(np.float32)
    logits[[31000, 17, 2999]] = logits.max() + 1.0                  # a three-way tie: the lowest id wins
    for kw in (dict(temperature=0.0), dict(top_k=1), dict(temperature=-1.0)):
        assert (b2.sample(logits, 8, **kw) == 17).all(), kw
    # top-k keeps the lowest ids among ties at the threshold
    t = b2.sample(logits, 4000, top_k=2, temperature=5.0, top_p=1.0, min_p=0.0)
    assert set(t.tolist()) == {17, 2999}


def test_every_decode_entry_point_uses_the_slot_s_sampler(b2, tiny_models):
    m = b2.Model(tiny_models["tiny-llama-q4_k_m"])
    V = m.info.vocab
    ctx = b2.Context(m, 256, n_slots=4)
    rng = np.random.default_rng(3)
    prompt = rng.integers(3, V, size=40).tolist()
    # greedy reference of every path
    g_tok, g_logits = ctx.prefill(prompt, slot=1)
    # a hot sampler on slot 1 only: its token must come from the distribution of the logits the call returns
    ctx.set_sampler(1, temperature=2.0, top_k=64, top_p=1.0, min_p=0.0, seed=99)
    seen = set()
    for trial in range(6):
        ctx.set_sampler(1, temperature=2.0, top_k=64, top_p=1.0, min_p=0.0, seed=99 + trial)
        tok, lg = ctx.prefill(prompt, slot=1)
        assert np.array_equal(lg, g_logits)
        assert tok in target_distribution(lg, temperature=2.0, top_k=64, top_p=1.0, min_p=0.0)
        seen.add(tok)
        # reproducible, and equal to the stand-alone sampler on the same logits with the draw index = the new position
        assert tok == b2.sample(lg, 1, temperature=2.0, top_k=64, top_p=1.0, min_p=0.0, seed=99 + trial, first_draw=len(prompt))[0]
    assert len(seen) > 1, "six seeds drew the same token at temperature 2"
    # batched decode: slot 1 samples, slot 2 (same tokens, no sampler) stays greedy
    ctx.prefill(prompt, slot=2)
    ids, lg = ctx.decode_batch([g_tok, g_tok], [40, 40], slots=[1, 2])
    assert ids[1] == int(np.argmax(lg[1]))
    assert ids[0] == b2.sample(lg[0], 1, temperature=2.0, top_k=64, top_p=1.0, min_p=0.0, seed=104, first_draw=41)[0]
    # slot 0: b2_ctx_decode and the device-resident loop
    ctx.set_sampler(0, temperature=1.5, top_k=50, top_p=1.0, min_p=0.0, seed=5)
    tok, lg = ctx.decode(prompt, 0)
    assert tok == b2.sample(lg, 1, temperature=1.5, top_k=50, top_p=1.0, min_p=0.0, seed=5, first_draw=40)[0]
    run_a = ctx.generate_greedy(12)
    tok2, _ = ctx.decode(prompt, 0)
    assert tok2 == tok and ctx.generate_greedy(12) == run_a          # same seed, same positions: the same continuation
    ctx.set_sampler(0, greedy=True)
    tok3, lg3 = ctx.decode(prompt, 0)
    assert tok3 == int(np.argmax(lg3))
    ctx.free(); m.free()
