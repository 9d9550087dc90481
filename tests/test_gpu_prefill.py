"""The prompt-phase path (b2_ctx_prefill: tcgen05 mat-muls on fp16 tiles + causal attention over the chunk) against
the oracle's token-by-token CPU decode of the same GGUF file.

The path feeds the tensor cores the reference's own quantised activations (Q8_K / Q8_0 blocks, d * q rounded to fp16)
and the dequantised weights rounded to fp16, so it differs from the reference by 2^-11 roundings plus the usual
rounding-flip amplification; the end-to-end tolerance is the decode path's (tests/test_gpu_model.py): max-abs <= 3 %
of the logit range, KL <= 1e-3 nat."""
import numpy as np
import pytest

from oracle import ref_model as rm

pytestmark = pytest.mark.gpu
MODELS = ["tiny-llama-q4_k_m", "tiny-llama3-q4_k_m", "tiny-llama-q5_k_m", "tiny-qwen2-q8_0"]
LOGIT_TOL = 0.03
KL_TOL = 1e-3


def kl(p_logits, q_logits):
    p = p_logits.astype(np.float64); q = q_logits.astype(np.float64)
    p -= p.max(); q -= q.max()
    lp = p - np.log(np.exp(p).sum()); lq = q - np.log(np.exp(q).sum())
    return float((np.exp(lp) * (lp - lq)).sum())


def close(got, want):
    return np.abs(got - want).max() <= LOGIT_TOL * np.abs(want).max() and kl(want, got) < KL_TOL


@pytest.mark.parametrize("name", MODELS)
@pytest.mark.parametrize("n", [48, 64, 65, 200, 333])
def test_prefill_logits_match_oracle(b2, tiny_models, name, n):
    path = tiny_models[name]
    ref = rm.RefGGUFModel(path, n_ctx=512)
    m = b2.Model(path)
    ctx = b2.Context(m, 512)
    toks = np.random.default_rng(n).integers(3, m.info.vocab, size=n).tolist()
    for p, t in enumerate(toks):
        tok, lr = ref.decode(t, p, want_logits=p == n - 1)
    got_tok, lg = ctx.prefill(toks)
    assert close(lg, lr), (float(np.abs(lg - lr).max() / np.abs(lr).max()), kl(lr, lg))
    top2 = np.partition(lr, -2)[-2:]
    assert got_tok == tok or float(top2.max() - top2.min()) <= 2 * LOGIT_TOL * np.abs(lr).max()
    # the cache the prompt path wrote serves the per-token kernel: continue token by token on both sides
    nxt = tok
    for i in range(8):
        tok_r, lr = ref.decode(nxt, n + i)
        tok_g, lg = ctx.decode([nxt], n + i)
        assert close(lg, lr), i
        nxt = tok_r
    ctx.free(); m.free(); ref.close()


@pytest.mark.parametrize("name", MODELS[:2])
def test_prefill_chunks_and_prefix_reuse_are_bit_identical(b2, tiny_models, name, monkeypatch):
    """Rows are independent in every kernel of the path and key blocks are aligned to absolute positions, so cutting
    the prompt into chunks (or prefilling on top of an existing prefix) reproduces the one-pass logits bit for bit."""
    path = tiny_models[name]
    m = b2.Model(path)
    toks = np.random.default_rng(5).integers(3, m.info.vocab, size=300).tolist()
    ctx = b2.Context(m, 512)
    t_all, l_all = ctx.prefill(toks)
    ctx.prefill(toks[:130], want_logits=False)
    t_two, l_two = ctx.prefill(toks[130:], pos0=130)
    ctx.free()
    assert t_two == t_all and np.array_equal(l_two, l_all)
    monkeypatch.setenv("B2_PREFILL_CHUNK", "160")   # > 128 rows per chunk: passes of <= 128 rows split K (another summation order)
    ctx = b2.Context(m, 512)
    t_ch, l_ch = ctx.prefill(toks)
    ctx.free(); m.free()
    assert t_ch == t_all and np.array_equal(l_ch, l_all)


def test_prefill_into_a_slot_then_batched_decode(b2, tiny_models):
    path = tiny_models["tiny-llama-q4_k_m"]
    ref = rm.RefGGUFModel(path, n_ctx=256)
    m = b2.Model(path)
    ctx = b2.Context(m, 256, n_slots=3)
    toks = np.random.default_rng(8).integers(3, m.info.vocab, size=100).tolist()
    other = np.random.default_rng(9).integers(3, m.info.vocab, size=70).tolist()
    for p, t in enumerate(toks):
        tok, lr = ref.decode(t, p, want_logits=p == len(toks) - 1)
    ctx.prefill(other, slot=1, want_logits=False)       # a neighbour sequence in another slot
    got_tok, lg = ctx.prefill(toks, slot=2)
    assert close(lg, lr)
    nxt = tok
    for i in range(6):
        tok_r, lr = ref.decode(nxt, 100 + i)
        ids, lgb = ctx.decode_batch([nxt, 5], [100 + i, 70 + i], slots=[2, 1])
        assert close(lgb[0], lr), i
        nxt = tok_r
    ctx.free(); m.free(); ref.close()


def test_decode_of_a_long_prompt_takes_the_prompt_path(b2, tiny_models):
    """b2_ctx_decode with >= 48 tokens runs the same kernels as b2_ctx_prefill on slot 0."""
    path = tiny_models["tiny-llama-q4_k_m"]
    m = b2.Model(path)
    toks = np.random.default_rng(2).integers(3, m.info.vocab, size=150).tolist()
    ctx = b2.Context(m, 256)
    t1, l1 = ctx.decode(toks, 0)
    t2, l2 = ctx.prefill(toks)
    assert t1 == t2 and np.array_equal(l1, l2)
    ids = ctx.generate_greedy(4)          # device-resident state (token, position) was left consistent
    assert len(ids) == 4
    ctx.free(); m.free()


def test_prefill_rejects_out_of_range(b2, tiny_models):
    m = b2.Model(tiny_models["tiny-llama-q4_k_m"])
    ctx = b2.Context(m, 64)
    with pytest.raises(b2.B2Error):
        ctx.prefill([1] * 65)
    with pytest.raises(b2.B2Error):
        ctx.prefill([1] * 10, pos0=60)
    with pytest.raises(b2.B2Error):
        ctx.prefill([m.info.vocab])
    with pytest.raises(b2.B2Error):
        ctx.prefill([1, 2], slot=1)
    ctx.free(); m.free()


@pytest.mark.parametrize("name", MODELS[:2] + MODELS[3:])
def test_several_prompts_in_one_pass(b2, tiny_models, name):
    """b2_ctx_prefill_multi: prompts of different lengths (one of them a tail after a reused prefix) share one pass over
    the weights; every sequence's last-position logits match the oracle's own decode of that sequence, and the caches
    serve the decode steps that follow."""
    path = tiny_models[name]
    m = b2.Model(path)
    ctx = b2.Context(m, 512, n_slots=8)
    ref = rm.RefGGUFModel(path, n_ctx=512)
    rng = np.random.default_rng(17)
    lens = [70, 33, 300, 5, 129, 64]
    seqs = [rng.integers(3, m.info.vocab, size=n).tolist() for n in lens]
    slots = [3, 1, 7, 2, 5, 6]
    ctx.prefill(seqs[2][:100], slot=7, want_logits=False)            # sequence 2 already has a 100-token prefix cached
    prompts = [s if i != 2 else s[100:] for i, s in enumerate(seqs)]
    pos0 = [0 if i != 2 else 100 for i in range(len(seqs))]
    ids, lg = ctx.prefill_multi(prompts, slots, pos0)
    want = []
    for i, s in enumerate(seqs):
        for p, t in enumerate(s):
            tok, lr = ref.decode(t, p, want_logits=p == len(s) - 1)
        want.append((tok, lr))
        assert close(lg[i], lr), (i, float(np.abs(lg[i] - lr).max() / np.abs(lr).max()))
        assert ids[i] == int(np.argmax(lg[i]))
    # one decode step of all six together on the caches the pass wrote
    nxt = [w[0] for w in want]
    ids2, lg2 = ctx.decode_batch(nxt, lens, slots=slots)
    for i, s in enumerate(seqs):
        for p, t in enumerate(s + [nxt[i]]):
            tok, lr = ref.decode(t, p, want_logits=p == len(s))
        assert close(lg2[i], lr), i
    with pytest.raises(b2.B2Error):
        ctx.prefill_multi([[1, 2], [3]], [1, 1])                     # the same slot twice
    with pytest.raises(b2.B2Error):
        ctx.prefill_multi([[1] * 300, [2] * 300], [1, 2])            # more rows than one pass takes (n_ctx = 512)
    ctx.free(); m.free(); ref.close()
