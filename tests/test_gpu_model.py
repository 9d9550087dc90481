"""End-to-end decode through b2_model_load / b2_ctx_decode against the oracle's C restatement of the
ggml CPU path (same GGUF file, same token ids)."""
import numpy as np
import pytest

from oracle import ref_model as rm

pytestmark = pytest.mark.gpu
MODELS = ["tiny-llama-q4_k_m", "tiny-llama3-q4_k_m", "tiny-llama-q5_k_m", "tiny-qwen2-q8_0"]


def kl(p_logits, q_logits):
    p = p_logits.astype(np.float64); q = q_logits.astype(np.float64)
    p -= p.max(); q -= q.max()
    lp = p - np.log(np.exp(p).sum()); lq = q - np.log(np.exp(q).sum())
    return float((np.exp(lp) * (lp - lq)).sum())


@pytest.mark.parametrize("name", MODELS)
def test_upload_is_byte_identical(b2, tiny_models, name):
    import gguf
    m = b2.Model(tiny_models[name])
    r = gguf.GGUFReader(tiny_models[name])
    for t in r.tensors:
        assert m.tensor_device_bytes(t.name) == bytes(np.ascontiguousarray(t.data).view(np.uint8)), t.name
    assert m.info.stream_bytes > 0
    m.free()


@pytest.mark.parametrize("name", MODELS)
def test_logits_and_greedy_match_oracle(b2, tiny_models, name):
    path = tiny_models[name]
    ref = rm.RefGGUFModel(path, n_ctx=512)
    m = b2.Model(path)
    ctx = b2.Context(m, 512)
    rng = np.random.default_rng(11)
    for trial in range(3):
        prompt = rng.integers(3, m.info.vocab, size=[5, 33, 128][trial]).tolist()
        n_new = 48
        want_ids, want_logits = ref.greedy(prompt, n_new)
        got_ids, got_logits = ctx.greedy(prompt, n_new)
        # logits tolerance (north_star: "logits within a stated fp tolerance (max-abs and KL)"):
        # integer block sums are identical; fp32 summation order and device expf/cosf differ.
        scale = np.abs(want_logits).max()
        assert np.abs(got_logits - want_logits).max() <= 2e-4 * scale, (np.abs(got_logits - want_logits).max(), scale)
        assert kl(want_logits, got_logits) < 1e-7
        # greedy ids identical unless the oracle's own top-2 margin is inside that tolerance
        if got_ids != want_ids:
            i = next(k for k in range(n_new) if got_ids[k] != want_ids[k])
            # replay the oracle to the divergence point and look at its margin
            ref2 = rm.RefGGUFModel(path, n_ctx=512)
            seq = prompt + want_ids[:i]
            for p, t in enumerate(seq):
                _, lg = ref2.decode(t, p, want_logits=(p == len(seq) - 1))
            top2 = np.sort(lg)[-2:]
            pytest.fail(f"{name}: greedy diverged at generated token {i}: got {got_ids[i]} want {want_ids[i]}; "
                        f"oracle top-1/top-2 margin {top2[1] - top2[0]:.3e}")
    ctx.free(); m.free(); ref.close()


@pytest.mark.parametrize("name", MODELS[:2])
def test_hidden_state_and_position_restart(b2, tiny_models, name):
    path = tiny_models[name]
    ref = rm.RefGGUFModel(path, n_ctx=256)
    m = b2.Model(path)
    ctx = b2.Context(m, 256)
    toks = [1, 17, 300, 5, 9]
    for p, t in enumerate(toks):
        ref.decode(t, p, want_logits=False)
    ctx.decode(toks, 0, want_logits=False)
    h_ref, h_got = ref.hidden(), ctx.hidden()
    assert np.abs(h_got - h_ref).max() <= 1e-4 * np.abs(h_ref).max()
    # decoding the same prefix again from position 0 reproduces the same logits bit for bit (idempotence)
    n1, l1 = ctx.decode(toks, 0)
    n2, l2 = ctx.decode(toks, 0)
    assert n1 == n2 and np.array_equal(l1, l2)
    # token-by-token == all at once
    for p, t in enumerate(toks):
        n3, l3 = ctx.decode([t], p)
    assert n3 == n1 and np.array_equal(l3, l1)
    ctx.free(); m.free(); ref.close()


def test_decode_rejects_out_of_range(b2, tiny_models):
    m = b2.Model(tiny_models["tiny-llama-q4_k_m"])
    ctx = b2.Context(m, 64)
    assert ctx.n_ctx == 64
    with pytest.raises(b2.B2Error):
        ctx.decode([m.info.vocab], 0)
    with pytest.raises(b2.B2Error):
        ctx.decode([-1], 0)
    with pytest.raises(b2.B2Error):
        ctx.decode([1] * 65, 0)
    with pytest.raises(b2.B2Error):
        ctx.decode([1], 64)
    ctx.decode([1] * 64, 0)
    with pytest.raises(b2.B2Error):
        ctx.generate_greedy(1)
    ctx.free(); m.free()


def test_graph_and_eager_paths_agree(b2, tiny_models, monkeypatch):
    """CUDA-graph + programmatic-dependent-launch replay gives the same bits as plain stream launches."""
    path = tiny_models["tiny-llama-q4_k_m"]
    m = b2.Model(path)
    prompt = list(range(3, 40))
    ctx = b2.Context(m, 256)
    ids_g, lg_g = ctx.greedy(prompt, 16)
    ctx.free()
    monkeypatch.setenv("B2_NO_GRAPH", "1")
    monkeypatch.setenv("B2_NO_PDL", "1")
    ctx = b2.Context(m, 256)
    ids_e, lg_e = ctx.greedy(prompt, 16)
    ctx.free(); m.free()
    assert ids_g == ids_e and np.array_equal(lg_g, lg_e)
