"""Paged KV cache: with a pool smaller than n_slots x n_ctx the slots' positions live in pages of
256 mapped on demand.  Sequences whose pages are scattered and recycled decode the same logits as the oracle on every
path that touches the cache: prompt pass (KV store, causal attention), wide and narrow decode batches (tensor-core
and int8 mat-mul epilogues, batched attention), pages mapped in the middle of a generation, exhaustion and release."""
import numpy as np
import pytest

from oracle import ref_model as rm

pytestmark = pytest.mark.gpu
TOL = 0.03


def close(got, want):
    return np.abs(got - want).max() <= TOL * np.abs(want).max()


@pytest.mark.parametrize("name", ["tiny-llama-q4_k_m", "tiny-qwen2-q8_0"])
def test_scattered_and_recycled_pages_decode_like_the_oracle(b2, tiny_models, monkeypatch, name):
    monkeypatch.setenv("B2_KV_POOL_TOKENS", str(6 * 256))          # slot 0 keeps 2 pages; 4 for everybody else
    path = tiny_models[name]
    m = b2.Model(path)
    ctx = b2.Context(m, 512, n_slots=4)
    assert ctx.kv_pages() == (6, 4, 256, True)
    ref = rm.RefGGUFModel(path, n_ctx=512)
    rng = np.random.default_rng(12)

    def oracle(tokens):
        for p, t in enumerate(tokens):
            tok, lg = ref.decode(t, p, want_logits=p == len(tokens) - 1)
        return tok, lg

    seq_a = rng.integers(3, m.info.vocab, size=250).tolist()        # slot 2: one page, will grow into a second
    seq_b = rng.integers(3, m.info.vocab, size=300).tolist()        # slot 1: two pages
    seq_c = rng.integers(3, m.info.vocab, size=270).tolist()        # slot 3: two recycled pages
    _, la = ctx.prefill(seq_a, slot=2)
    assert close(la, oracle(seq_a)[1]) and ctx.kv_pages()[1] == 3
    # generation across the page boundary at position 256: narrow batches (int8 kernel), the page is mapped on the way
    ext = rng.integers(3, m.info.vocab, size=12).tolist()
    for i, t in enumerate(ext):
        ids, lg = ctx.decode_batch([t], [250 + i], slots=[2])
    seq_a = seq_a + ext
    assert close(lg[0], oracle(seq_a)[1]) and ctx.kv_pages()[1] == 2
    _, lb = ctx.prefill(seq_b, slot=1)
    assert close(lb, oracle(seq_b)[1]) and ctx.kv_pages()[1] == 0
    with pytest.raises(b2.B2Error) as e:
        ctx.prefill(seq_c[:10], slot=3)                             # no page left
    assert e.value.args[0] == -8 or "pool" in str(e.value)
    ctx.slot_truncate(1, 0)                                         # sequence B ends: its pages return
    assert ctx.kv_pages()[1] == 2
    _, lc = ctx.prefill(seq_c, slot=3)                              # ... and are reused, in another order
    assert close(lc, oracle(seq_c)[1]) and ctx.kv_pages()[1] == 0
    # both live sequences advance together (two columns: int8 kernel), then as part of a wide batch (tcgen05 pass)
    ta, tc_ = int(rng.integers(3, m.info.vocab)), int(rng.integers(3, m.info.vocab))
    ids, lg = ctx.decode_batch([ta, tc_], [len(seq_a), len(seq_c)], slots=[2, 3])
    seq_a.append(ta); seq_c.append(tc_)
    assert close(lg[0], oracle(seq_a)[1]) and close(lg[1], oracle(seq_c)[1])
    filler = rng.integers(3, m.info.vocab, size=30).tolist()        # 30 more columns: a fresh sequence in slot 0
    ta, tc_ = int(rng.integers(3, m.info.vocab)), int(rng.integers(3, m.info.vocab))
    ids, lg = ctx.decode_batch([ta, tc_] + filler, [len(seq_a), len(seq_c)] + list(range(30)), slots=[2, 3] + [0] * 30)
    seq_a.append(ta); seq_c.append(tc_)
    assert close(lg[0], oracle(seq_a)[1]) and close(lg[1], oracle(seq_c)[1]) and close(lg[31], oracle(filler)[1])
    # cutting a sequence back to a shared prefix keeps the pages that hold it
    ctx.slot_truncate(2, 256)
    assert ctx.kv_pages()[1] == 1
    ctx.slot_truncate(2, 0)
    assert ctx.kv_pages()[1] == 2
    ctx.free(); m.free(); ref.close()


def test_default_pool_is_statically_mapped(b2, tiny_models):
    m = b2.Model(tiny_models["tiny-llama-q4_k_m"])
    ctx = b2.Context(m, 300, n_slots=3)
    total, free, page, dynamic = ctx.kv_pages()
    assert (total, page, dynamic) == (6, 256, False)                # 2 pages per slot, every slot at full length
    ctx.slot_truncate(1, 0)                                         # a no-op
    ctx.decode_batch([5, 6], [299, 0], slots=[2, 1])
    ctx.free(); m.free()


def test_position_beyond_an_unwritten_page_is_refused(b2, tiny_models, monkeypatch):
    """On an on-demand pool a pass maps only the page of the position it writes; attention then walks every page below
    it.  Decoding at position 300 of an empty slot must be turned away on the host (B2_ERANGE), not read page -1."""
    monkeypatch.setenv("B2_KV_POOL_TOKENS", str(6 * 256))
    m = b2.Model(tiny_models["tiny-llama-q4_k_m"])
    ctx = b2.Context(m, 512, n_slots=4)                              # the model's trained context is 512: 2 pages per slot
    assert ctx.kv_pages()[3]
    free_before = ctx.kv_pages()[1]
    with pytest.raises(b2.B2Error) as e:
        ctx.decode_batch([5], [300], slots=[1])
    assert e.value.code == -6 and "never written" in str(e.value)
    with pytest.raises(b2.B2Error):
        ctx.prefill([5, 6, 7], slot=2, pos0=300)
    assert ctx.kv_pages()[1] == free_before                          # nothing was mapped by the refused calls
    ctx.decode_batch([5], [0], slots=[1])                            # from position 0 upwards it works
    ctx.free(); m.free()
