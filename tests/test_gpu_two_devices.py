"""One process, two GPUs: kernel attributes (the opt-in to more than 48 KB of dynamic shared memory) and the cached SM
counts are per device, so a model loaded on device 1 after one on device 0 must run every path -- per-token kernel, int8
batch pass, tensor-core prompt pass -- and give the same numbers.  Skipped on a box with a single GPU."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_second_device_in_the_same_process(b2, tiny_models):
    if b2.device_count() < 2:
        pytest.skip("needs two GPUs")
    path = tiny_models["tiny-llama-q4_k_m"]
    rng = np.random.default_rng(21)
    got = []
    for dev in (0, 1):
        m = b2.Model(path, dev)
        ctx = b2.Context(m, 256, n_slots=4)
        prompt = rng.integers(3, m.info.vocab, size=5).tolist() if dev == 0 else prompt
        long_prompt = rng.integers(3, m.info.vocab, size=40).tolist() if dev == 0 else long_prompt
        tok, lg = ctx.decode(prompt, 0)                                  # per-token kernel
        ids = ctx.generate_greedy(6)
        bids, blg = ctx.decode_batch([5, 6, 7], [0, 0, 0], slots=[1, 2, 3])   # int8 batch pass
        ptok, plg = ctx.prefill(long_prompt, slot=1)                     # tensor-core prompt pass
        got.append((tok, lg, ids, bids, blg, ptok, plg))
        ctx.free(); m.free()
    a, b = got
    assert a[0] == b[0] and a[2] == b[2] and a[3] == b[3] and a[5] == b[5]
    assert np.array_equal(a[1], b[1]) and np.array_equal(a[4], b[4]) and np.array_equal(a[6], b[6])
