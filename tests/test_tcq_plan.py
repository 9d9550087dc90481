"""The stream-K decomposition of the fused-dequantisation mat-mul (tcq_gemm.cu) as host arithmetic: b2_tcq_plan.

The kernel's five roles (weight producer, activation producer, MMA issuer, dequantising warps, epilogue) each walk the
CTA's share on their own, and the epilogue derives its output slice from the first CTA that touches a tile -- so the
arithmetic has to hold for every shape: shares partition the (tile, block column) sequence, they differ by at most one
step, a tile's CTAs are consecutive, no tile needs more slices than the plan reserves.  No GPU needed."""
import importlib

import numpy as np
import pytest

b2 = importlib.import_module("llms-on-kubernetes_b200")

SHAPES = [
    ([4096, 1024, 1024], 4096),     # Llama-3.1-8B Q | K | V
    ([4096], 4096),                 # O
    ([14336, 14336], 4096),         # gate | up
    ([4096], 14336),                # down
    ([128256], 4096),               # lm_head
    ([8192 // 8, 1024 // 8, 1024 // 8], 8192),   # a 70B shard at TP 8
    ([200], 256), ([128], 256), ([512], 14336), ([2048, 256, 256], 2048), ([1], 256),
]


@pytest.mark.parametrize("seg_rows,K", SHAPES)
@pytest.mark.parametrize("n_sms", [148, 132, 7, 1])
def test_shares_partition_the_work_and_slices_suffice(seg_rows, K, n_sms):
    p = b2.tcq_plan(seg_rows, K, n_sms)
    nblk = K // 256
    tiles = sum(-(-r // 128) for r in seg_rows)
    assert p["tiles"] == tiles and 1 <= p["grid"] <= n_sms
    total = tiles * nblk
    sh = p["shares"]
    # contiguous, complete, balanced to one step
    assert sh[0, 0] == 0 and sh[-1, 1] == total and (sh[1:, 0] == sh[:-1, 1]).all()
    sizes = sh[:, 1] - sh[:, 0]
    assert sizes.min() >= 1 and sizes.max() - sizes.min() <= 1
    # per tile: the CTAs whose shares meet it are consecutive, start at first_cta, and number at most `slices`
    owner = np.empty(total, np.int64)
    for c, (lo, hi) in enumerate(sh):
        owner[lo:hi] = c
    worst = 0
    for t in range(tiles):
        ctas = np.unique(owner[t * nblk:(t + 1) * nblk])
        assert ctas[0] == p["first_cta"][t] and (np.diff(ctas) == 1).all()
        worst = max(worst, len(ctas))
    assert worst == p["slices"] and p["slices"] <= 8 or p["grid"] == 1


def test_skinny_matrices_use_fewer_ctas_rather_than_more_slices():
    # 4 tiles x 56 block columns on 148 SMs would cut a tile into 37 partial sums: the plan shrinks the grid instead
    p = b2.tcq_plan([512], 14336, 148)
    assert p["slices"] <= 8 and p["grid"] < 148


def test_bad_arguments_are_refused():
    with pytest.raises(b2.B2Error):
        b2.tcq_plan([128], 250)
    with pytest.raises(b2.B2Error):
        b2.tcq_plan([128, 128, 128, 128], 256)
    with pytest.raises(b2.B2Error):
        b2.tcq_plan([0], 256)
