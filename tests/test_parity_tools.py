"""oracle/parity.py -- the rules a GPU run is judged by -- checked on the CPU: the verdict logic on constructed cases, and the
measured noise floor (the oracle re-run with every residual update perturbed by one fp32 ulp) on a tiny model."""
import numpy as np

from oracle import parity as par
from oracle import ref_model as rm


def _logits(n=64, top=(10.0, 9.0), seed=0):
    lg = np.random.default_rng(seed).normal(0.0, 1.0, n).astype(np.float32)
    lg[3], lg[7] = top
    return lg


def test_logit_stats_of_equal_rows_is_zero_and_margin_is_top1_minus_top2():
    lg = _logits()
    st = par.logit_stats(lg, lg)
    assert st["err"] == 0.0 and abs(st["kl"]) < 1e-12
    assert abs(st["margin"] - 1.0 / st["range"]) < 1e-6


def test_identical_ids_pass_and_a_divergence_passes_only_inside_the_margin_rule():
    want_logits = [_logits(top=(10.0, 9.0)), _logits(top=(10.0, 9.98), seed=1)]
    want_ids = [3, 3]
    assert par.compare_greedy(want_ids, want_logits, [3, 3], want_logits[0])["ok"]
    # second step: the oracle's two best are 0.02 apart on a range of ~13: choosing the runner-up is a legitimate argmax
    r = par.compare_greedy(want_ids, want_logits, [3, 7])
    assert r["first_divergence"] == 1 and r["ok"] and r["oracle_gap_of_gpu_choice"] < 2 * par.LOGIT_TOL
    # first step: the runner-up is a full logit behind (~8 % of the range): not excusable
    r = par.compare_greedy(want_ids, want_logits, [7, 3])
    assert r["first_divergence"] == 0 and not r["ok"]
    # a logit row that is off by more than the tolerance fails even with identical ids
    off = want_logits[0].copy(); off[20] += 0.1 * (want_logits[0].max() - want_logits[0].min())
    assert not par.compare_greedy(want_ids, want_logits, [3, 3], off)["ok"]


def test_a_measured_floor_widens_the_bounds_and_never_narrows_them():
    assert par.tolerances(None) == (par.LOGIT_TOL, par.KL_TOL)
    assert par.tolerances({"err": 0.001, "kl": 1e-6}) == (par.LOGIT_TOL, par.KL_TOL)
    e, k = par.tolerances({"err": 0.045, "kl": 0.007})
    assert abs(e - 0.135) < 1e-9 and abs(k - 0.035) < 1e-9


def test_one_ulp_perturbation_moves_the_tiny_models_logits_a_little_and_switching_it_off_restores_them(tiny_models):
    path = tiny_models["tiny-llama-q4_k_m"]
    prompt = np.random.default_rng(2).integers(3, 1000, size=12).tolist()
    ref = rm.RefGGUFModel(path, n_ctx=64)
    ids, logits = par.oracle_greedy_trace(ref, prompt, 3)
    floor = par.sensitivity_floor(ref, rm.lib(), prompt, ids, logits, n_pos=2)
    assert 0.0 < floor["err"] < 0.01 and floor["kl"] < 1e-4           # 3 layers: far below the 3 % tolerance
    again = par.sensitivity_floor(ref, rm.lib(), prompt, ids, logits, n_pos=2)
    assert again == floor                                             # the perturbation is seeded
    ids2, logits2 = par.oracle_greedy_trace(ref, prompt, 3)           # and it is off again afterwards
    assert ids2 == ids and all(np.array_equal(a, b) for a, b in zip(logits, logits2))
    ref.close()
