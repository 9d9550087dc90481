"""Parity on the BASELINE.json configurations at their REAL shapes (synthetic GGUF files): the models the
chart serves (ramalama-models/helm-chart/values.yaml:4,12) are TinyLlama-1.1B-Chat Q4_K_M, and the GPU configurations
Qwen2.5-1.5B-Instruct Q8_0 and Llama-3.1-8B-Instruct Q4_K_M.  The tiny models of the other tests have 2-3 layers,
d = 512 and a 1 K vocabulary; here the persistent kernel runs its full grid with several tiles per CTA, K = 14336
Q6_K down projections, the 128 256-row Q6_K lm_head, GQA groups of 4 and 7, llama-3 rope factors at head_dim 128.

Per model and prompt (128 ids, seeded), against oracle/ggml_ref.c:
  (i)   the persistent decode kernel, token by token over the prompt, then 16 greedy ids: logits max-abs / KL at the
        last prompt position, ids identical up to the first near-tie of the oracle (index and margin reported);
  (ii)  b2_ctx_prefill (tcgen05 prompt pass): logits at the last prompt position;
  (iii) b2_ctx_decode_batch at 8 columns (int8 mma.sync pass) and 32 columns (tcgen05 pass): every column's logits
        for the second generated position.
The bound is max(3 % of the logit range, 3 x the drift of the ORACLE ITSELF under a one-ulp perturbation of its residual
stream) -- measured per model in the test (oracle/parity.py: sensitivity_floor): at 22-32 layers of random weights a
last-bit change already moves the logits by percents of their range, so no two correct implementations agree better.
The report goes to gpurun_out/r2_parity_configs.json when that directory exists.
"""
import json
import os

import numpy as np
import pytest

from oracle import parity as par
from oracle import ref_model as rm

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CONFIGS = [("tinyllama-1.1b-q4_k_m", 4), ("qwen2.5-1.5b-q8_0", 4), ("llama-3.1-8b-q4_k_m", 2)]   # (spec, prompts)
N_PROMPT, N_NEW = 128, 16
REPORT = {}


def model_file(synth, name):
    d = os.environ.get("B2_BENCH_DIR", "/tmp")
    path = os.path.join(d, f"b2_synth_{name}_s0.gguf")       # the file bench.py uses (seed 0)
    if not os.path.exists(path):
        synth.write_model(path + ".tmp", name, seed=0)
        os.replace(path + ".tmp", path)
    return path


def save_report():
    out = os.path.join(ROOT, "gpurun_out")
    if os.path.isdir(out):
        with open(os.path.join(out, "r2_parity_configs.json"), "w") as f:
            json.dump(REPORT, f, indent=1)


@pytest.mark.parametrize("name,n_prompts", CONFIGS)
def test_real_shape_parity(b2, synth, name, n_prompts, capsys):
    path = model_file(synth, name)
    rm.lib().ref_set_num_threads(len(os.sched_getaffinity(0)))
    ref = rm.RefGGUFModel(path, n_ctx=N_PROMPT + N_NEW + 8)
    m = b2.Model(path)
    V = m.info.vocab
    prompts, oracle = [], []
    for seed in range(n_prompts):
        p = np.random.default_rng(seed).integers(3, V, size=N_PROMPT).tolist()
        prompts.append(p)
        oracle.append(par.oracle_greedy_trace(ref, p, N_NEW))
    # the noise floor of this model at this depth and prompt length: the oracle against itself, one fp32 ulp apart
    floor = par.sensitivity_floor(ref, rm.lib(), prompts[0], oracle[0][0], oracle[0][1])
    tol_err, tol_kl = par.tolerances(floor)
    ref.close()
    rep = REPORT.setdefault(name, {})
    rep["oracle_one_ulp_drift"] = floor
    rep["tolerance"] = {"logits_err_over_range": tol_err, "kl": tol_kl}

    # (i) the persistent decode kernel
    ctx = b2.Context(m, N_PROMPT + N_NEW + 8)
    rows = []
    for p, (want_ids, want_logits) in zip(prompts, oracle):
        for pos, t in enumerate(p):
            tok, lg = ctx.decode([t], pos, want_logits=(pos == N_PROMPT - 1))
        got = [tok] + ctx.generate_greedy(N_NEW - 1)
        res = par.compare_greedy(want_ids, want_logits, got, lg, floor)
        rows.append(res)
        assert res["logits_err_over_range"] <= tol_err and res["kl"] <= tol_kl, (name, "decode kernel", res)
        assert res["ok"], f"{name}: greedy ids leave the oracle's at a step that is not a near-tie: {res}"
    assert ctx.launches_per_step == 1, "the persistent kernel did not run (one launch per token expected)"
    rep["decode_kernel"] = rows
    ctx.free()

    # (ii) the tcgen05 prompt pass, (iii) batched decode at 8 and 32 columns
    for cols in (8, 32):
        ctx = b2.Context(m, N_PROMPT + 8, n_slots=cols)
        pre = []
        for j in range(cols):
            nxt, lg = ctx.prefill(prompts[j % n_prompts], slot=j, want_logits=(j < n_prompts))
            if j < n_prompts:
                st = par.logit_stats(oracle[j][1][0], lg)
                pre.append({"err": st["err"], "kl": st["kl"], "first_id_equal": int(nxt) == oracle[j][0][0], "oracle_margin": st["margin"]})
                assert st["err"] <= tol_err and st["kl"] <= tol_kl, (name, "prefill", st)
                assert int(nxt) == oracle[j][0][0] or st["margin"] <= 2 * tol_err, (name, "prefill first id", st)
        if cols == 8:
            rep["prefill_tcgen05"] = pre
        # every sequence advances by the oracle's first generated id; its logits are the oracle's second-step logits
        toks = [oracle[j % n_prompts][0][0] for j in range(cols)]
        ids, lg = ctx.decode_batch(toks, [N_PROMPT] * cols, slots=list(range(cols)))
        bat = []
        for j in range(cols):
            st = par.logit_stats(oracle[j % n_prompts][1][1], lg[j])
            bat.append({"err": st["err"], "kl": st["kl"]})
            assert st["err"] <= tol_err and st["kl"] <= tol_kl, (name, f"batch {cols}", j, st)
        rep[f"decode_batch_{cols}"] = {"max_err": max(b["err"] for b in bat), "max_kl": max(b["kl"] for b in bat)}
        if cols == 32 and name == "llama-3.1-8b-q4_k_m":
            # (iv) the same 32 columns through the pass whose tcgen05 kernels dequantise the raw blocks themselves
            # (tcq_gemm.cu; B2_TCQ=1 prefers it although the fp16 copy is resident): same operand values, so the logits
            # differ from the fp16-copy pass by summation order only
            os.environ["B2_TCQ"] = "1"
            try:
                ids_q, lg_q = ctx.decode_batch(toks, [N_PROMPT] * cols, slots=list(range(cols)))
            finally:
                del os.environ["B2_TCQ"]
            worst = 0.0
            for j in range(cols):
                st = par.logit_stats(oracle[j % n_prompts][1][1], lg_q[j])
                assert st["err"] <= tol_err and st["kl"] <= tol_kl, (name, "fused-dequantisation batch", j, st)
                rng_j = float(lg[j].max() - lg[j].min())
                worst = max(worst, float(np.abs(lg_q[j] - lg[j]).max()) / rng_j)
            rep["decode_batch_32_fused_dequant"] = {"max_diff_vs_fp16_copy_pass_over_range": worst}
            assert worst <= tol_err
        ctx.free()
    m.free()
    save_report()
    with capsys.disabled():
        d = rep["decode_kernel"]
        print(f"\n  {name}: oracle one-ulp drift {floor['err']:.4f} / kl {floor['kl']:.2e} -> tolerance {tol_err:.4f} / {tol_kl:.2e}")
        print(f"  {name}: decode kernel err {max(r['logits_err_over_range'] for r in d):.4f} kl {max(r['kl'] for r in d):.2e} "
              f"divergences {[(r['first_divergence'], r['margin_at_divergence']) for r in d]}; "
              f"batch 8 {rep['decode_batch_8']}, batch 32 {rep['decode_batch_32']}")
