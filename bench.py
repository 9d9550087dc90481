#!/usr/bin/env python
"""bench.py -- BASELINE.json's metric: batch-1 greedy decode tok/s of Llama-3.1-8B Q4_K_M (synthetic GGUF at
the real shape) and the fraction of the B200 HBM roofline it reaches.

  python bench.py --gpus N --steps K --warmup W            # our arm  (torchrun for N > 1: one replica per GPU)
  python bench.py --impl reference --steps K --warmup W    # CPU arm: the oracle's restatement of the ggml CPU path

A "step" is one decode step = one generated token (per replica).  `value` is device-resident greedy
generation (token fed back on the GPU, CUDA-event timed); `e2e` is the same loop through b2_ctx_decode with
host buffers: every step copies the token id + position host->device from pinned memory and reads the sampled
id back.  N > 1 = independent replicas (the path shards by request; no collective) -> "scaling": "weak".
"""
from __future__ import annotations

import argparse
import importlib
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
PKG = "llms-on-kubernetes_b200"
METRIC = "decode tok/s + HBM GB/s % of peak, Llama-3.1-8B Q4_K_M, 1/2/4/8 B200"


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


def tensor_peak():
    """Dense bf16 TFLOP/s: the sustained figure (the prompt pass is one long back-to-back run of mat-muls)."""
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            j = json.load(open(p))
            return float(j.get("bf16_tflops_sustained") or j["bf16_tflops"]), "measured, sustained (MEASURED_PEAKS.json)"
        except Exception:
            pass
    return 2250.0, "fallback (nominal dense bf16, B200_PROFILING.md)"


def ncu_traffic(model: str):
    """DRAM bytes one step_kernel launch moved, from the committed ncu capture (None if there is none for this model)."""
    try:
        return float(json.load(open(os.path.join(ROOT, "profiles", "step_traffic.json")))[model]["dram_bytes_per_launch"])
    except Exception:
        return None


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while the timed region runs."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.gpu, self.lines, self.proc = gpu_index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "200"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))
            except ValueError:
                continue
            for n, v in zip(names, f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def load_synth_without_package():
    """synth_gguf.py by file path: importing the package would map libb2.so into the CPU arm's process."""
    import importlib.util
    spec = importlib.util.spec_from_file_location("b2_synth_gguf", os.path.join(ROOT, PKG, "synth_gguf.py"))
    mod = importlib.util.module_from_spec(spec)
    sys.modules["b2_synth_gguf"] = mod
    spec.loader.exec_module(mod)
    return mod


def host_threads() -> int:
    """Threads the CPU arm uses: every core this process may run on (B2_REF_THREADS overrides).  torch.distributed.run
    exports OMP_NUM_THREADS=1 to its workers, so the inherited value is not used."""
    if os.environ.get("B2_REF_THREADS"):
        return max(1, int(os.environ["B2_REF_THREADS"]))
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return max(1, os.cpu_count() or 1)


def bench_config(workload: str, world: int, tp: bool) -> dict:
    """The `config` object of the JSON line: identical for the B200 arm and the CPU arm of the same command line."""
    return {"workload": workload,
            "parallelism": (f"tensor parallel x{world}: Q/K/V by heads, gate/up by rows, O/down along K, all-reduce per half-layer; "
                            "roofline.achieved is ONE rank's weight stream") if tp
            else f"replicas x{world}" if world > 1 else "single GPU",
            "l2": "inputs exceed L2 (4.6 GB of weights streamed per step vs 126 MB)", "graph": True}


def find_llama_bench():
    """SURVEY 8c tier T4: a real llama.cpp on the box ($PATH or baseline/_ref/) is the ground-truth CPU baseline."""
    import shutil
    for name in ("llama-bench",):
        p = shutil.which(name)
        if p:
            return p
    base = os.path.join(ROOT, "baseline", "_ref")
    for dirpath, _, files in os.walk(base):
        if "llama-bench" in files:
            return os.path.join(dirpath, "llama-bench")
    return None


def llama_bench_decode(exe: str, path: str, n_tokens: int, threads: int):
    """tok/s of `llama-bench -p 0 -n n_tokens -ngl 0` on the same GGUF, or None if the run or its output is unusable."""
    try:
        out = subprocess.run([exe, "-m", path, "-p", "0", "-n", str(n_tokens), "-t", str(threads), "-ngl", "0", "-o", "json"],
                             capture_output=True, text=True, timeout=240)
        rows = json.loads(out.stdout)
        for r in rows:
            if int(r.get("n_gen", 0)) > 0 and float(r.get("avg_ts", 0)) > 0:
                return float(r["avg_ts"])
    except Exception:
        return None
    return None


def ensure_model(synth, spec_name: str, seed: int, rank: int, barrier) -> str:
    d = os.environ.get("B2_BENCH_DIR", "/tmp")
    path = os.path.join(d, f"b2_synth_{spec_name}_s{seed}.gguf")
    if rank == 0 and not os.path.exists(path):
        tmp = path + ".tmp"
        synth.write_model(tmp, spec_name, seed=seed)
        os.replace(tmp, path)
    barrier()
    return path


def cpu_decode_sample(path: str, n_tokens: int, warm: int):
    """Oracle port (oracle/ggml_ref.c, all host threads): decode n_tokens from an empty KV cache."""
    from oracle import ref_model as rm
    rm.lib().ref_set_num_threads(host_threads())
    ref = rm.RefGGUFModel(path, n_ctx=max(64, n_tokens + warm + 1))
    tok, pos = 1, 0
    for _ in range(warm):
        tok, _ = ref.decode(tok, pos, want_logits=False); pos += 1
    t0 = time.perf_counter()
    for _ in range(n_tokens):
        tok, _ = ref.decode(tok, pos, want_logits=False); pos += 1
    dt = time.perf_counter() - t0
    ref.close()
    return n_tokens / dt, dt, rm.lib().ref_num_threads()


def parity_check(b2, model, path: str, seed: int, n_prompt: int = 16, n_new: int = 8) -> dict:
    """Outside the timed region: the timed path (one persistent-kernel launch per token, through b2_ctx_decode) against the
    oracle on a short prompt of the benchmarked model -- logits after the prompt and the greedy ids that follow."""
    from oracle import parity as par
    from oracle import ref_model as rm
    rm.lib().ref_set_num_threads(host_threads())
    rng = np.random.default_rng(seed + 1000)
    prompt = rng.integers(3, model.info.vocab, size=n_prompt).tolist()
    ref = rm.RefGGUFModel(path, n_ctx=n_prompt + n_new + 8)
    want_ids, want_logits = par.oracle_greedy_trace(ref, prompt, n_new)
    floor = par.sensitivity_floor(ref, rm.lib(), prompt, want_ids, want_logits)   # the oracle against itself, one ulp apart
    ref.close()
    ctx = b2.Context(model, max(64, n_prompt + n_new + 8))
    tok, logits = None, None
    for pos, t in enumerate(prompt):          # token by token: every position runs the decode kernel that is timed below
        tok, logits = ctx.decode([t], pos, want_logits=(pos == n_prompt - 1))
    got = [tok] + (ctx.generate_greedy(n_new - 1) if n_new > 1 else [])
    ctx.free()
    res = par.compare_greedy(want_ids, want_logits, got, logits, floor)
    res.update({"prompt_tokens": n_prompt, "new_tokens": n_new, "oracle": "oracle/ggml_ref.c",
                "tolerance": "max(3 % of the logit range, 3 x the oracle's own drift under a one-ulp perturbation); KL: max(1e-3, 5 x)"})
    return res


def bench_batched(b2, synth, spec, model, args, world, barrier, max_over_ranks, gpu_index):
    """The other decode / prefill configurations of SURVEY 8d (not the headline): `--batch B` = one token of each of B
    sequences per step through b2_ctx_decode_batch, `--prefill T` = a T-token prompt through b2_ctx_decode.  Both go
    through the C ABI with host buffers (token ids in, sampled ids out, every step), so value == e2e."""
    info = model.info
    rng = np.random.default_rng(args.seed)
    peak, peak_src = peaks()
    if args.prefill > 0:
        T = args.prefill
        ctx = b2.Context(model, T + 8)
        prompt = rng.integers(3, info.vocab, size=T).tolist()
        for _ in range(max(1, min(args.warmup, 3))):
            ctx.decode(prompt, 0, want_logits=False)
        reps = max(1, min(args.steps, 8))
        barrier(); ctx.sync()
        sampler = ClockSampler(gpu_index); sampler.start()
        t0 = time.perf_counter()
        for _ in range(reps):
            ctx.decode(prompt, 0, want_logits=False)
        ctx.sync()
        dt = max_over_ranks(time.perf_counter() - t0)
        clocks = sampler.stop()
        ctx.free()
        value = world * reps * T / dt
        # mat-mul flops of the prompt (2 x parameters in the layer matrices x T) + attention (4 x head_dim x n_head x T^2 / 2)
        flops = 2.0 * synth.matmul_params(spec) * T + 2.0 * spec.n_layer * spec.n_head * spec.head_dim * float(T) * T
        tf_peak, tf_src = tensor_peak()
        tc = T >= 24   # b2_ctx_decode takes the tensor-core pass from 24 rows when the fp16 copy is resident
        return {"metric": "prefill tok/s", "value": value, "unit": "tok/s", "n_gpus": world, "steps": reps, "warmup": min(args.warmup, 3),
                "ms_per_step": dt / reps * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "bf16 x bf16 tcgen05.mma, fp32 accumulate (weights dequantised to bf16 tiles)" if tc else
                         "int8 x int4..6 mma.sync (u8.s8 -> s32), fp32 scales", "data": "synthetic",
                "config": {"workload": f"{spec.name} {spec.ftype} synthetic GGUF, prefill of {T} tokens" +
                                       (" in chunks of up to 2048 tokens on the tcgen05 path" if tc else " in 32-token chunks"),
                           "parallelism": f"replicas x{world}" if world > 1 else "single GPU"},
                "e2e": {"value": value, "unit": "tok/s", "h2d_bytes_per_step": 4 * T, "d2h_bytes_per_step": 4},
                "gpu_launches": None, "clocks": clocks,
                "roofline": {"bound": "tensor", "achieved": flops * reps / dt / 1e12, "peak": tf_peak, "unit": "TFLOP/s",
                             "frac": flops * reps / dt / 1e12 / tf_peak, "traffic": None, "peak_source": tf_src,
                             "flops_per_prompt": flops,
                             "note": "whole prompt pass: dequantisation to bf16, mat-muls, attention, elementwise kernels and the last position's lm_head"}}
    B = args.batch
    n_ctx = args.prompt + args.steps + args.warmup + 8
    ctx = b2.Context(model, n_ctx, n_slots=B)
    toks = []
    for s in range(B):
        prompt = rng.integers(3, info.vocab, size=args.prompt).tolist()
        for p0 in range(0, args.prompt, 32):
            chunk = prompt[p0:p0 + 32]
            ids, _ = ctx.decode_batch(chunk, list(range(p0, p0 + len(chunk))), slots=[s] * len(chunk), last_only=True, want_logits=False)
        toks.append(ids[0])
    pos = args.prompt
    slots = list(range(B))
    for _ in range(args.warmup):
        toks, _ = ctx.decode_batch(toks, [pos] * B, slots=slots, want_logits=False); pos += 1
    barrier(); ctx.sync()
    sampler = ClockSampler(gpu_index); sampler.start()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        toks, _ = ctx.decode_batch(toks, [pos] * B, slots=slots, want_logits=False); pos += 1
    ctx.sync()
    dt = max_over_ranks(time.perf_counter() - t0)
    clocks = sampler.stop()
    ctx.free()
    value = world * B * args.steps / dt
    step_bytes = info.stream_bytes + B * (args.prompt + args.warmup + args.steps // 2) * synth.kv_bytes_per_token(spec)
    gbs = step_bytes * args.steps / dt / 1e9
    tc = B >= 24   # from 24 columns b2_ctx_decode_batch takes the tensor-core pass when the fp16 copy is resident (tc_min_rows)
    return {"metric": f"decode tok/s at batch {B}", "value": value, "unit": "tok/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "fp16 x fp16 tcgen05.mma, fp32 accumulate (weights dequantised to fp16 tiles, activations quantised as the reference does)" if tc
                     else "int8 x int4..6 mma.sync (u8.s8 -> s32), fp32 scales", "data": "synthetic",
            "config": {"workload": f"{spec.name} {spec.ftype} synthetic GGUF, batch-{B} greedy decode at ctx {args.prompt}",
                       "parallelism": f"replicas x{world}" if world > 1 else "single GPU",
                       "l2": "inputs exceed L2 (4.6 GB of weights streamed per step vs 126 MB)"},
            "e2e": {"value": value, "unit": "tok/s", "h2d_bytes_per_step": 12 * B, "d2h_bytes_per_step": 4 * B},
            "gpu_launches": None, "clocks": clocks,
            "roofline": {"bound": "hbm", "achieved": gbs, "peak": peak, "unit": "GB/s", "frac": gbs / peak, "traffic": None,
                         "peak_source": peak_src, "bytes_per_step": step_bytes}}


_real_stdout = [None]


def emit(obj: dict):
    """The one JSON line of the run, on the process's original stdout."""
    line = json.dumps(obj)
    if _real_stdout[0] is not None:
        _real_stdout[0].write(line + "\n")
        _real_stdout[0].flush()
    else:
        print(line, flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=256)
    ap.add_argument("--warmup", type=int, default=16)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--model", default="llama-3.1-8b-q4_k_m")
    ap.add_argument("--prompt", type=int, default=128, help="tokens in the KV cache when the timed decode starts")
    ap.add_argument("--seed", type=int, default=0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-parity", action="store_true", help="skip the oracle check of the timed path (16-token prompt, 8 greedy ids) before timing")
    ap.add_argument("--dist-selftest", action="store_true", help="CPU-only check of the multi-rank aggregation (gloo)")
    ap.add_argument("--batch", type=int, default=1, help="> 1: batched decode of this many sequences (SURVEY 8d config 3: 8, 32)")
    ap.add_argument("--prefill", type=int, default=0, help="> 0: time the prefill of a prompt of this many tokens instead")
    ap.add_argument("--tp", action="store_true", help="under torchrun: ONE model sharded over all ranks (tensor parallel, SURVEY 8e "
                                                      "config 5, e.g. --model llama-3.1-70b-q4_k_m) instead of one replica per rank")
    args = ap.parse_args()
    if args.warmup < 3:
        args.warmup = 3
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    # the CPU arm must not map the product library: synth_gguf.py is loaded by path there
    synth = load_synth_without_package() if args.impl == "reference" else importlib.import_module(PKG + ".synth_gguf")
    spec = synth.SPECS[args.model]
    workload = f"{spec.name} {spec.ftype} synthetic GGUF, batch-1 greedy decode at ctx {args.prompt}"

    # ------------------------------------------------------------------ reference (CPU) arm
    if args.impl == "reference":
        if rank != 0:
            return
        os.environ["OMP_NUM_THREADS"] = str(host_threads())   # before the OpenMP runtime starts (the oracle library loads lazily)
        path = ensure_model(synth, args.model, args.seed, 0, lambda: None)
        tp = (args.tp or "70b" in args.model) and world > 1
        # bounded: one CPU token of the 8B model is ~0.1-1 s; cap the run near two minutes
        tps1, dt1, cores = cpu_decode_sample(path, 1, 1)
        budget = 120.0
        steps = max(1, min(args.steps, int(budget / max(dt1, 1e-3)) - args.warmup))
        warm = min(args.warmup, max(1, int(10.0 / max(dt1, 1e-3))))
        tps, dt, cores = cpu_decode_sample(path, steps, warm)
        kind = "port"
        sample = (f"{steps} decode steps from an empty KV cache after {warm} warm-up steps (KV read is <1% of step bytes), "
                  f"oracle/ggml_ref.c restatement of the ggml CPU path (llama.cpp itself is not on this machine), OpenMP {cores} threads")
        exe = find_llama_bench()
        if exe:
            real = llama_bench_decode(exe, path, min(steps, 64), host_threads())
            if real:
                tps, kind = real, "llama.cpp"
                sample = f"{exe} -p 0 -n {min(steps, 64)} -t {host_threads()} -ngl 0 on the same GGUF"
        line = {
            "impl": "reference", "metric": METRIC, "value": tps, "unit": "tok/s", "n_gpus": args.gpus, "steps": steps,
            "warmup": warm, "ms_per_step": 1000.0 / tps, "higher_is_better": True, "scaling": "strong" if tp else "weak", "vs_baseline": None,
            "dtype": "int8 x int4..8 dot, fp32 accumulate", "data": "synthetic",
            "config": bench_config(workload, world, tp),
            "cpu_baseline": {"value": tps, "unit": "tok/s", "cores": cores, "kind": kind, "sample": sample},
            "e2e": {"value": tps, "unit": "tok/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        }
        if cores <= 1 and host_threads() > 1:
            line["note"] = "the OpenMP team is one thread although more cores are available: do not form a ratio from this line"
        print(json.dumps(line))
        return

    # ------------------------------------------------------------------ B200 arm
    # Replicas only: the path shards by request, every rank serves its own copy of the model on its own GPU and no
    # collective touches the data path.  torch.distributed carries the barrier and the max-over-ranks of the timings.
    dist = None
    red_device = "cuda"
    if world > 1:
        import torch
        import torch.distributed as dist
        if args.dist_selftest:
            red_device = "cpu"
            dist.init_process_group("gloo")
        else:
            # NCCL's log (communicator ranks, transports) is wanted, but it goes to stdout, which must carry exactly one
            # JSON line: file descriptor 1 is pointed at stderr for the rest of the process and the line is written to a
            # duplicate of the original stdout (emit() below)
            if os.environ.get("NCCL_DEBUG", "").upper() not in ("INFO", "TRACE"):
                os.environ["NCCL_DEBUG"] = "INFO"      # (a WARN / VERSION left in the environment would hide the communicator lines)
            os.environ.setdefault("NCCL_DEBUG_SUBSYS", "INIT")
            os.environ.pop("NCCL_DEBUG_FILE", None)
            sys.stdout.flush()
            _real_stdout[0] = os.fdopen(os.dup(1), "w")
            os.dup2(2, 1)
            torch.cuda.set_device(local_rank)
            dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    def barrier():
        if dist is not None:
            dist.barrier()

    def max_over_ranks(v: float) -> float:
        if dist is None:
            return v
        import torch
        t = torch.tensor([v], dtype=torch.float64, device=red_device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    if args.dist_selftest:
        # CPU check of the multi-rank plumbing (tests/test_multi_rank.py, gloo): every rank "measures" a different time;
        # the job's throughput is all ranks' steps over the slowest rank's time
        barrier()
        ms = max_over_ranks(10.0 * (rank + 1))
        if rank == 0:
            emit({"selftest": True, "n_gpus": world, "ms": ms, "value": world * args.steps / (ms / 1e3), "scaling": "weak"})
        if dist is not None:
            dist.destroy_process_group()
        return

    b2 = importlib.import_module(PKG)  # raises if libb2.so is missing: no fallback
    path = ensure_model(synth, args.model, args.seed, rank, barrier)
    tp = (args.tp or "70b" in args.model) and world > 1   # BASELINE config 5: the 70B model is tensor parallel over the ranks
    model = b2.Model(path, device=local_rank, tp_rank=rank if tp else 0, tp_size=world if tp else 1)
    jobs = 1 if tp else world          # sequences decoded at once over the whole job
    if tp and (args.batch > 1 or args.prefill > 0):
        raise SystemExit("--tp measures batch-1 decode (a sharded context decodes one sequence)")
    if args.batch > 1 or args.prefill > 0:
        out = bench_batched(b2, synth, spec, model, args, world, barrier, max_over_ranks, local_rank)
        model.free()
        if rank == 0:
            emit(out)
        if dist is not None:
            dist.destroy_process_group()
        return
    parity = None
    if rank == 0 and world == 1 and not args.no_parity:
        parity = parity_check(b2, model, path, args.seed)
        if not parity["ok"]:
            raise SystemExit("bench.py: the decode path does not match the oracle: " + json.dumps(parity))
    n_ctx = max(512, args.prompt + 2 * (args.steps + args.warmup) + 64)
    ctx = b2.Context(model, n_ctx)
    if tp:
        box = [b2.allreduce_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(box, src=0)
        ctx.allreduce_init(box[0])       # collective: NCCL communicator + the step's CUDA graph
    info = model.info
    rng = np.random.default_rng(args.seed)
    prompt = rng.integers(3, info.vocab, size=args.prompt).tolist()

    # ---- device-resident decode: `value`
    tok, _ = ctx.decode(prompt, 0, want_logits=False)
    ctx.time_steps(args.warmup)
    barrier(); ctx.sync()
    sampler = ClockSampler(local_rank)
    sampler.start()
    ms = ctx.time_steps(args.steps)
    ctx.sync(); barrier()
    clocks = sampler.stop()
    ms = max_over_ranks(ms)
    value = jobs * args.steps / (ms / 1e3)
    ctx_mid = args.prompt + args.warmup + args.steps // 2
    step_bytes = info.stream_bytes + ctx_mid * synth.kv_bytes_per_token(spec)

    # ---- end to end through the C ABI with host buffers: `e2e`
    pos = args.prompt + args.warmup + args.steps
    for _ in range(args.warmup):
        tok, _ = ctx.decode([tok], pos, want_logits=False); pos += 1
    barrier(); ctx.sync()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        tok, _ = ctx.decode([tok], pos, want_logits=False); pos += 1
    ctx.sync()
    e2e_s = max_over_ranks(time.perf_counter() - t0)
    barrier()
    e2e = jobs * args.steps / e2e_s

    # ---- roofline of the dominant kernel.  With the persistent per-token kernel (step.cu) one launch IS one decode
    # step, so the kernel's average duration is the CUDA-event time of the timed region / steps and its algorithmic
    # bytes are the step's (W_stream + ctx * kv_bytes_per_token).  The mat-vec stages alone, launched one
    # kernel per stage back to back (gemv_kernel, same device code), are timed beside it as an explanation.
    peak, peak_src = peaks()
    launches_per_step = ctx.launches_per_step
    step_gbs = step_bytes * args.steps / (ms / 1e3) / 1e9
    roofline = {"bound": "hbm",
                "kernel": "step_kernel (persistent decode step: all mat-vecs, attention, sampling in one launch)"
                          if launches_per_step == 1 else "one launch per stage (gemv_kernel dominant)",
                "achieved": step_gbs, "peak": peak, "unit": "GB/s", "frac": step_gbs / peak,
                "traffic": ncu_traffic(args.model) if launches_per_step == 1 else None,
                "traffic_source": "profiles/step_traffic.json (ncu --set full: dram__bytes_read.sum + dram__bytes_write.sum of one launch)",
                "peak_source": peak_src, "bytes_per_launch": step_bytes, "us_per_launch": ms * 1e3 / args.steps,
                "launches_per_step": launches_per_step}
    try:
        reps = max(3, min(20, args.steps // 8))
        mv_ms, mv_bytes, mv_launches = ctx.time_matvec_only(reps)
        mv_gbs = mv_bytes * reps / (mv_ms / 1e3) / 1e9
        roofline["matvec_stages_alone"] = {"achieved": mv_gbs, "frac": mv_gbs / peak, "launches": mv_launches,
                                           "bytes_per_launch": mv_bytes / mv_launches, "us_per_launch": mv_ms * 1e3 / reps / mv_launches}
    except Exception as e:
        roofline["matvec_stages_alone"] = {"error": str(e)}

    out = {
        "metric": METRIC if not tp else f"decode tok/s, {spec.name} {spec.ftype} tensor-parallel", "value": value, "unit": "tok/s",
        "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "strong" if tp else "weak", "vs_baseline": None,
        "dtype": "int8 x int4..8 dot (mma.sync u8.s8 / dp4a -> s32), fp32 accumulate", "data": "synthetic",
        "config": bench_config(workload, world, tp),
        "e2e": {"value": e2e, "unit": "tok/s", "h2d_bytes_per_step": 8, "d2h_bytes_per_step": 4},
        "gpu_launches": ctx.launches_per_step * args.steps,
        "clocks": clocks, "roofline": roofline,
        "parity_checked": bool(parity and parity["ok"]), "parity": parity,
    }
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        try:
            tps1, dt1, cores = cpu_decode_sample(path, 1, 1)
            n = max(2, min(64, int(15.0 / max(dt1, 1e-3))))
            tps, dt, cores = cpu_decode_sample(path, n, 1)
            out["cpu_baseline"] = {"value": tps, "unit": "tok/s", "cores": cores, "kind": "port",
                                   "sample": f"{n} decode steps from an empty KV cache, oracle/ggml_ref.c (llama.cpp-equivalent CPU algorithm), OpenMP"}
        except Exception as e:  # the baseline is reported, never required
            out["cpu_baseline"] = {"value": None, "unit": "tok/s", "cores": 0, "kind": "port", "sample": f"failed: {e}"}
    ctx.free()
    model.free()
    if rank == 0:
        emit(out)
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
