"""Tensor-parallel decode on N GPUs of one box: parity of the sharded model against the oracle and
against a full replica, then decode tok/s of the Llama-3.1-8B shape at this TP degree.

  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tools/tp_check.py

One process per GPU; torch.distributed (gloo) only carries the 128-byte NCCL id and the barriers."""
import importlib, json, os, sys, time, traceback
import numpy as np
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
b2 = importlib.import_module("llms-on-kubernetes_b200")
sg = importlib.import_module("llms-on-kubernetes_b200.synth_gguf")

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ.get("LOCAL_RANK", 0))
dist.init_process_group("gloo")


def group_id():
    box = [b2.allreduce_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(box, src=0)
    return box[0]


def shard_context(path, n_ctx):
    m = b2.Model(path, local, rank, world)
    ctx = b2.Context(m, n_ctx)
    ctx.allreduce_init(group_id())
    return m, ctx


def main():
    tmp = os.environ.get("B2_BENCH_DIR", "/tmp")
    ok = True
    for name in ("tiny-llama-q4_k_m", "tiny-llama3-q4_k_m", "tiny-qwen2-q8_0"):
        if os.environ.get("B2_TP_BENCH_ONLY"):
            break
        spec = sg.SPECS[name]
        if spec.n_kv % world or spec.n_head % world:
            continue
        path = os.path.join(tmp, f"b2_tp_{name}.gguf")
        if rank == 0:
            sg.write_model(path, name, seed=3)
        dist.barrier()
        m, ctx = shard_context(path, 128)
        toks = np.random.default_rng(4).integers(3, spec.vocab, size=40).tolist()
        nxt, logits = ctx.decode(toks, 0)
        ids = ctx.generate_greedy(12)
        gathered = [None] * world
        dist.all_gather_object(gathered, (nxt, ids))
        if rank == 0:
            from oracle import ref_model as rm
            ref = rm.RefGGUFModel(path, n_ctx=128)
            for p, t in enumerate(toks):
                tok, lr = ref.decode(t, p, want_logits=p == len(toks) - 1)
            err = float(np.abs(logits - lr).max() / np.abs(lr).max())
            full = b2.Model(path, local)
            fctx = b2.Context(full, 128)
            fn, fl = None, None
            for p, t in enumerate(toks):
                fn, fl = fctx.decode([t], p)
            err_full = float(np.abs(logits - fl).max() / np.abs(fl).max())
            same = all(g == gathered[0] for g in gathered)
            print(json.dumps({"model": name, "tp": world, "logits_err_vs_oracle": err, "logits_err_vs_replica": err_full,
                              "ranks_agree": same, "first_token_matches_replica": nxt == fn}), flush=True)
            ok = ok and err <= 0.03 and err_full <= 0.03 and same
            fctx.free(); full.free(); ref.close()
        ctx.free(); m.free()
        dist.barrier()

    if os.environ.get("B2_TP_PARITY_ONLY"):
        return ok
    # throughput: the Llama-3.1-8B shape at this TP degree (the 70B file is 42 GB of synthetic blocks: not generated here)
    name = os.environ.get("B2_TP_MODEL", "llama-3.1-8b-q4_k_m")
    path = os.path.join(tmp, f"b2_synth_{name}_s0.gguf")
    if rank == 0 and not os.path.exists(path):
        sg.write_model(path + ".tmp", name, seed=0); os.replace(path + ".tmp", path)
    dist.barrier()
    m, ctx = shard_context(path, 512)
    prompt = np.random.default_rng(0).integers(3, sg.SPECS[name].vocab, size=16).tolist()
    ctx.decode(prompt, 0, want_logits=False)
    ctx.generate_greedy(16)
    ctx.sync(); dist.barrier()
    t0 = time.perf_counter()
    steps = int(os.environ.get("B2_TP_STEPS", "128"))
    ctx.generate_greedy(steps)
    ctx.sync()
    dt = time.perf_counter() - t0
    if rank == 0:
        print(json.dumps({"model": name, "tp": world, "decode_tok_per_s": steps / dt, "ms_per_token": dt / steps * 1e3,
                          "shard_stream_bytes": int(m.info.stream_bytes), "launches_per_step": ctx.launches_per_step,
                          "graph": not os.environ.get("B2_TP_NO_GRAPH"), "parity_ok": ok}), flush=True)
    ctx.free(); m.free()
    dist.barrier()
    return ok


try:
    good = main()
except Exception:
    traceback.print_exc()
    print(f"rank {rank}: FAILED", flush=True)
    good = False
dist.destroy_process_group()
sys.exit(0 if good else 1)
