"""Tensor parallelism: the model sharded over two GPUs of one box -- Q/K/V by heads and gate/up by
ff rows, O and down along K -- decodes the same logits as the oracle and as a full replica, and every rank samples
the same tokens.  One process per GPU under torch.distributed.run; skipped on a box with a single GPU."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


# every form of the exchange that was built: the default (reduce stage, partial rows pulled over NVLink),
# rows pushed into every rank's inbox, the consumer adding the rows up itself, and round 1's NCCL all-reduce path
# (captured in the step's CUDA graph, and eager)
FORMS = [{}, {"B2_TP_PUSH": "1"}, {"B2_TP_GATHER": "1"}, {"B2_TP_NCCL": "1"}, {"B2_TP_NCCL": "1", "B2_TP_NO_GRAPH": "1"}]


@pytest.mark.parametrize("form", FORMS, ids=lambda f: "+".join(sorted(f)) or "default")
def test_two_way_tensor_parallel_decode_matches_oracle_and_replica(b2, tmp_path, form):
    if b2.device_count() < 2:
        pytest.skip("needs two GPUs")
    env = dict(os.environ, B2_TP_PARITY_ONLY="1", B2_BENCH_DIR=str(tmp_path), **form)
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
                        "--master-port", "29517", os.path.join(ROOT, "tools", "tp_check.py")], env=env, capture_output=True, text=True,
                       timeout=600)
    lines = [ln for ln in r.stdout.splitlines() if ln.startswith("{")]
    assert r.returncode == 0 and len(lines) == 3, r.stdout[-2000:] + r.stderr[-2000:]


def test_shard_arguments_are_checked(b2, tiny_models):
    with pytest.raises(b2.B2Error):
        b2.Model(tiny_models["tiny-llama-q4_k_m"], 0, 2, 2)          # rank outside the group
    with pytest.raises(b2.B2Error):
        b2.Model(tiny_models["tiny-llama-q4_k_m"], 0, 0, 3)          # 8 heads / 2 kv heads do not divide over 3 ranks
    m = b2.Model(tiny_models["tiny-llama-q4_k_m"], 0, 1, 2)          # a shard loads on one GPU; it needs the group to decode
    full = b2.Model(tiny_models["tiny-llama-q4_k_m"])
    assert m.info.n_head * 2 == full.info.n_head and m.info.n_kv * 2 == full.info.n_kv and m.info.d_ff * 2 == full.info.d_ff
    assert m.info.vocab == full.info.vocab and m.info.stream_bytes < full.info.stream_bytes
    ctx = b2.Context(m, 64)
    with pytest.raises(b2.B2Error):
        ctx.decode([1, 2, 3], 0)                                      # no communicator yet
    fctx = b2.Context(full, 64)
    with pytest.raises(b2.B2Error):
        fctx.allreduce_init(b"\0" * 128)                              # a replica has no group
    fctx.free(); ctx.free(); m.free(); full.free()
