"""The N > 1 path of bench.py on CPU: world_size 2 over gloo.  The hot path shards by request (one replica per
GPU, no data-path collective -- DESIGN.md "replicas only"), so what there is to check is the launch contract:
RANK / WORLD_SIZE handling, barrier + max-over-ranks timing, one JSON line from rank 0, the reference arm
running on rank 0 alone."""
import json
import os
import socket
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def torchrun(n, *args, timeout=240):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}", "--master-addr", "127.0.0.1",
           "--master-port", str(free_port()), os.path.join(ROOT, "bench.py"), *args]
    return subprocess.run(cmd, capture_output=True, text=True, timeout=timeout, cwd=ROOT)


def test_two_ranks_aggregate_over_gloo():
    r = torchrun(2, "--gpus", "2", "--steps", "100", "--warmup", "3", "--dist-selftest")
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [ln for ln in r.stdout.splitlines() if ln.startswith("{")]
    assert len(lines) == 1, r.stdout          # rank 0 alone prints
    out = json.loads(lines[0])
    assert out["n_gpus"] == 2 and out["scaling"] == "weak"
    assert out["ms"] == 20.0                  # the slowest rank's time (ranks report 10 ms and 20 ms)
    assert abs(out["value"] - 2 * 100 / 0.020) < 1e-6   # all ranks' steps over that time


def test_single_process_selftest_needs_no_process_group():
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--steps", "10", "--dist-selftest"], capture_output=True, text=True,
                       timeout=120, cwd=ROOT)
    assert r.returncode == 0, r.stderr[-2000:]
    out = json.loads([ln for ln in r.stdout.splitlines() if ln.startswith("{")][0])
    assert out["n_gpus"] == 1 and out["ms"] == 10.0


def test_reference_arm_runs_on_rank_zero_only():
    """Under torchrun the CPU arm is rank 0's job; the other rank exits 0 without work or output."""
    env = dict(os.environ, RANK="1", LOCAL_RANK="1", WORLD_SIZE="2")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2", "--steps", "1", "--warmup", "1"],
                       capture_output=True, text=True, timeout=120, cwd=ROOT, env=env)
    assert r.returncode == 0 and r.stdout.strip() == "", (r.stdout, r.stderr[-500:])
