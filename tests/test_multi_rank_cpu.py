"""The N>1 path on CPU: two gloo ranks exercise the row plan, the handle exchange and the timing reduction;
bench.py's reference arm prints from rank 0 only."""
import json
import os
import subprocess
import sys

from kubernetes_b200 import sharding

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _torchrun(args, port, timeout=300):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
           "127.0.0.1", "--master-port", str(port)] + args
    return subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=timeout)


def test_plan_matches_engine_rules():
    assert sharding.plan(50000, 8) == [(i * 6400, 6400) for i in range(7)] + [(44800, 5200)]
    assert sharding.plan(1200, 2) == [(0, 768), (768, 432)]
    assert sharding.plan(100, 2) == [(0, 100), (100, 0)]
    assert sharding.plan(5000, 1) == [(0, 5000)]


def test_two_gloo_ranks_agree_on_plan_and_exchange():
    r = _torchrun([os.path.join("tests", "_dist_worker.py"), "1300"], 29631)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "rank 0 ok" in r.stdout and "rank 1 ok" in r.stdout


def test_reference_arm_prints_on_rank0_only():
    r = _torchrun(["bench.py", "--impl", "reference", "--gpus", "2", "--steps", "1", "--warmup", "0", "--config", "1",
                   "--cpu-sample", "200"], 29632)
    assert r.returncode == 0, r.stdout + r.stderr
    lines = [l for l in r.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1, r.stdout
    out = json.loads(lines[0])
    assert out["impl"] == "reference" and out["cpu_baseline"]["kind"] == "port" and out["value"] > 0
    assert out["e2e"]["h2d_bytes_per_step"] == 0
