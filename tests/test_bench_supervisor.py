"""bench.py runs multi-GPU jobs in a supervised child per rank: a run that stops making progress is
re-run once on the NCCL collectives and says so in its JSON line. Exercised here with a test double
for the worker (gloo rendezvous, no GPU)."""
import json
import os
import subprocess
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent


def _run(mode, port):
    env = dict(os.environ, B200ZERO_BENCH_FAKE=mode, B200ZERO_BENCH_QUIET_S="3", B200ZERO_BENCH_STARTUP_S="60",
               OMP_NUM_THREADS="1")
    res = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                          "--master-addr", "127.0.0.1", "--master-port", str(port), str(ROOT / "bench.py"),
                          "--gpus", "2", "--steps", "2", "--warmup", "1"],
                         env=env, capture_output=True, text=True, timeout=300, cwd=str(ROOT))
    lines = [ln for ln in res.stdout.splitlines() if ln.startswith("{")]
    return res, lines


def test_supervised_run_passes_through_when_healthy():
    from dist_utils import free_port
    res, lines = _run("ok", free_port())
    assert res.returncode == 0, res.stderr[-2000:]
    assert len(lines) == 1
    out = json.loads(lines[0])
    assert out["comm"] == "auto" and "fallback" not in out


@pytest.mark.parametrize("mode", ["hang", "crash"])
def test_stalled_or_failed_run_falls_back_to_nccl_once(mode):
    from dist_utils import free_port
    res, lines = _run(mode, free_port())
    assert res.returncode == 0, res.stderr[-2000:]
    assert len(lines) == 1, res.stdout                      # exactly one JSON line reaches the caller
    out = json.loads(lines[0])
    assert out["comm"] == "nccl" and "fallback" in out
    assert "attempt 'primary'" in res.stderr and "attempt 'fused-retry'" in res.stderr


def test_second_attempt_of_the_same_path_is_reported_when_only_the_first_stalls():
    from dist_utils import free_port
    res, lines = _run("hang-once", free_port())
    assert res.returncode == 0, res.stderr[-2000:]
    assert len(lines) == 1, res.stdout
    out = json.loads(lines[0])
    assert out["comm"] == "auto" and "retried" in out and "fallback" not in out
    assert "attempt 'primary'" in res.stderr and "fused-retry" not in res.stderr


def test_attempts_are_cut_off_by_the_total_budget():
    """No stall detection fires (quiet window longer than the budget): the shares of the total budget end each
    attempt, and the last one still gets to report."""
    from dist_utils import free_port
    env = dict(os.environ, B200ZERO_BENCH_FAKE="hang", B200ZERO_BENCH_QUIET_S="600", B200ZERO_BENCH_STARTUP_S="600",
               B200ZERO_BENCH_BUDGET_S="30", OMP_NUM_THREADS="1")
    res = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                          "--master-addr", "127.0.0.1", "--master-port", str(free_port()), str(ROOT / "bench.py"),
                          "--gpus", "2", "--steps", "2", "--warmup", "1"],
                         env=env, capture_output=True, text=True, timeout=300, cwd=str(ROOT))
    assert res.returncode == 0, res.stderr[-2000:]
    assert res.stderr.count("over-budget") >= 2
    out = json.loads([ln for ln in res.stdout.splitlines() if ln.startswith("{")][0])
    assert out["comm"] == "nccl" and "fallback" in out


def test_stalled_probe_keeps_the_measured_result():
    """The spike-to-halt probe runs after the timed regions; if it never returns, the parked throughput line is
    printed (once) instead of re-running the benchmark on the fallback path."""
    from dist_utils import free_port
    res, lines = _run("probe-hang", free_port())
    assert res.returncode == 0, res.stderr[-2000:]
    assert len(lines) == 1, res.stdout
    out = json.loads(lines[0])
    assert out["comm"] == "auto" and "fallback" not in out
    assert out["spike_to_halt_ms"] == {"error": "probe did not complete"}
