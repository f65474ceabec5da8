"""bench.py end to end on the CPU self-test path (tiny model, gloo): our arm and the stock-PyTorch library arm go
through the same harness and print the contract's keys. Numbers from this mode are not benchmark values."""
import json
import os
import subprocess
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
KEYS = {"metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
        "vs_baseline", "dtype", "data", "config", "clocks", "e2e", "gpu_launches", "impl"}


def _bench(extra, nproc=1, port=None):
    env = dict(os.environ, B200ZERO_BENCH_CPU="1", OMP_NUM_THREADS="2")
    tail = ["--model", "llama-tiny", "--seq-len", "64", "--steps", "2", "--warmup", "3", "--gpus", str(nproc)] + extra
    if nproc == 1:
        cmd = [sys.executable, str(ROOT / "bench.py")] + tail
    else:
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(nproc),
               "--master-addr", "127.0.0.1", "--master-port", str(port), str(ROOT / "bench.py")] + tail
    res = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=600, cwd=str(ROOT))
    assert res.returncode == 0, res.stderr[-3000:]
    lines = [ln for ln in res.stdout.splitlines() if ln.startswith("{")]
    assert len(lines) == 1, res.stdout
    return json.loads(lines[0])


def test_ours_single_process_line_has_every_contract_key():
    out = _bench([])
    assert KEYS <= set(out), KEYS - set(out)
    assert out["impl"] == "ours" and out["n_gpus"] == 1 and out["value"] > 0
    assert out["e2e"]["h2d_bytes_per_step"] == 2 * 2 * 64 * 8 and out["e2e"]["d2h_bytes_per_step"] == 20
    halt = out["spike_to_halt_ms"]
    assert halt["actions"] == {"nan_grad": "rollback", "nan_grad_mid": "rollback", "loss_spike_x50": "rollback"}
    assert halt["nan_grad_mid"]["from_fault_written"] > 0


@pytest.mark.parametrize("impl", ["torch", "ours"])
def test_two_ranks_under_torchrun(impl):
    from dist_utils import free_port
    out = _bench(["--impl", impl, "--recompute", "none"], nproc=2, port=free_port())
    assert KEYS <= set(out), KEYS - set(out)
    assert out["impl"] == impl and out["n_gpus"] == 2 and out["config"]["global_batch"] == 4
    assert out["config"]["activation_recompute"] == "none"
    assert "fallback" not in out
    if impl == "torch":
        assert out["gpu_launches"] == 0 and "unavailable" in out["spike_to_halt_ms"]


def test_library_arm_checkpoints_only_where_the_model_would_not_fit():
    """Same answer on every rank (no free-memory query): Llama-3-8B with 2 x 4096 tokens needs per-block checkpointing
    on one 180 GB GPU (fp32 master + gradient + two moments alone are 128 GB) and none from two GPUs on."""
    sys.path.insert(0, str(ROOT))
    from baseline.torch_zero3 import pick_checkpointing
    from distributed_llm_training_gpu_manager_b200.models import get_config
    c = get_config("llama3-8b")
    hbm = 183 * 2 ** 30
    assert pick_checkpointing(c, 1, 2, 4096, hbm) is True
    assert pick_checkpointing(c, 2, 2, 4096, hbm) is False and pick_checkpointing(c, 8, 2, 4096, hbm) is False
    assert pick_checkpointing(c, 1, 2, 4096, None) is False           # CPU self-test: nothing to fit into
