"""`python -m distributed_llm_training_gpu_manager_b200 …` — JSON on stdout for every sub-command."""
import json

from distributed_llm_training_gpu_manager_b200.__main__ import main


def _run(capsys, *argv):
    assert main(list(argv)) == 0
    return json.loads(capsys.readouterr().out)


def test_config_from_preset_with_overrides(capsys):
    out = _run(capsys, "config", "--preset", "7b", "--set", "zero_stage=2", "--set", "bf16_enabled=true",
               "--set", "fp16_enabled=false")
    ds = out["deepspeed_config"]
    assert ds["zero_optimization"]["stage"] == 2 and ds["bf16"]["enabled"] is True and "fp16" not in ds
    assert out["effective_batch_size"] == 2 * 16 * 4


def test_launch_dry_run_uses_builtin_worker(capsys):
    out = _run(capsys, "launch", "--preset", "13b", "--script", "builtin:llama2-13b", "--dry-run")
    assert out["status"] == "dry_run" and out["num_gpus"] == 8
    assert "builtin:llama2-13b" in out["command"] and "--deepspeed_config=" in out["command"]


def test_fleet_topology_and_selection_on_mocks(capsys):
    fleet = _run(capsys, "fleet", "--mock")
    assert fleet["total_gpus"] == 2 and fleet["devices"][1]["health"] == "warning"
    topo = _run(capsys, "topology", "--mock")
    assert topo["node_type"].startswith("HGX B200") and len(topo["links_per_gpu"]) == 8
    assert _run(capsys, "select", "--count", "1", "--mock")["selected"] == [0]
    assert _run(capsys, "select", "--count", "2", "--mock")["selected"] == []      # GPU 1 is over its memory threshold


def test_doctor_reports_every_check_and_an_overall_verdict():
    import json
    import subprocess
    import sys
    res = subprocess.run([sys.executable, "-m", "distributed_llm_training_gpu_manager_b200", "doctor", "--gpus", "2"],
                         capture_output=True, text=True, timeout=120)
    assert res.returncode == 0, res.stderr[-1500:]
    out = json.loads(res.stdout)
    assert set(out) == {"ready", "checks"}
    for name in ("kernel_library", "cuda_devices", "host_memory", "memlock_limit", "cpu_cores"):
        assert {"ok", "detail", "needed_for_default_job"} <= set(out["checks"][name]), name
    import torch
    if not torch.cuda.is_available():
        assert out["ready"] is False and out["checks"]["cuda_devices"]["detail"]["wanted"] == 2
