"""bench.py's spike-to-halt probe (second half of the BASELINE.json metric) on a tiny CPU engine: both
injected faults must end in a rollback decision and leave the engine usable."""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))


def test_measure_spike_to_halt_cpu():
    import bench
    from distributed_llm_training_gpu_manager_b200.ai_engine import DeepSpeedConfig, DeepSpeedLauncher
    from distributed_llm_training_gpu_manager_b200.engine import initialize
    from distributed_llm_training_gpu_manager_b200.models import build_torch_model, get_config
    ds = DeepSpeedLauncher().generate_config(DeepSpeedConfig(
        zero_stage=1, offload_optimizer="none", offload_param="none", bf16_enabled=False, fp16_enabled=False,
        train_micro_batch_size_per_gpu=2, gradient_accumulation_steps=1, learning_rate=1e-3, warmup_steps=1,
        wall_clock_breakdown=False))
    mc = get_config("gpt2-tiny")
    torch.manual_seed(0)
    eng = initialize(build_torch_model(mc), ds, device=torch.device("cpu"))
    ids = torch.randint(0, mc.vocab_size, (2, 32))
    res = bench.measure_spike_to_halt(eng, (ids, ids), [])
    for name in ("nan_grad", "loss_spike_x50"):
        assert res[name]["action"] == "rollback", res
        assert 0 < res[name]["from_backward_end"] <= res[name]["from_forward_end"]
    assert any(a.startswith("divergence") for a in res["nan_grad"]["alerts"])
    assert any(a.startswith("spike") or a.startswith("divergence") for a in res["loss_spike_x50"]["alerts"])
    # the engine still trains afterwards
    eng.backward(eng(ids, ids))
    m = eng.step()
    assert m is not None and m["loss"] == m["loss"]


def test_measure_spike_to_halt_native_engine_cpu():
    import bench
    from distributed_llm_training_gpu_manager_b200.ai_engine import DeepSpeedConfig, DeepSpeedLauncher
    from distributed_llm_training_gpu_manager_b200.models import get_config
    from distributed_llm_training_gpu_manager_b200.models.native_llama import NativeLlamaEngine
    ds = DeepSpeedLauncher().generate_config(DeepSpeedConfig(
        zero_stage=3, offload_optimizer="none", offload_param="none", bf16_enabled=False, fp16_enabled=False,
        train_micro_batch_size_per_gpu=2, gradient_accumulation_steps=1, learning_rate=1e-3, warmup_steps=1,
        wall_clock_breakdown=False))
    mc = get_config("llama-tiny")
    eng = NativeLlamaEngine(mc, ds, seq_len=64, device=torch.device("cpu"), seed=1)
    ids = torch.randint(0, mc.vocab_size, (2, 64))
    res = bench.measure_spike_to_halt(eng, (ids, ids), [])
    assert res["nan_grad"]["action"] == "rollback" and res["nan_grad"]["alerts"] == ["divergence/critical"], res
    assert res["loss_spike_x50"]["action"] == "rollback" and "spike/critical" in res["loss_spike_x50"]["alerts"], res
