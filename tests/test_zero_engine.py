"""ZeRO engine on CPU/gloo, world_size 2: sharded training must reproduce the single-process
result (this is the oracle tier for the GPU kernels — SURVEY.md §4 item 2), checkpoints must
round-trip and re-partition, and the monitor → halt → rollback → resume loop must work."""
import json
import os

import pytest
import torch

from distributed_llm_training_gpu_manager_b200.ai_engine import (DeepSpeedConfig, DeepSpeedLauncher, OffloadDevice,
                                                                  ZeROStage)
from distributed_llm_training_gpu_manager_b200.engine import initialize
from distributed_llm_training_gpu_manager_b200.engine import checkpoint as ckpt
from distributed_llm_training_gpu_manager_b200.engine.trainer import synthetic_batches
from distributed_llm_training_gpu_manager_b200.models import build_torch_model, get_config

from dist_utils import free_port, run_ranks

VOCAB, SEQ, MB = 512, 32, 2


def ds_dict(stage, gas=1, clip=1.0, threshold=1000, act_ckpt=False, lr=3e-3):
    cfg = DeepSpeedConfig(zero_stage=ZeROStage(stage), offload_optimizer=OffloadDevice.NONE,
                          offload_param=OffloadDevice.NONE, fp16_enabled=False, bf16_enabled=False,
                          train_micro_batch_size_per_gpu=MB, gradient_accumulation_steps=gas, gradient_clipping=clip,
                          learning_rate=lr, warmup_steps=2, total_training_steps=1000,
                          stage3_param_persistence_threshold=threshold, activation_checkpointing=act_ckpt,
                          wall_clock_breakdown=False)
    return DeepSpeedLauncher().generate_config(cfg)


def _model(name="gpt2-tiny"):
    torch.manual_seed(7)
    return build_torch_model(get_config(name))


def _train_rank(rank, world, stage, gas, steps, model_name, act_ckpt):
    eng = initialize(_model(model_name), ds_dict(stage, gas, act_ckpt=act_ckpt), device="cpu")
    data = synthetic_batches(VOCAB, MB, SEQ, "cpu", rank, world)
    losses = []
    while eng.global_steps < steps:
        loss = eng(*data(eng.micro_steps))
        eng.backward(loss)
        m = eng.step()
        if m:
            losses.append(m["loss"])
    full = {n: eng.full_param(n) for n, _ in eng.named_params}
    return losses, full, eng.layout.part_numel


def _train_single(stage_gas_steps, model_name, world=2):
    """Single process, same global batch: micro-batch = concat of every rank's micro-batch."""
    gas, steps = stage_gas_steps
    d = ds_dict(0, gas)
    d["train_micro_batch_size_per_gpu"] = MB * world
    eng = initialize(_model(model_name), d, device="cpu")
    srcs = [synthetic_batches(VOCAB, MB, SEQ, "cpu", r, world) for r in range(world)]
    losses = []
    while eng.global_steps < steps:
        parts = [s(eng.micro_steps) for s in srcs]
        ids = torch.cat([p[0] for p in parts]); lab = torch.cat([p[1] for p in parts])
        loss = eng(ids, lab)
        eng.backward(loss)
        m = eng.step()
        if m:
            losses.append(m["loss"])
    return losses, {n: eng.full_param(n) for n, _ in eng.named_params}


@pytest.mark.parametrize("stage,gas,model,act", [(1, 1, "gpt2-tiny", False), (2, 2, "gpt2-tiny", False),
                                                 (3, 1, "gpt2-tiny", True), (3, 2, "llama-tiny", False),
                                                 (0, 1, "llama-tiny", False)])
def test_sharded_matches_single_process(stage, gas, model, act):
    steps = 4
    ref_losses, ref_params = _train_single((gas, steps), model)
    out = run_ranks(_train_rank, 2, (stage, gas, steps, model, act))
    for losses, params, part_numel in out:
        if stage >= 1:
            assert part_numel > 0, "test config must actually partition something"
        assert losses == pytest.approx(ref_losses, rel=2e-4, abs=2e-5)
        for n, p in params.items():
            assert torch.allclose(p, ref_params[n], rtol=2e-4, atol=2e-5), n


def _ckpt_rank(rank, world, stage, tmp):
    eng = initialize(_model(), ds_dict(stage), device="cpu")
    data = synthetic_batches(VOCAB, MB, SEQ, "cpu", rank, world)
    for _ in range(3):
        eng.backward(eng(*data(eng.micro_steps))); eng.step()
    eng.save_checkpoint(tmp)
    cont = []
    for _ in range(2):
        eng.backward(eng(*data(eng.micro_steps))); cont.append(eng.step()["loss"])
    eng2 = initialize(_model(), ds_dict(stage), device="cpu")
    path, _ = eng2.load_checkpoint(tmp)
    assert path is not None and eng2.global_steps == 3
    again = []
    for _ in range(2):
        eng2.backward(eng2(*data(eng2.micro_steps))); again.append(eng2.step()["loss"])
    return cont, again


@pytest.mark.parametrize("stage", [1, 3])
def test_checkpoint_roundtrip_and_layout(stage, tmp_path):
    out = run_ranks(_ckpt_rank, 2, (stage, str(tmp_path)))
    for cont, again in out:
        assert again == pytest.approx(cont, rel=1e-5)
    tag = (tmp_path / "latest").read_text().strip()
    assert tag == "global_step3"
    names = sorted(os.listdir(tmp_path / tag))
    assert "manifest.json" in names
    for r in range(2):
        assert f"zero_pp_rank_{r}_mp_rank_00_optim_states.pt" in names
    if stage == 3:
        assert "zero_pp_rank_0_mp_rank_00_model_states.pt" in names and "zero_pp_rank_1_mp_rank_00_model_states.pt" in names
    else:
        assert "mp_rank_00_model_states.pt" in names
    assert ckpt.verify_checkpoint(str(tmp_path), tag) == (True, "ok")
    sd = torch.load(tmp_path / tag / "zero_pp_rank_0_mp_rank_00_optim_states.pt", weights_only=False)
    osd = sd["optimizer_state_dict"]
    assert osd["zero_stage"] == stage and osd["partition_count"] == 2
    assert ("fp32_flat_groups" if stage == 3 else "single_partition_of_fp32_groups") in osd
    # corrupt one byte -> detected
    f = tmp_path / tag / "zero_pp_rank_1_mp_rank_00_optim_states.pt"
    raw = bytearray(f.read_bytes()); raw[len(raw) // 2] ^= 0xFF; f.write_bytes(bytes(raw))
    ok, why = ckpt.verify_checkpoint(str(tmp_path), tag)
    assert not ok and "checksum" in why
    assert ckpt.find_rollback_tag(str(tmp_path)) is None


def test_elastic_resume_world_2_to_1(tmp_path):
    run_ranks(_ckpt_rank, 2, (2, str(tmp_path)))
    eng = initialize(_model(), ds_dict(2), device="cpu")          # world size 1 now
    path, _ = eng.load_checkpoint(str(tmp_path))
    assert path is not None and eng.global_steps == 3
    sds = ckpt._load_optim_files(tmp_path / "global_step3", 2, False)
    full = ckpt.consolidate_fp32(sds, "master")
    for n, _ in eng.named_params:
        assert torch.allclose(eng.full_param(n), full[n])


def test_launcher_spike_halt_rollback_resume(tmp_path):
    """BASELINE.json config 1: tiny GPT-2, ZeRO-1, CPU/gloo world_size 2 through the launcher;
    an injected loss spike must raise a CRITICAL alert, roll back to the last good checkpoint and
    the run must finish with a loss curve that re-joins."""
    cfg = DeepSpeedConfig(model_name="tiny", zero_stage=ZeROStage.OPTIMIZER_STATE, offload_optimizer=OffloadDevice.NONE,
                          offload_param=OffloadDevice.NONE, fp16_enabled=False, train_micro_batch_size_per_gpu=4,
                          gradient_accumulation_steps=1, learning_rate=3e-3, warmup_steps=5,
                          total_training_steps=200, num_gpus=2, master_port=free_port(),
                          stage3_param_persistence_threshold=1000, wall_clock_breakdown=False)
    launcher = DeepSpeedLauncher()
    result_file = tmp_path / "result.json"
    env_backup = dict(os.environ)
    os.environ.update(B200ZERO_INJECT_SPIKE_AT="26:40", B200ZERO_BACKEND="gloo", MASTER_ADDR="127.0.0.1")
    try:
        res = launcher.launch(cfg, "builtin:gpt2-tiny",
                              ["--steps=40", f"--save_dir={tmp_path / 'ckpt'}", "--save_every=10",
                               f"--result={result_file}", "--seq_len=32", "--window=20"])
        assert res.status == "launched", res.details
        rec = launcher.registry.wait(res.job_id, timeout=600)
    finally:
        os.environ.clear(); os.environ.update(env_backup)
    log = launcher.registry.tail(res.job_id) or ""
    assert rec is not None and rec.returncode == 0, log
    out = json.loads(result_file.read_text())
    assert out["status"] == "finished" and out["final_step"] == 40
    assert out["rollbacks"] == 1 and "spike" in out["alerts"], out
    assert "rolled back to global_step20" in log
    assert out["steps_run"] == 40 + 6            # steps 21..26 were replayed
    assert out["halt_latency_ms"] and out["halt_latency_ms"][0] < 50
    assert out["losses"][-1] < out["losses"][0]


def test_supervisor_restarts_after_rank_kill_and_resumes(tmp_path):
    """BASELINE.json config 4 (CPU tier): a rank is SIGKILLed mid-run (spot preemption of one
    worker); the supervisor tears the job down, relaunches it, and the workers resume from the
    newest verified checkpoint and finish."""
    cfg = DeepSpeedConfig(model_name="tiny", zero_stage=ZeROStage.GRADIENT_PARTITIONING,
                          offload_optimizer=OffloadDevice.NONE, offload_param=OffloadDevice.NONE, fp16_enabled=False,
                          train_micro_batch_size_per_gpu=4, gradient_accumulation_steps=1, learning_rate=3e-3,
                          warmup_steps=5, total_training_steps=200, num_gpus=2, master_port=free_port(),
                          wall_clock_breakdown=False)
    launcher = DeepSpeedLauncher()
    result_file = tmp_path / "result.json"
    env_backup = dict(os.environ)
    os.environ.update(B200ZERO_KILL_RANK_AT="17:1", B200ZERO_BACKEND="gloo", MASTER_ADDR="127.0.0.1",
                      B200ZERO_MAX_RESTARTS="2", B200ZERO_INJECT_NAN_GRAD_AT="23")
    try:
        res = launcher.launch(cfg, "builtin:gpt2-tiny",
                              ["--steps=30", f"--save_dir={tmp_path / 'ckpt'}", "--save_every=5",
                               f"--result={result_file}", "--seq_len=32", "--window=20"])
        rec = launcher.registry.wait(res.job_id, timeout=600)
    finally:
        os.environ.clear(); os.environ.update(env_backup)
    log = launcher.registry.tail(res.job_id) or ""
    assert rec is not None and rec.returncode == 0, log
    assert rec.last_step == 30 and rec.heartbeat_dir          # progress is visible through the registry (heartbeat files)
    assert "restart 1/2" in log and "resumed from global_step15" in log, log
    out = json.loads(result_file.read_text())
    assert out["status"] == "finished" and out["final_step"] == 30
    # the NaN gradient injected at step 23 (after the restart) was caught, the update skipped, and
    # the run rolled back to the last good checkpoint and carried on
    assert "divergence" in out["alerts"] and out["rollbacks"] == 1, out


def test_zero_to_fp32_consolidates_all_ranks(tmp_path):
    from distributed_llm_training_gpu_manager_b200.engine.zero_to_fp32 import consolidate
    run_ranks(_ckpt_rank, 2, (3, str(tmp_path)))
    sd, tag = consolidate(str(tmp_path))
    assert tag == "global_step3" and (tmp_path / "zero_to_fp32.py").exists()
    ref = build_torch_model(get_config("gpt2-tiny")).state_dict()
    assert set(sd) == {k for k in ref if not k.endswith("lm_head.weight")} or set(sd) <= set(ref)
    for k, v in sd.items():
        assert v.shape == ref[k].shape and v.dtype == torch.float32


def test_user_script_through_native_launcher(tmp_path):
    """examples/train_gpt2.py launched the way the reference launches a DeepSpeed script."""
    cfg = DeepSpeedConfig(model_name="ex", zero_stage=ZeROStage.GRADIENT_PARTITIONING, offload_optimizer=OffloadDevice.NONE,
                          offload_param=OffloadDevice.NONE, fp16_enabled=False, train_micro_batch_size_per_gpu=2,
                          gradient_accumulation_steps=2, num_gpus=2, master_port=free_port(), wall_clock_breakdown=False)
    launcher = DeepSpeedLauncher()
    script = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "examples", "train_gpt2.py")
    env_backup = dict(os.environ)
    os.environ.update(B200ZERO_BACKEND="gloo", MASTER_ADDR="127.0.0.1")
    try:
        res = launcher.launch(cfg, script, ["--steps", "12"])
        rec = launcher.registry.wait(res.job_id, timeout=300)
    finally:
        os.environ.clear(); os.environ.update(env_backup)
    log = launcher.registry.tail(res.job_id) or ""
    assert rec is not None and rec.returncode == 0, log
    assert "step 10 loss" in log
