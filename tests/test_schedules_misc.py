"""LR schedules, fp16 loss scaler, elastic batch solver, engine config parsing, fleet sampler."""

import pytest

from distributed_llm_training_gpu_manager_b200.ai_engine import (DeepSpeedConfig, DeepSpeedLauncher, FleetSampler,
                                                                  GPUManager)
from distributed_llm_training_gpu_manager_b200.engine.config import load_config
from distributed_llm_training_gpu_manager_b200.engine.schedules import (DynamicLossScaler, LRSchedule,
                                                                         elastic_batch_plan)


def test_warmup_decay_lr_shape():
    s = LRSchedule("WarmupDecayLR", {"warmup_min_lr": 0, "warmup_max_lr": 1e-3, "warmup_num_steps": 100,
                                     "total_num_steps": 1000}, 1e-3)
    assert s.lr_at(0) == 0.0
    assert 0 < s.lr_at(10) < s.lr_at(50) < 1e-3
    assert s.lr_at(100) == pytest.approx(1e-3)
    assert s.lr_at(550) == pytest.approx(0.5e-3)
    assert s.lr_at(1000) == 0.0 and s.lr_at(5000) == 0.0
    lin = LRSchedule("WarmupLR", {"warmup_max_lr": 2e-3, "warmup_num_steps": 10, "warmup_type": "linear"}, 2e-3)
    assert lin.lr_at(5) == pytest.approx(1e-3) and lin.lr_at(500) == 2e-3
    cos = LRSchedule("WarmupCosineLR", {"warmup_max_lr": 1.0, "warmup_num_steps": 0, "total_num_steps": 100}, 1.0)
    assert cos.lr_at(50) == pytest.approx(0.5 * (1 + 1e-4), rel=1e-3)
    with pytest.raises(ValueError):
        LRSchedule("Nope", {}, 1.0).lr_at(5)
    s.last_step = 42
    s2 = LRSchedule("WarmupDecayLR", s.p, 1e-3)
    s2.load_state_dict(s.state_dict())
    assert s2.last_step == 42


def test_dynamic_loss_scaler():
    sc = DynamicLossScaler(True, 0, init_power=4, window=3, hysteresis=2, min_scale=1)
    assert sc.scale == 16
    sc.update(True)            # first overflow is absorbed by hysteresis
    assert sc.scale == 16
    sc.update(True)
    assert sc.scale == 8
    for _ in range(3):
        sc.update(False)
    assert sc.scale == 16 and sc.skipped == 2
    static = DynamicLossScaler(True, 128.0)
    static.update(True)
    assert static.scale == 128.0
    off = DynamicLossScaler(False)
    assert off.scale == 1.0
    sd = sc.state_dict()
    sc2 = DynamicLossScaler(True, 0)
    sc2.load_state_dict(sd)
    assert sc2.scale == sc.scale


def test_elastic_plan_matches_emitted_block():
    cfg = DeepSpeedLauncher().generate_config(DeepSpeedConfig(elastic_training=True, num_gpus=8,
                                                              train_micro_batch_size_per_gpu=4,
                                                              gradient_accumulation_steps=8))
    batch, gpus = elastic_batch_plan(cfg["elasticity"], micro_batches=[1, 2, 4])
    assert batch <= cfg["elasticity"]["max_train_batch_size"] and batch % 4 == 0
    assert 1 in gpus and 8 in gpus and all(batch % n == 0 for n in gpus)


def test_engine_config_consumes_every_emitted_key():
    ds = DeepSpeedLauncher().generate_config(DeepSpeedConfig(zero_stage=3, bf16_enabled=True, cpu_checkpointing=True,
                                                             gradient_clipping=0.5, reduce_bucket_size=123456))
    c = load_config(ds)
    assert (c.stage, c.precision, c.grad_clip, c.reduce_bucket_size) == (3, "bf16", 0.5, 123456)
    assert c.offload_optimizer == "cpu" and c.offload_param == "cpu" and c.act_ckpt and c.cpu_checkpointing
    assert c.persistence_threshold == 100000 and c.sched_type == "WarmupDecayLR" and c.betas == (0.9, 0.999)
    import torch
    assert c.reduce_dtype == torch.bfloat16           # "fp16" comm dtype is not applied to bf16 gradients
    c1 = load_config(DeepSpeedLauncher().generate_config(DeepSpeedConfig(zero_stage=1)))
    assert c1.offload_param == "none" and c1.precision == "fp16" and c1.reduce_dtype == torch.float16
    assert c1.train_batch_size(8) == 4 * 8 * 8


def test_fleet_sampler_ring_buffer():
    class Fake(GPUManager):
        def get_fleet_status(self, xml_str=None):
            return self.get_mock_fleet()
    fs = FleetSampler(Fake(), interval_s=0.01, capacity=5).start()
    import time
    time.sleep(0.2)
    fs.stop()
    assert 1 <= len(fs.snapshots) <= 5 and fs.latest().total_gpus == 2
    assert set(fs.clocks_summary()) >= {"sm_mhz", "sm_max_mhz", "reasons"}


def test_preset_architectures_cover_the_launcher_presets():
    """The launcher's 7b / 13b / 70b presets (ref ai_engine/deepspeed_launcher.py:369-407) each map to an
    architecture whose size matches its name and whose head_dim is the 128 the attention kernels serve."""
    from distributed_llm_training_gpu_manager_b200.ai_engine import DeepSpeedLauncher
    from distributed_llm_training_gpu_manager_b200.models import PRESET_MODELS, get_config
    presets = DeepSpeedLauncher.presets()
    assert set(PRESET_MODELS) == set(presets)
    want = {"7b": 6.7e9, "13b": 13.0e9, "70b": 70.6e9}
    for name, model in PRESET_MODELS.items():
        cfg = get_config(model)
        assert abs(cfg.num_params() - want[name]) / want[name] < 0.01, (name, cfg.num_params())
        assert cfg.head_dim == 128 and cfg.n_heads % cfg.n_kv_heads == 0
        assert cfg.ffn % 8 == 0 and cfg.padded_vocab % 256 == 0
