"""Host AdamW (C++) vs torch.optim.AdamW, and the offloaded engine step vs the on-device one."""
import pytest
import torch

from distributed_llm_training_gpu_manager_b200.engine.offload import host_adamw


@pytest.mark.parametrize("gdt", [torch.float32, torch.bfloat16])
def test_host_adam_matches_torch(gdt):
    torch.manual_seed(0)
    n = (1 << 17) + 37
    p0 = torch.randn(n)
    ref = torch.nn.Parameter(p0.clone())
    opt = torch.optim.AdamW([ref], lr=1e-2, weight_decay=0.01)
    master, m, v = p0.clone(), torch.zeros(n), torch.zeros(n)
    out = torch.empty(n, dtype=torch.bfloat16)
    for step in range(1, 4):
        g = torch.randn(n).to(gdt)
        ref.grad = g.float() * 0.5
        opt.step()
        host_adamw(master, m, v, g, out, lr=1e-2, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=0.01, step=step,
                   grad_scale=0.5, threads=4)
    assert torch.allclose(master, ref.data, rtol=1e-5, atol=1e-6)
    assert torch.equal(out, master.bfloat16())


@pytest.mark.gpu
def test_offloaded_engine_matches_device_engine():
    from distributed_llm_training_gpu_manager_b200.ai_engine import DeepSpeedConfig, DeepSpeedLauncher
    from distributed_llm_training_gpu_manager_b200.models import get_config
    from distributed_llm_training_gpu_manager_b200.models.native_llama import NativeLlamaEngine

    def ds(off):
        return DeepSpeedLauncher().generate_config(DeepSpeedConfig(
            zero_stage=3, offload_optimizer=off, offload_param="none", bf16_enabled=True, fp16_enabled=False,
            train_micro_batch_size_per_gpu=2, gradient_accumulation_steps=1, learning_rate=1e-3, warmup_steps=1,
            wall_clock_breakdown=False))
    mc = get_config("llama-tiny")
    a = NativeLlamaEngine(mc, ds("none"), seq_len=256, device=torch.device("cuda", 0))
    b = NativeLlamaEngine(mc, ds("cpu"), seq_len=256, device=torch.device("cuda", 0))
    assert b.master.device.type == "cpu" and b.master.is_pinned()
    g = torch.Generator().manual_seed(0)
    for _ in range(3):
        ids = torch.randint(0, mc.vocab_size, (2, 256), generator=g).cuda()
        ma, mb = a.train_step(ids, ids), b.train_step(ids, ids)
        assert ma["loss"] == pytest.approx(mb["loss"], rel=2e-3)
    # Adam moves every element by ~lr per step, so elements whose gradient is ~0 may differ in sign
    # between two non-bit-identical backward passes; compare in the mean
    diff = (a.master.cpu() - b.master).abs()
    assert diff.mean().item() < 3e-4 and (diff > 5e-3).float().mean().item() < 1e-3
