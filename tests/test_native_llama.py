"""Native (hand-written kernel) Llama executor vs the nn.Module oracle with identical weights."""
import pytest
import torch

from distributed_llm_training_gpu_manager_b200.ai_engine import DeepSpeedConfig, DeepSpeedLauncher
from distributed_llm_training_gpu_manager_b200.engine import initialize
from distributed_llm_training_gpu_manager_b200.models import build_torch_model, get_config
from distributed_llm_training_gpu_manager_b200.models.native_llama import NativeLlamaEngine


def _ds(stage, bf16):
    return DeepSpeedLauncher().generate_config(DeepSpeedConfig(
        zero_stage=stage, offload_optimizer="none", offload_param="none", fp16_enabled=False, bf16_enabled=bf16,
        train_micro_batch_size_per_gpu=2, gradient_accumulation_steps=1, learning_rate=1e-3, warmup_steps=1,
        wall_clock_breakdown=False))


def _pair(dev, stage, bf16, seq):
    mc = get_config("llama-tiny")
    eng = NativeLlamaEngine(mc, _ds(stage, bf16), seq_len=seq, device=dev)
    m = build_torch_model(mc)
    m.load_state_dict({n: eng.full_param(n).cpu() for n in eng.layout.by_name})
    ref = initialize(m, _ds(0, False), device="cpu")   # fp32 oracle on CPU
    return mc, eng, ref


@pytest.mark.parametrize("stage", [0, 3])
def test_native_matches_module_cpu(stage):
    mc, eng, ref = _pair("cpu", stage, False, 32)
    torch.manual_seed(0)
    for _ in range(3):
        ids = torch.randint(0, mc.vocab_size, (2, 32)); lab = torch.randint(0, mc.vocab_size, (2, 32))
        a = eng.train_step(ids, lab)
        ref.backward(ref(ids, lab)); b = ref.step()
        assert a["loss"] == pytest.approx(b["loss"], rel=1e-4)
        assert a["grad_norm"] == pytest.approx(b["grad_norm"], rel=1e-3)


@pytest.mark.gpu
@pytest.mark.parametrize("stage", [0, 3])
def test_native_matches_module_gpu(stage):
    mc, eng, ref = _pair(torch.device("cuda", 0), stage, True, 256)
    torch.manual_seed(0)
    first = None
    for it in range(4):
        ids = torch.randint(0, mc.vocab_size, (2, 256)); lab = torch.randint(0, mc.vocab_size, (2, 256))
        a = eng.train_step(ids.cuda(), lab.cuda())
        ref.backward(ref(ids, lab)); b = ref.step()
        # bf16 kernels vs fp32 oracle: loss within 1 %, grad norm within 5 %
        assert a["loss"] == pytest.approx(b["loss"], rel=1e-2), (it, a, b)
        assert a["grad_norm"] == pytest.approx(b["grad_norm"], rel=5e-2), (it, a, b)
        first = first or a["loss"]
    assert not a["nonfinite"]


@pytest.mark.gpu
def test_native_loss_decreases_and_skips_nonfinite():
    mc = get_config("llama-tiny")
    eng = NativeLlamaEngine(mc, _ds(3, True), seq_len=256, device=torch.device("cuda", 0))
    ids = torch.randint(0, mc.vocab_size, (2, 256)).cuda()
    losses = [eng.train_step(ids, ids)["loss"] for _ in range(12)]
    assert losses[-1] < losses[0] - 0.5
    before = eng.master.clone()
    eng.fault_hooks["post_backward"] = lambda e: e.shard_grads.__setitem__(0, float("nan"))
    m = eng.train_step(ids, ids)
    assert m["nonfinite"] and torch.equal(before, eng.master)        # predicated update: nothing moved


@pytest.mark.gpu
@pytest.mark.parametrize("recompute", ["selective", "full", "none"])
def test_cpu_checkpointing_offloads_and_matches(recompute):
    """activation_checkpointing.cpu_checkpointing: saved activations live in pinned host memory
    between forward and backward; results must not change."""
    def ds(cpu):
        return DeepSpeedLauncher().generate_config(DeepSpeedConfig(
            zero_stage=3, offload_optimizer="none", offload_param="none", fp16_enabled=False, bf16_enabled=True,
            train_micro_batch_size_per_gpu=2, gradient_accumulation_steps=1, learning_rate=1e-3, warmup_steps=1,
            cpu_checkpointing=cpu, wall_clock_breakdown=False))
    mc = get_config("llama-tiny")
    a = NativeLlamaEngine(mc, ds(False), seq_len=256, device=torch.device("cuda", 0), recompute=recompute)
    b = NativeLlamaEngine(mc, ds(True), seq_len=256, device=torch.device("cuda", 0), recompute=recompute)
    assert b._act_offload is not None and a._act_offload is None
    g = torch.Generator().manual_seed(0)
    for _ in range(3):
        ids = torch.randint(0, mc.vocab_size, (2, 256), generator=g).cuda()
        ma, mb = a.train_step(ids, ids), b.train_step(ids, ids)
        assert ma["loss"] == pytest.approx(mb["loss"], rel=2e-3)
        assert ma["grad_norm"] == pytest.approx(mb["grad_norm"], rel=2e-2)


def test_capacity_planning_70b_on_8_b200():
    """BASELINE.json config 5 sizing: Llama-3-70B ZeRO-3 on 8 x 180 GB — what fits where."""
    from distributed_llm_training_gpu_manager_b200.models.native_llama import activation_bytes, static_bytes
    GB = 2 ** 30
    c70, c8 = get_config("llama3-70b"), get_config("llama3-8b")
    st = static_bytes(c70, 8)
    assert 69e9 < st["params_total"] < 72e9
    assert 140 * GB < st["device_total"] < 170 * GB          # fits in 180 GB with all state on the GPU ...
    act = activation_bytes(c70, 1, 4096)
    assert st["device_total"] + act["full"] < 178 * GB       # ... only with full recompute
    assert st["device_total"] + act["selective"] > 178 * GB
    off = static_bytes(c70, 8, offload_optimizer=True)       # host offload moves 12 B/param off the GPU
    assert off["device_total"] < 60 * GB and off["host_pinned"] > 95 * GB
    assert off["device_total"] + act["light"] < 178 * GB
    # the 8B headline config: one GPU holds everything, light recompute fits (measured peak 153 GB)
    s8, a8 = static_bytes(c8, 1), activation_bytes(c8, 2, 4096)
    assert 120 * GB < s8["device_total"] < 140 * GB and s8["device_total"] + a8["light"] < 172 * GB


def test_eval_loss_matches_forward_and_leaves_state_alone():
    from distributed_llm_training_gpu_manager_b200.engine import initialize
    ds = DeepSpeedLauncher().generate_config(DeepSpeedConfig(
        zero_stage=3, offload_optimizer="none", offload_param="none", bf16_enabled=False, fp16_enabled=False,
        train_micro_batch_size_per_gpu=2, gradient_accumulation_steps=1, learning_rate=1e-3, warmup_steps=1,
        wall_clock_breakdown=False))
    mc = get_config("llama-tiny")
    eng = NativeLlamaEngine(mc, ds, seq_len=64, device=torch.device("cpu"), seed=4)
    g = torch.Generator().manual_seed(0)
    ids = torch.randint(0, mc.vocab_size, (2, 64), generator=g)
    eng.train_step(ids, ids)
    before = eng.grad_view("lm_head.weight").clone()
    ev = float(eng.eval_loss(ids, ids))
    assert torch.equal(before, eng.grad_view("lm_head.weight"))          # evaluation does not touch gradients
    assert eng._saved is None
    assert ev == pytest.approx(float(eng.forward(ids, ids)), rel=1e-6)
    # the generic nn.Module engine offers the same call
    torch.manual_seed(0)
    mod = initialize(build_torch_model(get_config("gpt2-tiny")), ds, device=torch.device("cpu"))
    x = torch.randint(0, 512, (2, 32), generator=g)
    a = float(mod.eval_loss(x, x))
    assert a == pytest.approx(float(mod(x, x)), rel=1e-5) and mod.module.training
