"""Is the serial schedule itself run-to-run deterministic, and does the overlapped optimizer differ from it by more
than that? (flash-attention backward accumulates dQ with bulk reductions, i.e. in arrival order.)"""
import os
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from distributed_llm_training_gpu_manager_b200.ai_engine import DeepSpeedConfig, DeepSpeedLauncher  # noqa: E402
from distributed_llm_training_gpu_manager_b200.models import get_config  # noqa: E402
from distributed_llm_training_gpu_manager_b200.models.native_llama import NativeLlamaEngine  # noqa: E402

ds = DeepSpeedLauncher().generate_config(DeepSpeedConfig(
    zero_stage=3, offload_optimizer="none", offload_param="none", bf16_enabled=True, fp16_enabled=False,
    train_micro_batch_size_per_gpu=2, gradient_accumulation_steps=1, learning_rate=1e-3, warmup_steps=1,
    wall_clock_breakdown=False))
mc = get_config("llama-tiny", n_layers=4)


def run(flag, sync_after_step=False, steps=8):
    os.environ["B200ZERO_ADAM_OVERLAP"] = flag
    eng = NativeLlamaEngine(mc, ds, seq_len=256, device=torch.device("cuda", 0), seed=3)
    g = torch.Generator().manual_seed(0)
    losses = []
    for _ in range(steps):
        ids = torch.randint(0, mc.vocab_size, (2, 256), generator=g).cuda()
        losses.append(eng.train_step(ids, ids)["loss"])
        if sync_after_step:
            torch.cuda.synchronize()
    return losses, eng.full_param("blocks.1.w_down.weight").clone()


a, wa = run("0")
b, wb = run("0")
print("serial vs serial      : losses equal", a == b, "weights equal", torch.equal(wa, wb), "max|dw|", float((wa - wb).abs().max()))
c, wc = run("1", sync_after_step=True)
print("serial vs overlap+sync: losses equal", a == c, "weights equal", torch.equal(wa, wc), "max|dw|", float((wa - wc).abs().max()))
d, wd = run("1")
print("serial vs overlap     : losses equal", a == d, "weights equal", torch.equal(wa, wd), "max|dw|", float((wa - wd).abs().max()))
print("losses serial ", a)
print("losses overlap", d)
print("rel diff of last loss", abs(a[-1] - d[-1]) / abs(a[-1]), "serial-serial", abs(a[-1] - b[-1]) / abs(a[-1]))
