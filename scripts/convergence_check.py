"""End-to-end numerics check at a non-toy size: train Llama-1B-shaped models on the learnable synthetic
stream with (a) the native sm_100a executor and (b) the nn.Module oracle on cuBLAS/SDPA, same data,
same config, and record both loss curves (gpurun_out/convergence_llama1b.json)."""
import json, os, sys, time
from pathlib import Path
import torch
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from distributed_llm_training_gpu_manager_b200.ai_engine import DeepSpeedConfig, DeepSpeedLauncher
from distributed_llm_training_gpu_manager_b200.engine import initialize
from distributed_llm_training_gpu_manager_b200.engine.trainer import synthetic_batches
from distributed_llm_training_gpu_manager_b200.models import build_torch_model, get_config
from distributed_llm_training_gpu_manager_b200.models.native_llama import NativeLlamaEngine

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 120
ref_steps = int(sys.argv[2]) if len(sys.argv) > 2 else 40
S, MB = 1024, 4
mc = get_config("llama-1b", vocab_size=32000, n_layers=8)
ds = DeepSpeedLauncher().generate_config(DeepSpeedConfig(
    zero_stage=3, offload_optimizer="none", offload_param="none", bf16_enabled=True, fp16_enabled=False,
    train_micro_batch_size_per_gpu=MB, gradient_accumulation_steps=1, learning_rate=3e-4, warmup_steps=10,
    total_training_steps=1000, wall_clock_breakdown=False))
dev = torch.device("cuda", 0)
data = synthetic_batches(mc.vocab_size, MB, S, dev, learnable=True)
out = {"model": mc.to_dict(), "seq": S, "micro_batch": MB}
eng = NativeLlamaEngine(mc, ds, seq_len=S, device=dev)
t0 = time.time()
out["native"] = [eng.train_step(*data(i))["loss"] for i in range(steps)]
out["native_s"] = time.time() - t0
weights = {n: eng.full_param(n) for n in eng.layout.by_name} if False else None
del eng
torch.cuda.empty_cache()
# oracle: identical init (same seeded ParamSource), module path
eng0 = NativeLlamaEngine(mc, ds, seq_len=S, device=dev)
m = build_torch_model(mc)
m.load_state_dict({n: eng0.full_param(n).cpu() for n in eng0.layout.by_name})
del eng0
torch.cuda.empty_cache()
ref = initialize(m, ds, device=dev)
losses = []
for i in range(ref_steps):
    ref.backward(ref(*data(i)))
    losses.append(ref.step()["loss"])
out["module_oracle"] = losses
n = min(ref_steps, steps)
out["max_abs_loss_diff_first_%d" % n] = max(abs(a - b) for a, b in zip(out["native"][:n], losses[:n]))
os.makedirs("gpurun_out", exist_ok=True)
json.dump(out, open("gpurun_out/convergence_llama1b.json", "w"), indent=1)
print("native first/last:", out["native"][0], out["native"][-1], "oracle last:", losses[-1],
      "max diff:", out["max_abs_loss_diff_first_%d" % n])
