#!/usr/bin/env python
"""A user training script in the DeepSpeed shape — what the reference's launcher expects you to bring
(it ships none). Launch it exactly as the reference would launch a DeepSpeed script:

    python -m distributed_llm_training_gpu_manager_b200.engine.launch --num_gpus=2 examples/train_gpt2.py \
        --deepspeed_config=/tmp/ds_config.json --steps 50

or from Python: ``DeepSpeedLauncher().launch(cfg, "examples/train_gpt2.py", ["--steps", "50"])``.
The config JSON is the one ``DeepSpeedLauncher.generate_config`` emits (ZeRO stage, precision, LR
schedule, clipping, activation checkpointing … are all honoured by the engine)."""
import argparse
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_llm_training_gpu_manager_b200.ai_engine import LossSpikeMonitor, TrainingMetrics  # noqa: E402
from distributed_llm_training_gpu_manager_b200.engine import initialize  # noqa: E402
from distributed_llm_training_gpu_manager_b200.engine.worker import init_distributed  # noqa: E402
from distributed_llm_training_gpu_manager_b200.models import build_torch_model, get_config  # noqa: E402


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--local_rank", type=int, default=0)
    ap.add_argument("--deepspeed_config", required=True)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--model", default="gpt2-tiny")
    ns = ap.parse_args()
    rank, world, device = init_distributed(os.environ.get("B200ZERO_BACKEND", "auto"))
    mcfg = get_config(ns.model)
    engine = initialize(model=build_torch_model(mcfg), config=ns.deepspeed_config, device=device)
    monitor = LossSpikeMonitor()
    g = torch.Generator().manual_seed(rank)
    while engine.global_steps < ns.steps:
        ids = torch.randint(0, mcfg.vocab_size, (engine.cfg.micro_batch, 64), generator=g).to(device)
        loss = engine(ids, ids)
        engine.backward(loss)
        m = engine.step()
        if m is not None:
            alerts = monitor.ingest(TrainingMetrics(step=m["step"], loss=m["loss"], gradient_norm=m["grad_norm"],
                                                    learning_rate=m["lr"]))
            if rank == 0 and (m["step"] % 10 == 0 or alerts):
                print(f"step {m['step']} loss {m['loss']:.4f} alerts {[a.alert_type for a in alerts]}", flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
