"""Packaged training worker (``builtin:<model>``): what the reference expects the user to bring
(its launcher starts an un-shipped ``training_script``). One process per rank."""
from __future__ import annotations

import argparse
import json
import os
import sys

import torch
import torch.distributed as dist


def init_distributed(backend: str = "auto"):
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    use_cuda = torch.cuda.is_available() and backend != "gloo"
    if use_cuda:
        dev_index = int(os.environ.get("B200ZERO_DEVICE", os.environ.get("LOCAL_RANK", "0")))
        torch.cuda.set_device(dev_index % torch.cuda.device_count())
        device = torch.device("cuda", torch.cuda.current_device())
    else:
        device = torch.device("cpu")
    if world > 1 and not dist.is_initialized():
        be = "nccl" if use_cuda else "gloo"
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        kw = {"device_id": device} if use_cuda else {}
        dist.init_process_group(be, rank=rank, world_size=world, **kw)
    return rank, world, device


def main(argv=None) -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="gpt2-tiny")
    ap.add_argument("--local_rank", type=int, default=0)
    ap.add_argument("--deepspeed_config", required=True)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--seq_len", type=int, default=0)
    ap.add_argument("--save_dir", default=None)
    ap.add_argument("--save_every", type=int, default=0)
    ap.add_argument("--metrics", default=None)
    ap.add_argument("--result", default=None)
    ap.add_argument("--engine", default="auto", choices=["auto", "module", "native"])
    ap.add_argument("--comm", default="auto")
    ap.add_argument("--window", type=int, default=100)
    ap.add_argument("--seed", type=int, default=1234)
    ap.add_argument("--data", default=None, help="flat binary token file (uint16/uint32); default: synthetic tokens")
    ap.add_argument("--data_dtype", default="auto", help="auto | uint16 | uint32 | int32 | int64")
    ap.add_argument("--init_hf", default=None,
                    help="directory of a Hugging Face Llama / GPT-2 checkpoint to start from (architecture taken from its config)")
    ap.add_argument("--eval_data", default=None, help="token file for the validation loss")
    ap.add_argument("--eval_every", type=int, default=0)
    ap.add_argument("--eval_steps", type=int, default=4)
    ns = ap.parse_args(argv)
    if os.environ.get("B200ZERO_HANG_DUMP"):  # debugging aid: dump all stacks and exit after N seconds
        import faulthandler
        faulthandler.dump_traceback_later(int(os.environ["B200ZERO_HANG_DUMP"]), exit=True)

    rank, world, device = init_distributed(os.environ.get("B200ZERO_BACKEND", "auto"))
    from ..ai_engine.loss_monitor import MonitorConfig
    from ..models import build_torch_model, get_config
    from .config import load_config
    from .trainer import Trainer, TrainerConfig, synthetic_batches

    hf_named = None
    if ns.init_hf:
        # every rank reads the checkpoint and keeps its own rows (load_full_params); fine for one node with host RAM to spare
        from transformers import AutoModelForCausalLM
        from ..models import hf as hfmod
        hf_model = AutoModelForCausalLM.from_pretrained(ns.init_hf, torch_dtype=torch.float32)
        if hf_model.config.model_type == "gpt2":
            mcfg = hfmod.gpt2_config_from_hf(hf_model.config)
            hf_named = hfmod.gpt2_from_hf_state_dict(hf_model.state_dict(), mcfg)
        else:
            mcfg = hfmod.config_from_hf(hf_model.config)
            hf_named = hfmod.from_hf_state_dict(hf_model.state_dict(), mcfg)
        del hf_model
    else:
        mcfg = get_config(ns.model)
    ecfg = load_config(ns.deepspeed_config)
    seq = ns.seq_len or min(mcfg.max_seq, 128)
    torch.manual_seed(ns.seed)
    use_native = ns.engine == "native" or (ns.engine == "auto" and mcfg.family == "llama" and device.type == "cuda"
                                           and ecfg.precision == "bf16")
    if use_native:
        from ..models.native_llama import NativeLlamaEngine
        engine = NativeLlamaEngine(mcfg, ecfg, seq_len=seq, device=device, comm=ns.comm, seed=ns.seed)
    else:
        from .zero import ZeroEngine
        engine = ZeroEngine(build_torch_model(mcfg), ecfg, device=device, comm=ns.comm)
    if hf_named is not None:
        if engine.source is not None:
            engine.load_full_params(hf_named)
        else:       # nn.Module engine: parameters are views of the engine's buffers, loaded through the same call
            engine.load_full_params({k: v for k, v in hf_named.items() if k in engine.layout.by_name}, strict=False)
        del hf_named
    if ns.data:
        from .data import token_batches
        batches = token_batches(ns.data, ecfg.micro_batch, seq, device, rank, world, seed=ns.seed,
                                dtype=ns.data_dtype, vocab_size=mcfg.vocab_size)
    else:
        batches = synthetic_batches(mcfg.vocab_size, ecfg.micro_batch, seq, device, rank, world, seed=ns.seed)
    eval_batches = None
    if ns.eval_data and ns.eval_every:
        from .data import token_batches
        eval_batches = token_batches(ns.eval_data, ecfg.micro_batch, seq, device, rank, world, seed=ns.seed + 1,
                                     shuffle=False, dtype=ns.data_dtype, vocab_size=mcfg.vocab_size, prefetch=0)
    tcfg = TrainerConfig(total_steps=ns.steps, save_dir=ns.save_dir, save_every=ns.save_every,
                         metrics_path=ns.metrics, monitor=MonitorConfig(window_size=ns.window),
                         log_every=min(10, max(1, ecfg.steps_per_print)),   # steps_per_print caps the log interval
                         eval_batches=eval_batches, eval_every=ns.eval_every, eval_steps=ns.eval_steps)
    res = Trainer(engine, batches, tcfg).run()
    if rank == 0 and ns.result:
        with open(ns.result, "w") as f:
            json.dump({"status": res.status, "steps_run": res.steps_run, "final_step": res.final_step,
                       "rollbacks": res.rollbacks, "alerts": [a["alert_type"] for a in res.alerts],
                       "halt_latency_ms": res.halt_latency_ms,
                       "losses": [h["loss"] for h in res.history], "steps": [h["step"] for h in res.history]}, f)
    if dist.is_initialized():
        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
