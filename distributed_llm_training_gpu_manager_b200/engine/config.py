"""Engine-side view of the DeepSpeed-schema JSON that ``DeepSpeedLauncher.generate_config``
emits (every key of ref ``ai_engine/deepspeed_launcher.py:124-238`` is consumed here; the
reference only writes them to a file for the DeepSpeed CLI). Unknown keys are kept in
``raw`` so a config survives a save/load round trip unchanged."""
from __future__ import annotations

import json
from dataclasses import dataclass, field
from typing import Any, Dict, Optional, Union

import torch


@dataclass
class EngineConfig:
    micro_batch: int = 1
    grad_accum: int = 1
    grad_clip: float = 0.0
    steps_per_print: int = 100
    wall_clock_breakdown: bool = False
    # zero
    stage: int = 0
    reduce_bucket_size: int = 500_000_000
    allgather_bucket_size: int = 500_000_000
    overlap_comm: bool = True
    reduce_scatter: bool = True
    contiguous_gradients: bool = True
    max_live_parameters: int = 1_000_000_000
    max_reuse_distance: int = 1_000_000_000
    prefetch_bucket_size: int = 500_000_000
    persistence_threshold: int = 100_000
    gather_16bit_on_save: bool = True
    sub_group_size: int = 1_000_000_000
    offload_optimizer: str = "none"
    offload_param: str = "none"
    pin_memory: bool = True
    nvme_path: Optional[str] = None       # offload_optimizer.nvme_path (DeepSpeed key); default $B200ZERO_NVME_PATH or tmp
    offload_buffer_count: int = 4         # offload_optimizer.buffer_count: staging slots of the NVMe tier
    # precision
    precision: str = "fp32"  # fp32 | fp16 | bf16
    loss_scale: float = 0.0
    loss_scale_window: int = 1000
    initial_scale_power: int = 16
    hysteresis: int = 2
    min_loss_scale: float = 1.0
    comm_dtype: Optional[str] = None
    prescale_gradients: bool = False
    predivide_factor: float = 1.0
    grad_accum_dtype: Optional[str] = None  # None -> compute dtype when grad_accum == 1 else fp32
    # optimizer / schedule
    lr: float = 1e-3
    betas: tuple = (0.9, 0.999)
    eps: float = 1e-8
    weight_decay: float = 0.01
    sched_type: Optional[str] = None
    sched_params: Dict[str, Any] = field(default_factory=dict)
    # activation checkpointing
    act_ckpt: bool = False
    cpu_checkpointing: bool = False
    # elasticity
    elastic: Dict[str, Any] = field(default_factory=dict)
    raw: Dict[str, Any] = field(default_factory=dict)

    @property
    def compute_dtype(self) -> torch.dtype:
        return {"fp32": torch.float32, "fp16": torch.float16, "bf16": torch.bfloat16}[self.precision]

    @property
    def reduce_dtype(self) -> torch.dtype:
        # The reference emits communication_data_type "fp16" unconditionally, even under bf16
        # (SURVEY.md §2.3); reducing bf16 gradients in fp16 would overflow, so fp16 is honoured
        # only for fp16 compute and otherwise mapped to the compute dtype.
        name = (self.comm_dtype or "").lower()
        if name in ("fp32", "float32"):
            return torch.float32
        if name in ("bf16", "bfloat16") and self.precision != "fp16":
            return torch.bfloat16 if self.precision == "bf16" else torch.float32
        if name in ("fp16", "float16") and self.precision == "fp16":
            return torch.float16
        return self.compute_dtype

    @property
    def accum_dtype(self) -> torch.dtype:
        if self.grad_accum_dtype:
            return {"fp32": torch.float32, "bf16": torch.bfloat16, "fp16": torch.float16}[self.grad_accum_dtype]
        if self.precision == "fp32" or self.grad_accum > 1:
            return torch.float32
        return self.compute_dtype

    def train_batch_size(self, world_size: int) -> int:
        return self.micro_batch * self.grad_accum * world_size


# How the native engine treats every key the launcher can emit (ref ``deepspeed_launcher.py:124-238``).
# "honoured" = changes what the engine does; "fixed" = the engine has one behaviour and says which; "n/a" = nothing to
# configure in this design. ``python -m <pkg> config --explain`` prints the table for a concrete config, and
# tests/test_round2_fixes.py fails if the launcher learns to emit a key that is missing here.
KEY_HANDLING: Dict[str, str] = {
    "train_micro_batch_size_per_gpu": "honoured: tokens per rank per micro-step; the native executor is built for it",
    "gradient_accumulation_steps": "honoured: micro-steps per optimizer step; gradients accumulate in fp32 when > 1",
    "gradient_clipping": "honoured: global-norm clipping, coefficient computed on the device (0 disables it)",
    "steps_per_print": "honoured: caps the packaged worker's log interval",
    "wall_clock_breakdown": "honoured: per-phase CUDA-event timers in the step metrics",
    "dump_state": "n/a: state is inspected through engine.memory_report() and the checkpoints",
    "communication_data_type": "honoured for fp16 compute; bf16 / fp32 compute reduce in their own dtype (an fp16 sum of "
                               "bf16 gradients would overflow)",
    "prescale_gradients": "honoured on the NCCL / gloo path; the fused reduce kernel sums in fp32 and scales once",
    "gradient_predivide_factor": "honoured together with prescale_gradients",
    "optimizer.type": "fixed: AdamW with fp32 master weights (device kernel, host C++ or NVMe-streamed)",
    "optimizer.params.lr": "honoured", "optimizer.params.betas": "honoured", "optimizer.params.eps": "honoured",
    "optimizer.params.weight_decay": "honoured (decoupled)",
    "scheduler.type": "honoured: WarmupLR, WarmupDecayLR, WarmupCosineLR, constant",
    "scheduler.params.warmup_min_lr": "honoured", "scheduler.params.warmup_max_lr": "honoured",
    "scheduler.params.warmup_num_steps": "honoured", "scheduler.params.total_num_steps": "honoured",
    "fp16.enabled": "honoured: fp16 compute with the dynamic loss scaler; collectives run on NCCL (the fused kernels "
                    "reduce bf16)",
    "fp16.loss_scale": "honoured (0 = dynamic)", "fp16.loss_scale_window": "honoured",
    "fp16.initial_scale_power": "honoured", "fp16.hysteresis": "honoured", "fp16.min_loss_scale": "honoured",
    "bf16.enabled": "honoured: the B200 default; enables the native kernels and the fused collectives",
    "zero_optimization.stage": "honoured: 0 / 1 / 2 / 3",
    "zero_optimization.allgather_partitions": "fixed: updated rows are always all-gathered (stage <= 2: pushed by the "
                                              "optimizer kernel)",
    "zero_optimization.allgather_bucket_size": "fixed: one unit (transformer block) per gather; audited against the size",
    "zero_optimization.overlap_comm": "honoured: communication on a side stream, prefetch one block ahead",
    "zero_optimization.reduce_scatter": "honoured: false = all-reduce and keep own rows (NCCL / gloo path only)",
    "zero_optimization.reduce_bucket_size": "fixed: one unit per reduction; audited against the size",
    "zero_optimization.contiguous_gradients": "fixed: gradients always live in flat per-unit buffers",
    "zero_optimization.round_robin_gradients": "n/a: every tensor is row-partitioned over all ranks, so every rank owns "
                                               "an equal share of every bucket by construction",
    "zero_optimization.stage3_max_live_parameters": "honoured: number of gathered blocks kept resident (2 ... all)",
    "zero_optimization.stage3_max_reuse_distance": "honoured: below one block nothing is kept beyond the pipeline's two",
    "zero_optimization.stage3_prefetch_bucket_size": "honoured: 0 = gather on demand, > 0 = prefetch one block ahead",
    "zero_optimization.stage3_param_persistence_threshold": "honoured: smaller tensors are replicated, not partitioned",
    "zero_optimization.stage3_gather_16bit_weights_on_model_save": "honoured",
    "zero_optimization.sub_group_size": "honoured: caps the tile of the host optimizer",
    "zero_optimization.offload_optimizer.device": "honoured: none / cpu / nvme",
    "zero_optimization.offload_optimizer.pin_memory": "honoured",
    "zero_optimization.offload_optimizer.buffer_count": "honoured: staging slots of the NVMe tier (the cpu tier "
                                                        "double-buffers)",
    "zero_optimization.offload_optimizer.fast_init": "n/a: host state is initialised from the device shard in one pass",
    "zero_optimization.offload_param.device": "honoured: cpu (nvme is served from the same pinned-host tier)",
    "zero_optimization.offload_param.pin_memory": "honoured through offload_optimizer.pin_memory",
    "zero_optimization.offload_param.buffer_count": "fixed: one device staging buffer per resident gathered block + root",
    "zero_optimization.offload_param.buffer_size": "fixed: a staging buffer holds one unit's rows of the shard",
    "zero_optimization.offload_param.max_in_cpu": "n/a: the whole 16-bit shard lives on the host (2 bytes per parameter "
                                                  "of this rank)",
    "activation_checkpointing.partition_activations": "n/a at tensor-parallel degree 1 (nothing to partition across)",
    "activation_checkpointing.cpu_checkpointing": "honoured: saved activations go to pinned host memory on a side "
                                                  "stream and come back one block ahead of backward",
    "activation_checkpointing.contiguous_memory_optimization": "n/a: saved activations are per-block tensors",
    "activation_checkpointing.number_checkpoints": "fixed: one checkpoint per transformer block",
    "activation_checkpointing.synchronize_checkpoint_boundary": "n/a: ordering is by streams and events",
    "activation_checkpointing.profile": "n/a: use wall_clock_breakdown / B200ZERO_NVTX",
    "elasticity.enabled": "honoured: engine.schedules.elastic_batch_plan + re-partition on load at another world size",
    "elasticity.max_train_batch_size": "honoured", "elasticity.min_train_batch_size": "honoured",
    "elasticity.min_gpus": "honoured", "elasticity.max_gpus": "honoured", "elasticity.prefer_larger_batch_size": "honoured",
}


def key_paths(d: Dict[str, Any], prefix: str = ""):
    for k, v in d.items():
        if isinstance(v, dict):
            yield from key_paths(v, f"{prefix}{k}.")
        else:
            yield f"{prefix}{k}"


def explain(config: Dict[str, Any]) -> Dict[str, str]:
    """key path -> how this engine treats it, for every key present in ``config``."""
    return {p: KEY_HANDLING.get(p, "not interpreted by the native engine (kept verbatim in checkpoints)")
            for p in key_paths(config)}


def audit(c: EngineConfig, unit_numels=(), world_size: int = 1, fused: bool = False) -> list:
    """What this engine does with the keys of the emitted config that have no literal counterpart here
    (ref ``deepspeed_launcher.py:133-142,186-194``). Returns human-readable notes; the engine logs them once.
    Keys that change arithmetic or traffic are honoured in the engine itself (``reduce_scatter``,
    ``prescale_gradients`` / ``gradient_predivide_factor``, ``stage3_*``, ``sub_group_size``, offload blocks)."""
    notes = []
    biggest = max(unit_numels, default=0)
    if c.stage >= 2 and biggest > c.reduce_bucket_size:
        notes.append(f"reduce_bucket_size={c.reduce_bucket_size}: gradients are reduced one unit (transformer block) at a "
                     f"time and the largest unit has {biggest} elements; the unit is not split further")
    if c.stage >= 1 and c.stage <= 2 and biggest > c.allgather_bucket_size:
        notes.append(f"allgather_bucket_size={c.allgather_bucket_size}: updated rows are gathered one unit at a time "
                     f"(largest: {biggest} elements)")
    if not c.contiguous_gradients:
        notes.append("contiguous_gradients=false: gradients always live in flat per-unit buffers here; nothing to turn off")
    if not c.reduce_scatter and c.stage >= 1 and world_size > 1:
        notes.append("reduce_scatter=false: gradients are all-reduced and every rank keeps its rows; the fused backend "
                     "(which always reduce-scatters) is not used for this job")
    if (c.prescale_gradients or c.predivide_factor != 1.0) and fused:
        notes.append("prescale_gradients / gradient_predivide_factor: the fused reduce kernel sums in fp32 and scales "
                     "once afterwards, so there is nothing to pre-scale; honoured on the NCCL / gloo path")
    if c.stage >= 3 and c.prefetch_bucket_size <= 0:
        notes.append("stage3_prefetch_bucket_size=0: blocks are gathered synchronously, on demand")
    return notes


def load_config(src: Union[str, Dict[str, Any], EngineConfig]) -> EngineConfig:
    if isinstance(src, EngineConfig):
        return src
    if isinstance(src, str):
        with open(src) as f:
            src = json.load(f)
    d = dict(src)
    z = d.get("zero_optimization", {}) or {}
    c = EngineConfig(raw=d)
    c.micro_batch = int(d.get("train_micro_batch_size_per_gpu", 1))
    c.grad_accum = int(d.get("gradient_accumulation_steps", 1))
    c.grad_clip = float(d.get("gradient_clipping", 0.0))
    c.steps_per_print = int(d.get("steps_per_print", 100))
    c.wall_clock_breakdown = bool(d.get("wall_clock_breakdown", False))
    c.stage = int(z.get("stage", 0))
    c.reduce_bucket_size = int(z.get("reduce_bucket_size", c.reduce_bucket_size))
    c.allgather_bucket_size = int(z.get("allgather_bucket_size", c.allgather_bucket_size))
    c.overlap_comm = bool(z.get("overlap_comm", True))
    c.reduce_scatter = bool(z.get("reduce_scatter", True))
    c.contiguous_gradients = bool(z.get("contiguous_gradients", True))
    c.max_live_parameters = int(z.get("stage3_max_live_parameters", c.max_live_parameters))
    c.max_reuse_distance = int(z.get("stage3_max_reuse_distance", c.max_reuse_distance))
    c.prefetch_bucket_size = int(z.get("stage3_prefetch_bucket_size", c.prefetch_bucket_size))
    # below stage 3 nothing needs to stay resident, so only tiny tensors are left unpartitioned
    c.persistence_threshold = int(z.get("stage3_param_persistence_threshold",
                                        c.persistence_threshold if c.stage >= 3 else 4096))
    c.gather_16bit_on_save = bool(z.get("stage3_gather_16bit_weights_on_model_save", True))
    c.sub_group_size = int(z.get("sub_group_size", c.sub_group_size))
    oo, op = z.get("offload_optimizer") or {}, z.get("offload_param") or {}
    # offload blocks are emitted for any stage by the launcher; they only apply where meaningful
    c.offload_optimizer = oo.get("device", "none") if c.stage >= 1 else "none"
    c.offload_param = op.get("device", "none") if c.stage >= 3 else "none"
    c.pin_memory = bool(oo.get("pin_memory", True))
    c.nvme_path = oo.get("nvme_path")
    c.offload_buffer_count = int(oo.get("buffer_count", 4))
    if (d.get("bf16") or {}).get("enabled"):
        c.precision = "bf16"
    elif (d.get("fp16") or {}).get("enabled"):
        c.precision = "fp16"
        f16 = d["fp16"]
        c.loss_scale = float(f16.get("loss_scale", 0))
        c.loss_scale_window = int(f16.get("loss_scale_window", 1000))
        c.initial_scale_power = int(f16.get("initial_scale_power", 16))
        c.hysteresis = int(f16.get("hysteresis", 2))
        c.min_loss_scale = float(f16.get("min_loss_scale", 1))
    c.comm_dtype = d.get("communication_data_type")
    c.prescale_gradients = bool(d.get("prescale_gradients", False))
    c.predivide_factor = float(d.get("gradient_predivide_factor", 1.0))
    c.grad_accum_dtype = (d.get("data_types") or {}).get("grad_accum_dtype")
    opt = (d.get("optimizer") or {}).get("params", {})
    c.lr = float(opt.get("lr", c.lr))
    c.betas = tuple(opt.get("betas", c.betas))
    c.eps = float(opt.get("eps", c.eps))
    c.weight_decay = float(opt.get("weight_decay", c.weight_decay))
    sch = d.get("scheduler") or {}
    c.sched_type = sch.get("type")
    c.sched_params = dict(sch.get("params", {}))
    ac = d.get("activation_checkpointing")
    c.act_ckpt = ac is not None
    c.cpu_checkpointing = bool((ac or {}).get("cpu_checkpointing", False))
    c.elastic = dict(d.get("elasticity") or {})
    return c
