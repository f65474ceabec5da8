"""Training-job endpoints (routes and request models: ref ``backend/routers/training.py:16-153``).

``dry_run`` still defaults to true. Bad enum strings now produce HTTP 422 from request
validation instead of an unhandled ``ValueError`` → 500."""
from typing import List, Literal, Optional

from fastapi import APIRouter, HTTPException
from pydantic import BaseModel, Field

from ...ai_engine.deepspeed_launcher import (DeepSpeedConfig, DeepSpeedLauncher, LaunchResult, OffloadDevice,
                                             ZeROStage)

router = APIRouter()
launcher = DeepSpeedLauncher()

_Offload = Literal["none", "cpu", "nvme"]


class TrainingLaunchRequest(BaseModel):
    training_script: str = Field(..., description="Path to training script (or builtin:<model>)")
    model_name: str = Field(default="custom_model")
    script_args: Optional[List[str]] = None
    zero_stage: int = Field(default=3, ge=0, le=3)
    offload_optimizer: _Offload = "cpu"
    offload_param: _Offload = "cpu"
    micro_batch_size: int = Field(default=4, ge=1)
    gradient_accumulation_steps: int = Field(default=8, ge=1)
    fp16: bool = True
    bf16: bool = False
    learning_rate: float = Field(default=3e-5, gt=0)
    warmup_steps: int = 100
    total_steps: int = 10000
    num_gpus: int = Field(default=1, ge=1)
    num_nodes: int = Field(default=1, ge=1)
    activation_checkpointing: bool = True
    dry_run: bool = Field(default=True, description="If true, only generates config without launching")

    def to_config(self) -> DeepSpeedConfig:
        return DeepSpeedConfig(
            model_name=self.model_name, zero_stage=ZeROStage(self.zero_stage),
            offload_optimizer=OffloadDevice(self.offload_optimizer), offload_param=OffloadDevice(self.offload_param),
            train_micro_batch_size_per_gpu=self.micro_batch_size,
            gradient_accumulation_steps=self.gradient_accumulation_steps, fp16_enabled=self.fp16,
            bf16_enabled=self.bf16, learning_rate=self.learning_rate, warmup_steps=self.warmup_steps,
            total_training_steps=self.total_steps, num_gpus=self.num_gpus, num_nodes=self.num_nodes,
            activation_checkpointing=self.activation_checkpointing)


class PresetLaunchRequest(BaseModel):
    preset: str = Field(..., description="Preset name: 7b | 13b | 70b")
    training_script: str = Field(...)
    script_args: Optional[List[str]] = None
    dry_run: bool = True


def _all_presets():
    table = dict(launcher.presets())
    table.update(launcher.b200_presets())
    return table


@router.post("/launch", response_model=LaunchResult)
async def launch_training(request: TrainingLaunchRequest):
    return launcher.launch(request.to_config(), request.training_script, request.script_args, request.dry_run)


@router.post("/launch/preset", response_model=LaunchResult)
async def launch_from_preset(request: PresetLaunchRequest):
    table = _all_presets()
    if request.preset not in table:
        raise HTTPException(status_code=400,
                            detail=f"Unknown preset '{request.preset}'. Available: {list(table.keys())}")
    return launcher.launch(table[request.preset], request.training_script, request.script_args, request.dry_run)


@router.get("/presets")
async def list_presets():
    return {name: {"zero_stage": c.zero_stage.value, "offload_optimizer": c.offload_optimizer.value,
                   "offload_param": c.offload_param.value, "micro_batch_size": c.train_micro_batch_size_per_gpu,
                   "gradient_accumulation": c.gradient_accumulation_steps,
                   "precision": "bf16" if c.bf16_enabled else "fp16", "num_gpus": c.num_gpus,
                   "num_nodes": c.num_nodes, "activation_checkpointing": c.activation_checkpointing}
            for name, c in _all_presets().items()}


@router.post("/config/generate")
async def generate_config(request: TrainingLaunchRequest, explain: bool = False):
    """``?explain=true`` adds ``native_engine_handling``: key by key, what the built-in engine does with the emitted
    DeepSpeed key (honoured / fixed / not applicable, and why)."""
    cfg = request.to_config()
    # build the command against a named-but-unwritten path: generating a config must not
    # leave temp files behind (the reference writes one as a side effect, ref :274-275)
    path = f"/tmp/ds_config_{cfg.model_name}.json"
    ds = launcher.generate_config(cfg)
    out = {"deepspeed_config": ds,
           "launch_command": launcher.build_launch_command(cfg, request.training_script, request.script_args,
                                                           config_path=path),
           "effective_batch_size": cfg.effective_batch_size,
           "warnings": launcher.validate(cfg)}
    if explain:
        from ...engine.config import explain as _explain
        out["native_engine_handling"] = _explain(ds)
    return out
