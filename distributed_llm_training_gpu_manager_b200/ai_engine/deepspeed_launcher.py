"""ZeRO training launcher: typed config -> DeepSpeed-schema JSON -> launch command -> supervised job.

Public surface kept from the reference (``ai_engine/deepspeed_launcher.py:22-407``):
``ZeROStage``, ``OffloadDevice``, ``DeepSpeedConfig`` (same 37 fields and defaults,
ref ``:35-87``), ``LaunchResult`` (ref ``:90-100``) and ``DeepSpeedLauncher`` with
``generate_config / write_config / build_launch_command / launch / presets``.
The emitted JSON is key-for-key what the reference emits (golden-tested against
SURVEY.md Appendix A.1), because the native engine consumes exactly that schema.

What is different, deliberately:

* The default engine is *ours* (``engine="native"``): the command starts
  ``python -m <pkg>.engine.launch``, which accepts the same ``--num_gpus /
  --num_nodes / --master_addr / --master_port`` flags as the ``deepspeed`` CLI and
  spawns one process per GPU. ``engine="deepspeed"`` (or an explicit
  ``deepspeed_path``) reproduces the reference command string verbatim.
* Commands are built as argv lists (``shlex``-quoted only for display), so script
  arguments containing spaces survive (the reference's ``command.split()`` mangles
  them).
* Config files get collision-free names, and launched processes are kept in a
  :class:`JobRegistry` that drains their output to a log file and exposes
  status / stop / logs (the reference drops the ``Popen`` handle, ref ``:354-362``).
* Nonsense combinations the reference silently emits are still emitted (parity) but
  reported by :meth:`DeepSpeedLauncher.validate`.
"""
from __future__ import annotations

import itertools
import json
import os
import shlex
import signal
import subprocess
import sys
import tempfile
import threading
import time
from datetime import datetime, timezone
from enum import Enum
from pathlib import Path
from typing import Any, Callable, Dict, List, Optional, Sequence

from pydantic import BaseModel, Field

PKG = __name__.split(".")[0]
NATIVE_LAUNCH_MODULE = f"{PKG}.engine.launch"


def _utc_now() -> datetime:
    return datetime.now(timezone.utc).replace(tzinfo=None)


class ZeROStage(int, Enum):
    DISABLED = 0
    OPTIMIZER_STATE = 1
    GRADIENT_PARTITIONING = 2
    FULL_PARTITIONING = 3


class OffloadDevice(str, Enum):
    NONE = "none"
    CPU = "cpu"
    NVME = "nvme"


class DeepSpeedConfig(BaseModel):
    """Structured training configuration (field names and defaults: ref ``:35-87``)."""

    # partitioning
    zero_stage: ZeROStage = ZeROStage.FULL_PARTITIONING
    offload_optimizer: OffloadDevice = OffloadDevice.CPU
    offload_param: OffloadDevice = OffloadDevice.CPU
    # batch
    train_micro_batch_size_per_gpu: int = Field(default=4, ge=1)
    gradient_accumulation_steps: int = Field(default=8, ge=1)
    gradient_clipping: float = Field(default=1.0, ge=0)
    # precision
    fp16_enabled: bool = True
    bf16_enabled: bool = False
    fp16_loss_scale: float = 0
    fp16_loss_scale_window: int = 1000
    # schedule
    learning_rate: float = Field(default=3e-5, gt=0)
    lr_scheduler_type: str = "WarmupDecayLR"
    warmup_steps: int = 100
    total_training_steps: int = 10000
    # communication
    communication_data_type: str = "fp16"
    prescale_gradients: bool = False
    gradient_predivide_factor: float = 1.0
    # activation checkpointing
    activation_checkpointing: bool = True
    contiguous_memory_optimization: bool = True
    cpu_checkpointing: bool = False
    # stage-3 knobs
    stage3_max_live_parameters: int = 1_000_000_000
    stage3_max_reuse_distance: int = 1_000_000_000
    stage3_prefetch_bucket_size: int = 500_000_000
    stage3_param_persistence_threshold: int = 100_000
    stage3_gather_16bit_weights_on_model_save: bool = True
    reduce_bucket_size: int = 500_000_000
    # misc
    elastic_training: bool = False
    wall_clock_breakdown: bool = True
    dump_state: bool = False
    # topology
    model_name: str = "custom_model"
    num_gpus: int = Field(default=1, ge=1)
    num_nodes: int = Field(default=1, ge=1)
    master_addr: str = "localhost"
    master_port: int = 29500

    # -- derived quantities -------------------------------------------------
    @property
    def world_size(self) -> int:
        return self.num_gpus * self.num_nodes

    @property
    def effective_batch_size(self) -> int:
        return self.train_micro_batch_size_per_gpu * self.gradient_accumulation_steps * self.world_size

    @property
    def precision(self) -> str:
        return "bf16" if self.bf16_enabled else ("fp16" if self.fp16_enabled else "fp32")


class LaunchResult(BaseModel):
    job_id: str
    config_path: str
    command: str
    status: str = "launched"
    num_gpus: int
    num_nodes: int
    effective_batch_size: int
    timestamp: str = Field(default_factory=lambda: _utc_now().isoformat())
    details: Dict[str, Any] = Field(default_factory=dict)


# ---------------------------------------------------------------------------
# config sections — each returns the fragment it owns; generate_config stitches
# them in the reference's key order (Appendix A.1)
# ---------------------------------------------------------------------------
def _zero_section(c: DeepSpeedConfig) -> Dict[str, Any]:
    z: Dict[str, Any] = {"stage": int(c.zero_stage.value)}
    bucket = c.reduce_bucket_size
    z.update(allgather_partitions=True, allgather_bucket_size=bucket, overlap_comm=True, reduce_scatter=True,
             reduce_bucket_size=bucket, contiguous_gradients=True, round_robin_gradients=True)
    if c.zero_stage is ZeROStage.FULL_PARTITIONING:
        for key in ("stage3_max_live_parameters", "stage3_max_reuse_distance", "stage3_prefetch_bucket_size",
                    "stage3_param_persistence_threshold", "stage3_gather_16bit_weights_on_model_save"):
            z[key] = getattr(c, key)
        z["sub_group_size"] = 1_000_000_000
    # Offload blocks are emitted for any stage, exactly as the reference does (ref :197-212).
    if c.offload_optimizer is not OffloadDevice.NONE:
        z["offload_optimizer"] = {"device": c.offload_optimizer.value, "pin_memory": True, "buffer_count": 4,
                                  "fast_init": False}
    if c.offload_param is not OffloadDevice.NONE:
        z["offload_param"] = {"device": c.offload_param.value, "pin_memory": True, "buffer_count": 5,
                              "buffer_size": 1_000_000_000, "max_in_cpu": 1_000_000_000}
    return z


def _schedule_section(c: DeepSpeedConfig) -> Dict[str, Any]:
    return {"type": c.lr_scheduler_type,
            "params": {"warmup_min_lr": 0, "warmup_max_lr": c.learning_rate, "warmup_num_steps": c.warmup_steps,
                       "total_num_steps": c.total_training_steps}}


def _optimizer_section(c: DeepSpeedConfig) -> Dict[str, Any]:
    return {"type": "AdamW", "params": {"lr": c.learning_rate, "betas": [0.9, 0.999], "eps": 1e-8,
                                        "weight_decay": 0.01}}


def _precision_section(c: DeepSpeedConfig) -> Dict[str, Any]:
    if c.bf16_enabled:  # bf16 wins when both are set
        return {"bf16": {"enabled": True}}
    if c.fp16_enabled:
        return {"fp16": {"enabled": True, "loss_scale": c.fp16_loss_scale,
                         "loss_scale_window": c.fp16_loss_scale_window, "initial_scale_power": 16,
                         "hysteresis": 2, "min_loss_scale": 1}}
    return {}


def _activation_section(c: DeepSpeedConfig) -> Dict[str, Any]:
    if not c.activation_checkpointing:
        return {}
    return {"activation_checkpointing": {"partition_activations": True, "cpu_checkpointing": c.cpu_checkpointing,
                                         "contiguous_memory_optimization": c.contiguous_memory_optimization,
                                         "number_checkpoints": None, "synchronize_checkpoint_boundary": False,
                                         "profile": False}}


def _elastic_section(c: DeepSpeedConfig) -> Dict[str, Any]:
    if not c.elastic_training:
        return {}
    per_node = c.train_micro_batch_size_per_gpu * c.gradient_accumulation_steps * c.num_gpus
    return {"elasticity": {"enabled": True, "max_train_batch_size": per_node * 2,
                           "min_train_batch_size": c.train_micro_batch_size_per_gpu, "min_gpus": 1,
                           "max_gpus": c.world_size, "prefer_larger_batch_size": True}}


# ---------------------------------------------------------------------------
# job registry (new: the reference has none)
# ---------------------------------------------------------------------------
class JobRecord(BaseModel):
    job_id: str
    pid: Optional[int] = None
    argv: List[str] = Field(default_factory=list)
    config_path: str = ""
    log_path: str = ""
    status: str = "pending"  # pending | running | finished | failed | stopped
    returncode: Optional[int] = None
    started_at: str = Field(default_factory=lambda: _utc_now().isoformat())
    finished_at: Optional[str] = None
    restarts: int = 0
    heartbeat_dir: str = ""                        # every rank's engine writes its step number here after each optimizer step
    last_step: Optional[int] = None                # newest optimizer step any rank has completed (filled in by get / list)
    seconds_since_progress: Optional[float] = None


class JobRegistry:
    """Keeps child process handles, drains their output into per-job log files and
    answers status/stop/logs queries. Thread-safe."""

    def __init__(self, log_dir: Optional[str] = None):
        self.log_dir = Path(log_dir or os.path.join(tempfile.gettempdir(), "b200zero_jobs"))
        self._lock = threading.Lock()
        self._jobs: Dict[str, JobRecord] = {}
        self._procs: Dict[str, subprocess.Popen] = {}
        self._callbacks: Dict[str, List[Callable[[JobRecord], None]]] = {}

    def spawn(self, job_id: str, argv: Sequence[str], env: Dict[str, str], config_path: str = "") -> JobRecord:
        self.log_dir.mkdir(parents=True, exist_ok=True)
        log_path = self.log_dir / f"{job_id}.log"
        logf = open(log_path, "ab", buffering=0)
        proc = subprocess.Popen(list(argv), stdout=logf, stderr=subprocess.STDOUT, env=env, start_new_session=True)
        logf.close()  # the child holds its own descriptor; nothing to drain, nothing to block on
        rec = JobRecord(job_id=job_id, pid=proc.pid, argv=list(argv), config_path=config_path,
                        log_path=str(log_path), status="running", heartbeat_dir=env.get("B200ZERO_HEARTBEAT_DIR", ""))
        with self._lock:
            self._jobs[job_id] = rec
            self._procs[job_id] = proc
        threading.Thread(target=self._reap, args=(job_id,), daemon=True, name=f"reap-{job_id}").start()
        return rec

    def _reap(self, job_id: str) -> None:
        proc = self._procs[job_id]
        rc = proc.wait()
        with self._lock:
            rec = self._jobs[job_id]
            rec.returncode = rc
            rec.finished_at = _utc_now().isoformat()
            if rec.status != "stopped":
                rec.status = "finished" if rc == 0 else "failed"
            cbs = list(self._callbacks.get(job_id, ()))
        for cb in cbs:
            try:
                cb(rec)
            except Exception:
                pass

    def on_exit(self, job_id: str, cb: Callable[[JobRecord], None]) -> None:
        with self._lock:
            self._callbacks.setdefault(job_id, []).append(cb)

    @staticmethod
    def _with_progress(rec: JobRecord) -> JobRecord:
        """Copy of ``rec`` with the newest completed step and the time since then, read from the heartbeat files."""
        out = rec.model_copy()
        try:
            files = [os.path.join(rec.heartbeat_dir, f) for f in os.listdir(rec.heartbeat_dir)] if rec.heartbeat_dir else []
        except OSError:
            files = []
        best, newest = None, None
        for f in files:
            try:
                step = int(open(f).read().strip() or 0)
                mtime = os.path.getmtime(f)
            except (OSError, ValueError):
                continue
            best = step if best is None else max(best, step)
            newest = mtime if newest is None else max(newest, mtime)
        out.last_step = best
        if newest is not None and rec.status == "running":
            out.seconds_since_progress = round(max(0.0, time.time() - newest), 1)
        return out

    def get(self, job_id: str) -> Optional[JobRecord]:
        with self._lock:
            rec = self._jobs.get(job_id)
            return self._with_progress(rec) if rec else None

    def list(self) -> List[JobRecord]:
        with self._lock:
            return [self._with_progress(r) for r in self._jobs.values()]

    def stop(self, job_id: str, sig: int = signal.SIGTERM, grace_s: float = 30.0) -> Optional[JobRecord]:
        """SIGTERM the job's process group (workers take an emergency checkpoint on
        SIGTERM), escalate to SIGKILL after ``grace_s``."""
        with self._lock:
            proc = self._procs.get(job_id)
            rec = self._jobs.get(job_id)
            if proc is None or rec is None:
                return None
            if proc.poll() is None:
                rec.status = "stopped"
        if proc.poll() is None:
            try:
                os.killpg(proc.pid, sig)
            except (ProcessLookupError, PermissionError):
                pass
            deadline = time.time() + grace_s
            while proc.poll() is None and time.time() < deadline:
                time.sleep(0.05)
            if proc.poll() is None:
                try:
                    os.killpg(proc.pid, signal.SIGKILL)
                except (ProcessLookupError, PermissionError):
                    pass
                proc.wait()
        return self.get(job_id)

    def wait(self, job_id: str, timeout: Optional[float] = None) -> Optional[JobRecord]:
        proc = self._procs.get(job_id)
        if proc is None:
            return None
        try:
            proc.wait(timeout=timeout)
        except subprocess.TimeoutExpired:
            pass
        # let the reaper thread publish the final state
        for _ in range(100):
            rec = self.get(job_id)
            if rec and (rec.returncode is not None or proc.poll() is None):
                break
            time.sleep(0.01)
        return self.get(job_id)

    def tail(self, job_id: str, max_bytes: int = 65536) -> Optional[str]:
        rec = self.get(job_id)
        if rec is None or not rec.log_path or not os.path.exists(rec.log_path):
            return None
        with open(rec.log_path, "rb") as f:
            f.seek(0, os.SEEK_END)
            size = f.tell()
            f.seek(max(0, size - max_bytes))
            return f.read().decode(errors="replace")


_default_registry = JobRegistry()
_name_counter = itertools.count()


def default_registry() -> JobRegistry:
    return _default_registry


class DeepSpeedLauncher:
    """Generates ZeRO configs, builds launch commands and runs supervised jobs."""

    def __init__(self, deepspeed_path: str = "deepspeed", engine: Optional[str] = None,
                 registry: Optional[JobRegistry] = None, python: Optional[str] = None):
        self.deepspeed_path = deepspeed_path
        # An explicitly customised CLI path means "really use that CLI" (ref :111 seam).
        self.engine = engine or ("native" if deepspeed_path == "deepspeed" else "deepspeed")
        if self.engine not in ("native", "deepspeed"):
            raise ValueError(f"engine must be 'native' or 'deepspeed', got {self.engine!r}")
        self.registry = registry or default_registry()
        self.python = python or sys.executable

    # -- config -----------------------------------------------------------
    def generate_config(self, config: DeepSpeedConfig) -> Dict[str, Any]:
        out: Dict[str, Any] = {
            "train_micro_batch_size_per_gpu": config.train_micro_batch_size_per_gpu,
            "gradient_accumulation_steps": config.gradient_accumulation_steps,
            "gradient_clipping": config.gradient_clipping,
            "steps_per_print": 100,
            "wall_clock_breakdown": config.wall_clock_breakdown,
            "dump_state": config.dump_state,
            "zero_optimization": _zero_section(config),
            "scheduler": _schedule_section(config),
            "optimizer": _optimizer_section(config),
            "communication_data_type": config.communication_data_type,
            "prescale_gradients": config.prescale_gradients,
            "gradient_predivide_factor": config.gradient_predivide_factor,
        }
        for section in (_precision_section, _activation_section, _elastic_section):
            out.update(section(config))
        return out

    @staticmethod
    def validate(config: DeepSpeedConfig) -> List[str]:
        """Human-readable warnings for combinations the schema allows but an engine
        would reject or silently ignore (the reference emits them without comment)."""
        notes: List[str] = []
        if config.zero_stage.value < 3 and config.offload_param is not OffloadDevice.NONE:
            notes.append("offload_param is only meaningful with ZeRO stage 3; it will be ignored")
        if config.zero_stage.value == 0 and config.offload_optimizer is not OffloadDevice.NONE:
            notes.append("offload_optimizer requires ZeRO stage >= 1; it will be ignored")
        if config.bf16_enabled and config.fp16_enabled:
            notes.append("both bf16 and fp16 requested; bf16 takes precedence")
        if config.bf16_enabled and config.communication_data_type == "fp16":
            notes.append("communication_data_type 'fp16' with bf16 compute: the native engine reduces in bf16/fp32")
        if config.offload_optimizer is OffloadDevice.NVME:
            notes.append("offload_optimizer=nvme: state files go to offload_optimizer.nvme_path, else $B200ZERO_NVME_PATH, "
                         "else <tmp>/b200zero_nvme (the reference schema has no nvme_path field)")
        if config.offload_param is OffloadDevice.NVME:
            notes.append("offload_param=nvme is accepted in the schema; parameters are kept sharded in HBM by the native engine")
        if config.elastic_training and config.num_nodes > 1:
            notes.append("elasticity.max_train_batch_size ignores num_nodes (reference behaviour kept)")
        return notes

    def write_config(self, config: DeepSpeedConfig, output_path: Optional[str] = None) -> str:
        payload = self.generate_config(config)
        if output_path is None:
            stamp = _utc_now().strftime("%Y%m%d_%H%M%S")
            base = Path(tempfile.gettempdir()) / f"ds_config_{config.model_name}_{stamp}.json"
            output_path = str(base)
            # second-resolution stamps collide (ref :250); disambiguate instead of overwriting
            while os.path.exists(output_path):
                output_path = str(base.with_name(f"{base.stem}_{os.getpid()}_{next(_name_counter)}.json"))
        tmp = f"{output_path}.tmp{os.getpid()}"
        with open(tmp, "w") as f:
            json.dump(payload, f, indent=2)
        os.replace(tmp, output_path)
        return output_path

    # -- command ----------------------------------------------------------
    def build_launch_argv(self, config: DeepSpeedConfig, training_script: str,
                          script_args: Optional[List[str]] = None, config_path: Optional[str] = None,
                          engine: Optional[str] = None) -> List[str]:
        if config_path is None:
            config_path = self.write_config(config)
        engine = engine or self.engine
        argv: List[str] = ([self.deepspeed_path] if engine == "deepspeed"
                           else [self.python, "-m", NATIVE_LAUNCH_MODULE])
        if config.num_nodes > 1:
            argv += [f"--num_nodes={config.num_nodes}", f"--master_addr={config.master_addr}",
                     f"--master_port={config.master_port}"]
        argv += [f"--num_gpus={config.num_gpus}", training_script, f"--deepspeed_config={config_path}"]
        argv += list(script_args or [])
        return argv

    def build_launch_command(self, config: DeepSpeedConfig, training_script: str,
                             script_args: Optional[List[str]] = None, config_path: Optional[str] = None,
                             engine: Optional[str] = None) -> str:
        argv = self.build_launch_argv(config, training_script, script_args, config_path, engine)
        if (engine or self.engine) == "deepspeed":
            return " ".join(argv)  # byte-for-byte what the reference prints (ref :300)
        return shlex.join(argv)

    # -- launch -----------------------------------------------------------
    def launch(self, config: DeepSpeedConfig, training_script: str, script_args: Optional[List[str]] = None,
               dry_run: bool = False) -> LaunchResult:
        config_path = self.write_config(config)
        argv = self.build_launch_argv(config, training_script, script_args, config_path)
        command = " ".join(argv) if self.engine == "deepspeed" else shlex.join(argv)
        stamp = _utc_now().strftime("%Y%m%d_%H%M%S")
        job_id = f"ds_{config.model_name}_{stamp}"
        if self.registry.get(job_id) is not None:
            job_id = f"{job_id}_{next(_name_counter)}"
        result = LaunchResult(
            job_id=job_id, config_path=config_path, command=command, num_gpus=config.world_size,
            num_nodes=config.num_nodes, effective_batch_size=config.effective_batch_size,
            details={"zero_stage": config.zero_stage.value, "offload_optimizer": config.offload_optimizer.value,
                     "offload_param": config.offload_param.value, "precision": config.precision,
                     "activation_checkpointing": config.activation_checkpointing, "dry_run": dry_run})
        notes = self.validate(config)
        if notes:
            result.details["warnings"] = notes
        if dry_run:
            result.status = "dry_run"
            return result
        env = dict(os.environ, MASTER_ADDR=config.master_addr, MASTER_PORT=str(config.master_port),
                   B200ZERO_JOB_ID=job_id)
        if self.engine == "native":  # make `python -m <pkg>.engine.launch` importable from any cwd
            pkg_parent = str(Path(__file__).resolve().parents[2])
            env["PYTHONPATH"] = pkg_parent + (os.pathsep + env["PYTHONPATH"] if env.get("PYTHONPATH") else "")
        try:
            # progress of the job (newest completed optimizer step) is visible through the registry and the jobs API
            hb_dir = self.registry.log_dir / f"{job_id}.hb"
            hb_dir.mkdir(parents=True, exist_ok=True)
            env["B200ZERO_HEARTBEAT_DIR"] = str(hb_dir)
            rec = self.registry.spawn(job_id, argv, env, config_path)
            result.status = "launched"
            result.details.update(pid=rec.pid, log_path=rec.log_path, engine=self.engine)
        except Exception as exc:  # missing binary, bad permissions, ...
            result.status = "failed"
            result.details["error"] = str(exc)
        return result

    # -- presets ----------------------------------------------------------
    @staticmethod
    def presets() -> Dict[str, DeepSpeedConfig]:
        """Model-size presets (values: ref ``:369-407``)."""
        table = {
            "7b": dict(train_micro_batch_size_per_gpu=2, gradient_accumulation_steps=16,
                       offload_param=OffloadDevice.NONE, num_gpus=4),
            "13b": dict(train_micro_batch_size_per_gpu=1, gradient_accumulation_steps=32, num_gpus=8),
            "70b": dict(train_micro_batch_size_per_gpu=1, gradient_accumulation_steps=64, bf16_enabled=True,
                        fp16_enabled=False, cpu_checkpointing=True, num_gpus=8, num_nodes=2),
        }
        return {name: DeepSpeedConfig(model_name=f"{name}_model", **kw) for name, kw in table.items()}

    @staticmethod
    def b200_presets() -> Dict[str, DeepSpeedConfig]:
        """B200-sized variants (180 GB HBM3e per GPU: no host offload needed below 70B)."""
        common = dict(bf16_enabled=True, fp16_enabled=False, communication_data_type="bf16",
                      offload_optimizer=OffloadDevice.NONE, offload_param=OffloadDevice.NONE)
        return {
            "gpt2-xl": DeepSpeedConfig(model_name="gpt2_xl", zero_stage=ZeROStage.GRADIENT_PARTITIONING,
                                       train_micro_batch_size_per_gpu=8, gradient_accumulation_steps=1,
                                       num_gpus=8, **common),
            "llama3-8b": DeepSpeedConfig(model_name="llama3_8b", train_micro_batch_size_per_gpu=1,
                                         gradient_accumulation_steps=1, num_gpus=8, **common),
            "llama3-70b": DeepSpeedConfig(model_name="llama3_70b", train_micro_batch_size_per_gpu=1,
                                          gradient_accumulation_steps=16, num_gpus=8, bf16_enabled=True,
                                          fp16_enabled=False, communication_data_type="bf16",
                                          offload_optimizer=OffloadDevice.CPU, offload_param=OffloadDevice.NONE,
                                          cpu_checkpointing=True),
        }
