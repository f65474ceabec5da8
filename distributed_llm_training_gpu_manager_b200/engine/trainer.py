"""Training loop with in-process health monitoring, halt, checkpoint rollback and resume.

This closes the loop the reference only draws in its README (``F -->|Divergence Detected| B``,
``README.md:24``): there, ``LossSpikeMonitor.ingest`` returns alerts that nothing consumes
(SURVEY.md §3.3). Here every rank feeds its own monitor with the *globally reduced* metrics the
engine already produces each step (identical on all ranks, so every rank reaches the same
decision with no extra broadcast), and a CRITICAL alert stops the step loop in the same
iteration — the optimizer update of a non-finite step is additionally predicated off on the
device, so master weights are never corrupted.

Rollback picks the newest checkpoint older than the first bad step whose manifest verifies
(``checkpoint.find_rollback_tag``), reloads it, resets the monitor
(ref ``loss_monitor.py:273-280`` exists for exactly this) and continues. SIGTERM (spot
preemption, ``ai_engine/spot_resiliency.py``) finishes the current step, writes a
``preempt_step{N}`` checkpoint and exits 0.
"""
from __future__ import annotations

import gc
import json
import signal
import time
from dataclasses import dataclass, field
from typing import Any, Callable, Dict, List, Optional, Tuple

import torch
import torch.distributed as dist

from ..ai_engine.loss_monitor import AlertSeverity, HaltAction, HaltPolicy, LossSpikeMonitor, MonitorConfig
from . import checkpoint as ckpt
from .faults import FaultPlan


@dataclass
class TrainerConfig:
    total_steps: int = 100
    save_dir: Optional[str] = None
    save_every: int = 0
    keep_last: int = 3
    async_save: bool = True
    resume: bool = True
    log_every: int = 10
    monitor: MonitorConfig = field(default_factory=MonitorConfig)
    policy: HaltPolicy = field(default_factory=HaltPolicy)
    metrics_path: Optional[str] = None       # JSONL of per-step metrics (rank 0)
    report_url: Optional[str] = None         # POST metrics to /api/v1/monitoring/ingest/single
    eval_batches: Optional[Callable[[int], Tuple]] = None    # eval_batches(k) -> (args tuple) of the k-th validation batch
    eval_every: int = 0                      # validation every N optimizer steps (0 = never)
    eval_steps: int = 1                      # validation batches per evaluation
    gc_every: int = 0                        # > 0: Python's automatic cyclic GC is switched off for the run and a collection
    #                                          is made every N optimizer steps instead — on every rank at the same step, so
    #                                          the pauses coincide instead of each one stalling all ranks in turn


@dataclass
class TrainResult:
    status: str                               # finished | halted | preempted
    steps_run: int
    final_step: int
    rollbacks: int
    alerts: List[Dict[str, Any]]
    history: List[Dict[str, Any]]
    halt_latency_ms: List[float]


class Trainer:
    def __init__(self, engine, batches: Callable[[int], Tuple], cfg: TrainerConfig,
                 fault_plan: Optional[FaultPlan] = None):
        """``batches(global_micro_step) -> (args tuple)``: a deterministic, seekable data source
        (seekable so a rolled-back run continues from the right sample)."""
        self.engine, self.batches, self.cfg = engine, batches, cfg
        self.monitor = LossSpikeMonitor(cfg.monitor)
        self.faults = fault_plan or FaultPlan.from_env()
        self.faults.install(engine)
        self._preempted = False
        self._rollbacks = 0
        self._warn_spikes = 0
        self._alerts: List[Dict[str, Any]] = []
        self._history: List[Dict[str, Any]] = []
        self._halt_ms: List[float] = []
        try:
            signal.signal(signal.SIGTERM, self._on_sigterm)
        except ValueError:  # not the main thread
            pass

    def _on_sigterm(self, signum, frame) -> None:
        # Only a request: ranks can see the signal on different sides of a step boundary, and the emergency save is
        # collective. The engine carries the bit through the step's stats all-reduce; every rank then acts on the
        # agreed value in the same iteration (``metrics["preempt"]``).
        self._preempted = True
        self.engine.preempt_requested = True

    # ------------------------------------------------------------------
    def _maybe_resume(self) -> bool:
        c = self.cfg
        if not (c.resume and c.save_dir):
            return False
        # every rank must take the same branch: rank 0 looks at the filesystem, the rest follow
        tag = ckpt.find_rollback_tag(c.save_dir, None, deep=True) if self.engine.rank == 0 else None
        if dist.is_available() and dist.is_initialized() and self.engine.world_size > 1:
            box = [tag]
            dist.broadcast_object_list(box, src=0)
            tag = box[0]
        if tag is None:
            return False
        _, client = self.engine.load_checkpoint(c.save_dir, tag)
        if client.get("monitor"):
            self.monitor.load_state_dict(client["monitor"])
        self._log(f"resumed from {tag} at step {self.engine.global_steps}")
        return True

    def _save(self, tag: Optional[str] = None, sync: bool = False) -> None:
        c = self.cfg
        if not c.save_dir:
            return
        self.engine.save_checkpoint(c.save_dir, tag, client_state={"monitor": self.monitor.state_dict()},
                                    async_write=c.async_save and not sync)
        if c.keep_last and self.engine.rank == 0 and not (c.async_save and not sync):
            ckpt.prune_checkpoints(c.save_dir, c.keep_last)

    def _log(self, msg: str) -> None:
        if self.engine.rank == 0:
            print(f"[trainer] {msg}", flush=True)

    def _rollback(self, bad_step: int) -> bool:
        c = self.cfg
        if not c.save_dir or self._rollbacks >= c.policy.max_rollbacks:
            return False
        ckpt.wait_for_pending_writes()
        if dist.is_available() and dist.is_initialized() and self.engine.world_size > 1:
            dist.barrier()
        tag = ckpt.find_rollback_tag(c.save_dir, before_step=bad_step, deep=True) if self.engine.rank == 0 else None
        if dist.is_available() and dist.is_initialized() and self.engine.world_size > 1:
            box = [tag]
            dist.broadcast_object_list(box, src=0)
            tag = box[0]
        if tag is None:
            return False
        self.engine.load_checkpoint(c.save_dir, tag)
        self.monitor.reset()
        self._warn_spikes = 0
        self._rollbacks += 1
        self._log(f"rolled back to {tag} (fault at step {bad_step}); resuming at step {self.engine.global_steps}")
        return True

    def _validate(self) -> float:
        """Mean validation loss over ``eval_steps`` batches, averaged over the data-parallel ranks."""
        eng, c = self.engine, self.cfg
        tot = torch.zeros(1, dtype=torch.float32, device=eng.device)
        for k in range(max(c.eval_steps, 1)):
            tot += eng.eval_loss(*c.eval_batches(k)).detach().float().reshape(1)
        tot /= max(c.eval_steps, 1)
        if dist.is_available() and dist.is_initialized() and eng.world_size > 1:
            dist.all_reduce(tot)
            tot /= eng.world_size
        return float(tot)

    # ------------------------------------------------------------------
    def run(self) -> TrainResult:
        eng, c = self.engine, self.cfg
        resumed = self._maybe_resume()
        if c.save_dir and c.save_every and not resumed and eng.global_steps == 0:
            self._save(sync=True)  # step-0 anchor: a rollback target always exists
        status = "finished"
        steps_run = 0
        mf = open(c.metrics_path, "a") if (c.metrics_path and eng.rank == 0) else None
        gc_was_on = gc.isenabled()
        if c.gc_every > 0:
            gc.collect()
            gc.disable()
        try:
            return self._loop(eng, c, status, steps_run, mf)
        finally:
            if c.gc_every > 0 and gc_was_on:
                gc.enable()

    def _loop(self, eng, c, status, steps_run, mf) -> TrainResult:
        while eng.global_steps < c.total_steps:
            self.faults.at_step_start(eng)
            t_step0 = time.perf_counter()
            metrics = None
            while metrics is None:
                batch = self.batches(eng.micro_steps)
                loss = eng(*batch)
                eng.backward(loss)
                metrics = eng.step()
            t_detect0 = time.perf_counter()
            alerts = self.monitor.ingest_device_flags(
                step=metrics["step"], loss=metrics["loss"], nonfinite_grads=bool(metrics["nonfinite"]),
                grad_norm_sq=metrics["grad_norm_sq"], learning_rate=metrics["lr"])
            action = c.policy.decide(alerts, self._warn_spikes)
            self._warn_spikes = self._warn_spikes + 1 if any(
                a.alert_type == "spike" and a.severity is AlertSeverity.WARNING for a in alerts) else 0
            steps_run += 1
            rec = dict(metrics, step_ms=(time.perf_counter() - t_step0) * 1e3, action=action.value,
                       alerts=[a.alert_type for a in alerts])
            if c.eval_batches is not None and c.eval_every and metrics["step"] % c.eval_every == 0:
                rec["val_loss"] = self._validate()
            rec.pop("timers_ms", None)
            self._history.append(rec)
            if mf:
                mf.write(json.dumps(rec) + "\n")
                mf.flush()
            for a in alerts:
                self._alerts.append(a.model_dump())
            if eng.rank == 0 and (metrics["step"] % max(c.log_every, 1) == 0 or alerts):
                self._log(f"step {metrics['step']} loss {metrics['loss']:.4f} |g| {metrics['grad_norm']:.3f} "
                          f"lr {metrics['lr']:.2e} {rec['step_ms']:.0f} ms" +
                          (f" ALERTS {[a.alert_type + '/' + a.severity.value for a in alerts]} -> {action.value}"
                           if alerts else ""))
            if action is not HaltAction.CONTINUE:
                # halt decision reached: time from end-of-step metrics to here is the host part of
                # spike-to-halt; the device part is the predicated update inside the step
                self._halt_ms.append((time.perf_counter() - t_detect0) * 1e3)
                if action is HaltAction.ROLLBACK and self._rollback(metrics["step"]):
                    continue
                status = "halted"
                break
            if metrics.get("preempt", self._preempted):
                self._save(tag=f"preempt_step{eng.global_steps}", sync=True)
                status = "preempted"
                break
            if c.save_dir and c.save_every and eng.global_steps % c.save_every == 0:
                self._save()
            if c.gc_every > 0 and eng.global_steps % c.gc_every == 0:
                gc.collect()
        ckpt.wait_for_pending_writes()
        if mf:
            mf.close()
        return TrainResult(status, steps_run, eng.global_steps, self._rollbacks, self._alerts, self._history,
                           self._halt_ms)


def synthetic_batches(vocab: int, micro_batch: int, seq_len: int, device, rank: int = 0, world_size: int = 1,
                      seed: int = 1234, learnable: bool = True, pinned: bool = False) -> Callable[[int], Tuple]:
    """Deterministic synthetic token stream (no dataset on this box: BASELINE.json says synthetic).
    ``learnable``: next token = (token·5 + 7) mod vocab with 10 % noise, so the loss really falls
    and a spike is distinguishable from noise."""
    dev = torch.device(device)

    def get(micro_step: int):
        g = torch.Generator().manual_seed(seed + micro_step * 1009 + rank * 7919)
        first = torch.randint(0, vocab, (micro_batch, 1), generator=g)
        if learnable:
            toks = [first]
            for _ in range(seq_len):
                nxt = (toks[-1] * 5 + 7) % vocab
                noise = torch.rand(micro_batch, 1, generator=g) < 0.1
                toks.append(torch.where(noise, torch.randint(0, vocab, (micro_batch, 1), generator=g), nxt))
            seq = torch.cat(toks, 1)
        else:
            seq = torch.randint(0, vocab, (micro_batch, seq_len + 1), generator=g)
        ids, labels = seq[:, :-1].contiguous(), seq[:, 1:].contiguous()
        if pinned and dev.type == "cuda":
            ids, labels = ids.pin_memory(), labels.pin_memory()
        return ids.to(dev, non_blocking=True), labels.to(dev, non_blocking=True)

    return get
