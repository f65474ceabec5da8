"""Training-health anomaly detection: NaN/Inf divergence, loss spikes, plateaus,
gradient-norm explosions and learning-rate jumps — plus the *decision* layer the
reference lacks (halt / rollback policy) and O(1) rolling statistics.

API kept from the reference (``ai_engine/loss_monitor.py:23-310``): ``AlertSeverity``,
``SpikeAlert``, ``TrainingMetrics``, ``MonitorConfig``, ``MonitorState``,
``LossSpikeMonitor.{ingest,get_summary,get_loss_curve,reset}``. Detection order,
thresholds, cooldown rules and the quirks documented in SURVEY.md Appendix A.3 are
preserved in the default (``compat``) mode so existing dashboards see identical
alerts.

Re-designed:

* Rolling mean/σ come from a sliding-window Welford accumulator (O(1) per step with a
  periodic exact refresh) instead of ``statistics.mean/stdev`` over the deque
  (O(window) per step, ref ``:153-156``).
* Histories are bounded (``history_limit``); ``max_alerts_per_type`` is honoured (the
  reference defines it but never reads it, ref ``:65``).
* ``MonitorState`` really is persisted: :meth:`state_dict` / :meth:`load_state_dict`
  travel inside checkpoints so a resumed run keeps its window.
* :class:`HaltPolicy` turns alerts into ``continue | halt | rollback`` decisions; the
  training engine consults it every step (spike-to-halt path).
* :meth:`ingest_device_flags` accepts the non-finite flag and ‖g‖² partial produced
  inside the fused reduce-scatter kernel, so no extra pass over gradients is needed.
* Summaries are JSON-safe (no ``inf``/``nan`` leaks → no HTTP 500).
"""
from __future__ import annotations

import math
from collections import deque
from datetime import datetime, timezone
from enum import Enum
from typing import Any, Deque, Dict, List, Optional

from pydantic import BaseModel, Field


def _now() -> datetime:
    return datetime.now(timezone.utc).replace(tzinfo=None)


class AlertSeverity(str, Enum):
    INFO = "info"
    WARNING = "warning"
    CRITICAL = "critical"


class SpikeAlert(BaseModel):
    alert_id: str
    severity: AlertSeverity
    alert_type: str = Field(..., description="spike | divergence | plateau | gradient_explosion | lr_anomaly")
    step: int
    message: str
    metric_value: float
    threshold: float
    rolling_mean: float
    rolling_std: float
    remediation: List[str] = Field(default_factory=list)
    timestamp: str = Field(default_factory=lambda: _now().isoformat())


class TrainingMetrics(BaseModel):
    step: int
    loss: float
    learning_rate: float = 0.0
    gradient_norm: float = 0.0
    throughput_samples_per_sec: float = 0.0
    gpu_memory_used_mib: int = 0
    epoch: Optional[int] = None
    timestamp: str = Field(default_factory=lambda: _now().isoformat())


class MonitorConfig(BaseModel):
    window_size: int = Field(default=100, description="Rolling window for statistics")
    spike_sigma_threshold: float = Field(default=3.0, description="Std devs for spike detection")
    divergence_threshold: float = Field(default=1e6, description="Loss considered diverged")
    plateau_patience: int = Field(default=500, description="Steps without improvement for plateau")
    plateau_min_delta: float = Field(default=1e-4, description="Minimum improvement to reset plateau")
    gradient_norm_threshold: float = Field(default=100.0, description="Max gradient norm before alert")
    lr_spike_ratio: float = Field(default=10.0, description="LR jump ratio for alert")
    max_alerts_per_type: int = Field(default=50, description="Max alerts to store per type")
    cooldown_steps: int = Field(default=20, description="Min steps between same-type alerts")
    # --- additions (defaults keep reference behaviour) ---
    history_limit: int = Field(default=100_000, description="Max per-step metric records kept in memory")
    exclude_spikes_from_window: bool = Field(default=False, description="robust mode: spikes do not pollute the window")
    nan_gradient_is_divergence: bool = Field(default=False, description="treat a NaN grad norm as divergence")


class MonitorState(BaseModel):
    total_steps: int = 0
    best_loss: float = float("inf")
    best_loss_step: int = 0
    plateau_counter: int = 0
    alert_count: int = 0
    last_alert_step: Dict[str, int] = Field(default_factory=dict)


class RollingStats:
    """Sliding-window mean / sample standard deviation in O(1) per update."""

    def __init__(self, window: int):
        self.window = max(int(window), 1)
        self.values: Deque[float] = deque()
        self.mean = 0.0
        self.m2 = 0.0
        self._since_refresh = 0

    def __len__(self) -> int:
        return len(self.values)

    def push(self, x: float) -> None:
        x = float(x)
        n = len(self.values)
        if n < self.window:
            self.values.append(x)
            n += 1
            d = x - self.mean
            self.mean += d / n
            self.m2 += d * (x - self.mean)
        else:
            old = self.values.popleft()
            self.values.append(x)
            d = x - old
            new_mean = self.mean + d / n
            self.m2 += d * (x - new_mean + old - self.mean)
            self.mean = new_mean
        self._since_refresh += 1
        if self._since_refresh >= self.window or self.m2 < 0.0 or not math.isfinite(self.m2):
            self.refresh()

    def refresh(self) -> None:
        n = len(self.values)
        self._since_refresh = 0
        if n == 0:
            self.mean, self.m2 = 0.0, 0.0
            return
        self.mean = math.fsum(self.values) / n
        self.m2 = math.fsum((v - self.mean) ** 2 for v in self.values)

    @property
    def std(self) -> float:
        n = len(self.values)
        return math.sqrt(max(self.m2, 0.0) / (n - 1)) if n > 1 else 0.0

    def clear(self) -> None:
        self.values.clear()
        self.mean = self.m2 = 0.0
        self._since_refresh = 0


_REMEDIATION = {
    "nonfinite": ["Immediately reduce learning rate by 10x", "Check for data corruption in current batch",
                  "Enable gradient clipping if not already active",
                  "Restore from last checkpoint and retry with lower LR"],
    "diverged": ["Reduce learning rate by 5-10x", "Increase gradient clipping value",
                 "Check for numerical instability in model architecture"],
    "spike": ["Monitor next 10 steps for recovery", "If persistent, reduce learning rate",
              "Check system logs for GPU errors or data issues", "Consider reverting to previous checkpoint"],
    "plateau": ["Consider reducing learning rate (cosine annealing or step decay)",
                "Try increasing model capacity or data augmentation",
                "Evaluate if training has converged (check validation metrics)",
                "Experiment with different optimizer (switch AdamW ↔ LAMB)"],
    "gradient_explosion": ["Reduce gradient clipping value", "Lower learning rate",
                           "Check for exploding activations in specific layers"],
    "lr_anomaly": ["Verify LR scheduler configuration", "Check for warmup restart if using cosine with restarts"],
}


class LossSpikeMonitor:
    """Per-job stateful detector. Feed one :class:`TrainingMetrics` per optimizer step."""

    def __init__(self, config: Optional[MonitorConfig] = None):
        self.config = config or MonitorConfig()
        self.state = MonitorState()
        w = self.config.window_size
        self._loss = RollingStats(w)
        self._grad = RollingStats(w)
        self._lr = RollingStats(w)
        self._all_metrics: Deque[TrainingMetrics] = deque(maxlen=self.config.history_limit)
        self._all_alerts: List[SpikeAlert] = []
        self._stored_per_type: Dict[str, int] = {}

    # kept for callers that poked the reference's private deques
    @property
    def _loss_history(self) -> Deque[float]:
        return self._loss.values

    # ------------------------------------------------------------------ ingest
    def ingest(self, metrics: TrainingMetrics) -> List[SpikeAlert]:
        cfg, st = self.config, self.state
        fired: List[SpikeAlert] = []
        st.total_steps += 1
        self._all_metrics.append(metrics)
        loss, step = metrics.loss, metrics.step

        # (1) non-finite loss: report and stop processing this sample (window untouched)
        if math.isnan(loss) or math.isinf(loss):
            return [self._alert("divergence", AlertSeverity.CRITICAL, step, loss, 0, 0, 0,
                                f"Loss diverged at step {step}: {loss}", _REMEDIATION["nonfinite"])]

        if cfg.nan_gradient_is_divergence and math.isnan(metrics.gradient_norm):
            fired.append(self._alert("divergence", AlertSeverity.CRITICAL, step, metrics.gradient_norm, 0, 0, 0,
                                     f"Gradient norm is NaN at step {step}", _REMEDIATION["nonfinite"]))

        # (1b) finite but absurd
        if loss > cfg.divergence_threshold:
            fired.append(self._alert("divergence", AlertSeverity.CRITICAL, step, loss, cfg.divergence_threshold, 0, 0,
                                     f"Loss {loss:.4e} exceeds divergence threshold {cfg.divergence_threshold:.4e}",
                                     _REMEDIATION["diverged"]))

        # (2) spike against the rolling window
        is_spike = False
        if len(self._loss) >= 10:
            mean, std = self._loss.mean, self._loss.std
            floor_std = max(std, 1e-8)
            limit = mean + cfg.spike_sigma_threshold * floor_std
            if loss > limit:
                is_spike = True
                if self._cooled("spike", step):
                    sev = AlertSeverity.CRITICAL if loss > mean + 5 * floor_std else AlertSeverity.WARNING
                    fired.append(self._alert(
                        "spike", sev, step, loss, limit, mean, std,
                        f"Loss spike at step {step}: {loss:.6f} (mean={mean:.6f}, std={std:.6f}, threshold={limit:.6f})",
                        _REMEDIATION["spike"]))

        # (3) plateau bookkeeping
        if loss < st.best_loss - cfg.plateau_min_delta:
            st.best_loss, st.best_loss_step, st.plateau_counter = loss, step, 0
        else:
            st.plateau_counter += 1
        if st.plateau_counter >= cfg.plateau_patience and self._cooled("plateau", step):
            fired.append(self._alert(
                "plateau", AlertSeverity.WARNING, step, loss, st.best_loss, float(st.plateau_counter), 0,
                f"Training plateau: no improvement for {st.plateau_counter} steps "
                f"(best={st.best_loss:.6f} at step {st.best_loss_step})", _REMEDIATION["plateau"]))

        # (4) gradient norm (a NaN norm fails `> 0` and is skipped, as in the reference)
        gn = metrics.gradient_norm
        if gn > 0:
            if gn > cfg.gradient_norm_threshold and self._cooled("gradient_explosion", step):
                fired.append(self._alert(
                    "gradient_explosion", AlertSeverity.WARNING, step, gn, cfg.gradient_norm_threshold,
                    self._grad.mean if len(self._grad) else 0, 0,
                    f"Gradient norm explosion at step {step}: {gn:.4f} > {cfg.gradient_norm_threshold}",
                    _REMEDIATION["gradient_explosion"]))
            if math.isfinite(gn):
                self._grad.push(gn)

        # (5) learning-rate jump
        lr = metrics.learning_rate
        if lr > 0:
            if len(self._lr) >= 5:
                avg = self._lr.mean
                if lr > avg * cfg.lr_spike_ratio and self._cooled("lr_anomaly", step):
                    fired.append(self._alert(
                        "lr_anomaly", AlertSeverity.WARNING, step, lr, avg * cfg.lr_spike_ratio, avg, 0,
                        f"Learning rate jump at step {step}: {lr:.2e} (avg={avg:.2e})", _REMEDIATION["lr_anomaly"]))
            self._lr.push(lr)

        if not (cfg.exclude_spikes_from_window and is_spike):
            self._loss.push(loss)

        for a in fired:
            if self._stored_per_type.get(a.alert_type, 0) < cfg.max_alerts_per_type:
                self._all_alerts.append(a)
                self._stored_per_type[a.alert_type] = self._stored_per_type.get(a.alert_type, 0) + 1
        st.alert_count += len(fired)
        return fired

    def ingest_device_flags(self, step: int, loss: float, nonfinite_grads: bool, grad_norm_sq: float,
                            learning_rate: float = 0.0, **extra: Any) -> List[SpikeAlert]:
        """Entry point for the fused reduce-scatter kernel's by-products: the OR-reduced
        isfinite flag and the global ‖g‖². A non-finite gradient is
        reported as divergence even when the loss itself still looks finite."""
        gn = math.sqrt(grad_norm_sq) if (grad_norm_sq >= 0 and math.isfinite(grad_norm_sq)) else float("nan")
        alerts: List[SpikeAlert] = []
        if nonfinite_grads:
            alerts.append(self._alert("divergence", AlertSeverity.CRITICAL, step, float("nan"), 0, 0, 0,
                                      f"Non-finite gradients detected in-kernel at step {step}",
                                      _REMEDIATION["nonfinite"]))
            self.state.alert_count += 1
            gn = 0.0 if not math.isfinite(gn) else gn
        alerts += self.ingest(TrainingMetrics(step=step, loss=loss, learning_rate=learning_rate,
                                              gradient_norm=gn if math.isfinite(gn) else 0.0, **extra))
        return alerts

    # ------------------------------------------------------------------ queries
    def get_summary(self) -> Dict[str, Any]:
        vals = self._loss.values
        return {
            "total_steps": self.state.total_steps,
            "best_loss": self.state.best_loss,
            "best_loss_step": self.state.best_loss_step,
            "current_loss": vals[-1] if vals else None,
            "rolling_mean": self._loss.mean if vals else None,
            "rolling_std": self._loss.std if len(vals) > 1 else None,
            "plateau_counter": self.state.plateau_counter,
            "total_alerts": self.state.alert_count,
            "recent_alerts": [a.model_dump() for a in self._all_alerts[-10:]],
            "alert_counts_by_type": dict(self._stored_per_type),
        }

    def get_loss_curve(self) -> Dict[str, Any]:
        ms = list(self._all_metrics)
        return {
            "steps": [m.step for m in ms],
            "losses": [m.loss for m in ms],
            "learning_rates": [m.learning_rate for m in ms],
            "gradient_norms": [m.gradient_norm for m in ms],
            "spike_steps": [a.step for a in self._all_alerts if a.alert_type == "spike"],
        }

    def reset(self) -> None:
        """Forget everything (call after restoring a checkpoint; ref ``:273-280``)."""
        for r in (self._loss, self._grad, self._lr):
            r.clear()
        self._all_metrics.clear()
        self._all_alerts.clear()
        self._stored_per_type.clear()
        self.state = MonitorState()

    # ------------------------------------------------------------------ persistence
    def state_dict(self) -> Dict[str, Any]:
        return {"config": self.config.model_dump(), "state": self.state.model_dump(),
                "loss_window": list(self._loss.values), "grad_window": list(self._grad.values),
                "lr_window": list(self._lr.values)}

    def load_state_dict(self, sd: Dict[str, Any]) -> None:
        self.state = MonitorState(**sd.get("state", {}))
        for roll, key in ((self._loss, "loss_window"), (self._grad, "grad_window"), (self._lr, "lr_window")):
            roll.clear()
            for v in sd.get(key, []):
                roll.push(v)

    # ------------------------------------------------------------------ internals
    def _alert(self, kind: str, sev: AlertSeverity, step: int, value: float, threshold: float, mean: float,
               std: float, message: str, remediation: List[str]) -> SpikeAlert:
        self.state.last_alert_step[kind] = step
        return SpikeAlert(alert_id=f"{kind}_{step}_{_now().strftime('%H%M%S')}", severity=sev, alert_type=kind,
                          step=step, message=message, metric_value=value, threshold=threshold, rolling_mean=mean,
                          rolling_std=std, remediation=list(remediation))

    def _cooled(self, kind: str, step: int) -> bool:
        return step - self.state.last_alert_step.get(kind, -999) >= self.config.cooldown_steps

    # reference-compatible private names
    _can_alert = _cooled

    def _count_alerts_by_type(self) -> Dict[str, int]:
        return dict(self._stored_per_type)


# ---------------------------------------------------------------------------
# decision layer (new)
# ---------------------------------------------------------------------------
class HaltAction(str, Enum):
    CONTINUE = "continue"
    HALT = "halt"          # stop the job, keep state
    ROLLBACK = "rollback"  # stop, restore newest verified checkpoint older than the fault, resume


class HaltPolicy(BaseModel):
    """Maps alerts to an action. Defaults: any divergence or CRITICAL spike rolls back;
    ``spike_patience`` consecutive WARNING spikes roll back; the rest continue."""

    rollback_on_divergence: bool = True
    rollback_on_critical_spike: bool = True
    spike_patience: int = 3
    halt_on_gradient_explosion: bool = False
    max_rollbacks: int = 3

    def decide(self, alerts: List[SpikeAlert], consecutive_warning_spikes: int = 0) -> HaltAction:
        action = HaltAction.CONTINUE
        for a in alerts:
            if a.alert_type == "divergence" and self.rollback_on_divergence:
                return HaltAction.ROLLBACK
            if a.alert_type == "spike":
                if a.severity is AlertSeverity.CRITICAL and self.rollback_on_critical_spike:
                    return HaltAction.ROLLBACK
                if consecutive_warning_spikes + 1 >= self.spike_patience:
                    return HaltAction.ROLLBACK
            if a.alert_type == "gradient_explosion" and self.halt_on_gradient_explosion:
                action = HaltAction.HALT
        return action
