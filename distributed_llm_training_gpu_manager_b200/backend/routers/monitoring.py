"""Loss-monitoring endpoints (routes: ref ``backend/routers/monitoring.py:48-142``). Payloads
are NaN-safe via the app's default response class; ``/create`` exposes all monitor knobs."""
import threading
from typing import Dict, List, Optional

from fastapi import APIRouter, HTTPException
from pydantic import BaseModel

from ...ai_engine.loss_monitor import HaltPolicy, LossSpikeMonitor, MonitorConfig, TrainingMetrics

router = APIRouter()
_monitors: Dict[str, LossSpikeMonitor] = {}
_lock = threading.Lock()
_policy = HaltPolicy()


def _get_or_create_monitor(job_id: str, config: Optional[MonitorConfig] = None) -> LossSpikeMonitor:
    with _lock:
        mon = _monitors.get(job_id)
        if mon is None:
            mon = _monitors[job_id] = LossSpikeMonitor(config)
        return mon


def _require(job_id: str) -> LossSpikeMonitor:
    mon = _monitors.get(job_id)
    if mon is None:
        raise HTTPException(status_code=404, detail=f"No monitor for job '{job_id}'")
    return mon


class MonitorCreateRequest(BaseModel):
    job_id: str
    window_size: int = 100
    spike_sigma_threshold: float = 3.0
    divergence_threshold: float = 1e6
    plateau_patience: int = 500
    gradient_norm_threshold: float = 100.0
    # knobs the reference's request model left out
    plateau_min_delta: float = 1e-4
    lr_spike_ratio: float = 10.0
    max_alerts_per_type: int = 50
    cooldown_steps: int = 20


class MetricsIngestRequest(BaseModel):
    job_id: str
    metrics: List[TrainingMetrics]


class SingleMetricRequest(BaseModel):
    job_id: str
    step: int
    loss: float
    learning_rate: float = 0.0
    gradient_norm: float = 0.0
    throughput_samples_per_sec: float = 0.0
    gpu_memory_used_mib: int = 0


@router.post("/create")
async def create_monitor(request: MonitorCreateRequest):
    config = MonitorConfig(**request.model_dump(exclude={"job_id"}))
    existed = request.job_id in _monitors
    _get_or_create_monitor(request.job_id, config)
    return {"job_id": request.job_id, "status": "exists" if existed else "created", "config": config.model_dump()}


@router.post("/ingest")
async def ingest_metrics(request: MetricsIngestRequest):
    monitor = _get_or_create_monitor(request.job_id)
    alerts = [a for m in request.metrics for a in monitor.ingest(m)]
    return {"job_id": request.job_id, "ingested": len(request.metrics), "alerts_triggered": len(alerts),
            "alerts": [a.model_dump() for a in alerts], "action": _policy.decide(alerts).value}


@router.post("/ingest/single")
async def ingest_single_metric(request: SingleMetricRequest):
    monitor = _get_or_create_monitor(request.job_id)
    alerts = monitor.ingest(TrainingMetrics(**request.model_dump(exclude={"job_id"})))
    return {"job_id": request.job_id, "step": request.step, "alerts": [a.model_dump() for a in alerts],
            "healthy": not alerts, "action": _policy.decide(alerts).value}


@router.get("/summary/{job_id}")
async def get_monitor_summary(job_id: str):
    return _require(job_id).get_summary()


@router.get("/loss-curve/{job_id}")
async def get_loss_curve(job_id: str):
    return _require(job_id).get_loss_curve()


@router.delete("/reset/{job_id}")
async def reset_monitor(job_id: str):
    _require(job_id).reset()
    return {"job_id": job_id, "status": "reset"}


@router.get("/jobs")
async def list_monitored_jobs():
    return {"active_jobs": len(_monitors),
            "jobs": [{"job_id": jid, "total_steps": m.state.total_steps, "total_alerts": m.state.alert_count,
                      "best_loss": m.state.best_loss} for jid, m in _monitors.items()]}
