"""FastAPI application: GPU fleet, training jobs, loss monitoring, NVLink topology.

Same app metadata, CORS policy, prefixes and root/health payloads as the reference
(``backend/main.py:1-39``). Differences: the package imports correctly
(``uvicorn distributed_llm_training_gpu_manager_b200.backend.main:app`` — the reference's
``from ...ai_engine`` climbs above its top-level package), the NVLink
router is mounted (it is orphaned in the reference, ``backend/routers/nvlink.py``), a job
API and a Prometheus ``/metrics`` endpoint exist, and responses are NaN-safe."""
from fastapi import FastAPI
from fastapi.middleware.cors import CORSMiddleware
from fastapi.responses import PlainTextResponse

from ..utils.json_safe import SafeJSONResponse
from .routers import gpu, jobs, monitoring, nvlink, training

app = FastAPI(
    title="MLOps Platform API",
    description="GPU fleet management, DeepSpeed distributed training, and training health monitoring",
    version="1.0.0",
    default_response_class=SafeJSONResponse,
)

app.add_middleware(CORSMiddleware, allow_origins=["*"], allow_credentials=True, allow_methods=["*"],
                   allow_headers=["*"])

app.include_router(gpu.router, prefix="/api/v1/gpu", tags=["gpu-management"])
app.include_router(training.router, prefix="/api/v1/training", tags=["distributed-training"])
app.include_router(monitoring.router, prefix="/api/v1/monitoring", tags=["loss-monitoring"])
app.include_router(nvlink.router, prefix="/api/v1/nvlink", tags=["nvlink-topology"])
app.include_router(jobs.router, prefix="/api/v1/jobs", tags=["jobs"])


@app.get("/")
def root():
    return {
        "name": "MLOps Platform API",
        "version": "1.0.0",
        "features": ["GPU Fleet Management", "DeepSpeed ZeRO-3 Launcher", "Training Loss Monitor"],
        "engine": "b200-native ZeRO (sm_100a)",
    }


@app.get("/health")
def health_check():
    return {"status": "healthy"}


@app.get("/metrics", response_class=PlainTextResponse)
def prometheus_metrics():
    from ..utils.metrics import render_prometheus
    return render_prometheus(gpu.manager, monitoring._monitors, training.launcher.registry)
