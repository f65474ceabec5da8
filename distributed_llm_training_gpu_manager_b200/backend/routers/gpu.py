"""GPU fleet endpoints (routes and payloads: ref ``backend/routers/gpu.py:12-83``).

The mock fallback is reachable here: when no telemetry source answers, ``/fleet``,
``/devices`` and ``/alerts`` serve the mock fleet only if ``B200ZERO_MOCK_FALLBACK=1``;
otherwise they return the empty fleet with its "Unable to query" alert (the reference's
``except RuntimeError`` branches are dead code because its manager swallows everything,
SURVEY.md C13/C20)."""
import os

from fastapi import APIRouter, HTTPException

from ...ai_engine.gpu_manager import GPUFleetStatus, GPUManager

router = APIRouter()
manager = GPUManager()


def _fleet() -> GPUFleetStatus:
    fleet = manager.get_fleet_status()
    if fleet.total_gpus == 0 and os.environ.get("B200ZERO_MOCK_FALLBACK") == "1":
        return manager.get_mock_fleet()
    return fleet


@router.get("/fleet", response_model=GPUFleetStatus)
async def get_fleet_status():
    return _fleet()


@router.get("/fleet/mock", response_model=GPUFleetStatus)
async def get_mock_fleet():
    return manager.get_mock_fleet()


@router.get("/select")
async def select_best_gpu(required_memory_mib: int = 0, count: int = 1):
    """Best device for a new job; ``count>1`` returns an NVLink-aware placement."""
    try:
        if count > 1:
            from ...parallel.topology import discover_topology
            picked = manager.select_gpus(count, required_memory_mib,
                                         link_matrix=discover_topology().link_counts())
        else:
            one = manager.select_best_gpu(required_memory_mib)
            picked = [one] if one is not None else None
    except RuntimeError:
        pool = [d for d in manager.get_mock_fleet().devices
                if d.is_available and d.memory_free_mib >= required_memory_mib]
        picked = pool[:count] if len(pool) >= count else None
    if not picked:
        raise HTTPException(status_code=503, detail=f"No GPU available with {required_memory_mib} MiB free memory")
    gpu = picked[0]
    body = {"selected_gpu": gpu.model_dump(),
            "recommendation": f"Use GPU {gpu.index} ({gpu.name}) with {gpu.memory_free_mib} MiB free"}
    if count > 1:
        body["selected_gpus"] = [d.model_dump() for d in picked]
        body["cuda_visible_devices"] = ",".join(str(d.index) for d in picked)
    return body


@router.get("/devices/{index}")
async def get_device(index: int):
    for device in _fleet().devices:
        if device.index == index:
            return device.model_dump()
    raise HTTPException(status_code=404, detail=f"GPU {index} not found")


@router.get("/alerts")
async def get_gpu_alerts():
    fleet = _fleet()
    return {"total_alerts": len(fleet.alerts), "alerts": fleet.alerts,
            "devices_with_alerts": [{"gpu_index": d.index, "gpu_name": d.name, "health": d.health,
                                     "alerts": d.alerts} for d in fleet.devices if d.alerts]}
