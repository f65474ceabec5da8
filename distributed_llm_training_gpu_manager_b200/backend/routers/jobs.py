"""Job lifecycle endpoints — new: the reference launches and forgets (no registry, no
status/stop/logs)."""
from fastapi import APIRouter, HTTPException

from . import training

router = APIRouter()


def _registry():
    return training.launcher.registry


@router.get("")
@router.get("/")
async def list_jobs():
    jobs = _registry().list()
    return {"total": len(jobs), "jobs": [j.model_dump() for j in jobs]}


@router.get("/{job_id}")
async def job_status(job_id: str):
    rec = _registry().get(job_id)
    if rec is None:
        raise HTTPException(status_code=404, detail=f"Unknown job '{job_id}'")
    return rec.model_dump()


@router.post("/{job_id}/stop")
async def stop_job(job_id: str, grace_s: float = 30.0):
    rec = _registry().stop(job_id, grace_s=grace_s)
    if rec is None:
        raise HTTPException(status_code=404, detail=f"Unknown job '{job_id}'")
    return rec.model_dump()


@router.get("/{job_id}/logs")
async def job_logs(job_id: str, max_bytes: int = 65536):
    text = _registry().tail(job_id, max_bytes)
    if text is None:
        raise HTTPException(status_code=404, detail=f"No logs for job '{job_id}'")
    return {"job_id": job_id, "log": text}
