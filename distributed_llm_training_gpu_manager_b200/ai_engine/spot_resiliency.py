"""Spot / preemptible-instance resiliency: detect an impending preemption, trigger an
emergency checkpoint in every rank, and (optionally) relaunch from it.

The reference class (``ai_engine/spot_resiliency.py:5-49``) is a stub whose poller can
never fire and whose "checkpoint" is three ``print`` calls. The name, constructor and
coroutine entry points are kept (``SpotInstanceResiliencyManager(check_interval_sec)``,
``monitor_preemption_notices(launcher_ref)``, ``_simulate_interruption``,
``_execute_emergency_checkpoint``), but each now does real work:

* **Detection** — pluggable probes polled every ``check_interval_sec``: AWS IMDSv2
  ``spot/instance-action``, GCP ``instance/preempted``, a sentinel file, an environment
  variable, and ``SIGTERM`` delivered to this process. Endpoints are injectable so tests
  run a fake metadata server on 127.0.0.1.
* **Action** — the job's process group receives ``SIGTERM``; native workers install a
  handler (``engine/trainer.py``) that finishes the current step, writes a checkpoint
  tagged ``preempt_step{N}`` through the pinned async writer, and exits 0. The manager
  waits up to ``grace_seconds`` (the cloud's ~2 min warning, ref ``:9,35``).
* **Recovery** — ``on_checkpointed`` callback (the supervisor relaunches with ``--resume``).
"""
from __future__ import annotations

import asyncio
import json
import os
import signal
import time
import urllib.error
import urllib.request
from dataclasses import dataclass, field
from typing import Any, Awaitable, Callable, Dict, List, Optional

AWS_TOKEN_URL = "http://169.254.169.254/latest/api/token"
AWS_ACTION_URL = "http://169.254.169.254/latest/meta-data/spot/instance-action"
GCP_PREEMPTED_URL = "http://metadata.google.internal/computeMetadata/v1/instance/preempted"


@dataclass
class PreemptionNotice:
    source: str
    detail: str = ""
    deadline_s: float = 120.0
    detected_at: float = field(default_factory=time.time)


def _http_get(url: str, headers: Optional[Dict[str, str]] = None, method: str = "GET",
              timeout: float = 1.0) -> Optional[str]:
    req = urllib.request.Request(url, headers=headers or {}, method=method)
    try:
        with urllib.request.urlopen(req, timeout=timeout) as r:
            return r.read().decode(errors="replace")
    except (urllib.error.URLError, urllib.error.HTTPError, OSError, ValueError):
        return None


class AwsSpotProbe:
    """IMDSv2: PUT a session token, then GET ``spot/instance-action`` (404 until a notice exists)."""

    def __init__(self, action_url: str = AWS_ACTION_URL, token_url: Optional[str] = AWS_TOKEN_URL):
        self.action_url, self.token_url = action_url, token_url

    def __call__(self) -> Optional[PreemptionNotice]:
        headers = {}
        if self.token_url:
            tok = _http_get(self.token_url, {"X-aws-ec2-metadata-token-ttl-seconds": "60"}, method="PUT")
            if tok:
                headers["X-aws-ec2-metadata-token"] = tok.strip()
        body = _http_get(self.action_url, headers)
        if not body:
            return None
        try:
            info = json.loads(body)
        except ValueError:
            info = {"raw": body}
        return PreemptionNotice("aws", json.dumps(info))


class GcpPreemptProbe:
    def __init__(self, url: str = GCP_PREEMPTED_URL):
        self.url = url

    def __call__(self) -> Optional[PreemptionNotice]:
        body = _http_get(self.url, {"Metadata-Flavor": "Google"})
        if body and body.strip().upper() == "TRUE":
            return PreemptionNotice("gcp", "preempted=TRUE", deadline_s=30.0)
        return None


class FileProbe:
    """Fires when ``path`` exists — the fault-injection hook used by tests and drills."""

    def __init__(self, path: str):
        self.path = path

    def __call__(self) -> Optional[PreemptionNotice]:
        return PreemptionNotice("file", self.path) if os.path.exists(self.path) else None


class EnvProbe:
    def __init__(self, var: str = "B200ZERO_INJECT_PREEMPTION"):
        self.var = var

    def __call__(self) -> Optional[PreemptionNotice]:
        return PreemptionNotice("env", self.var) if os.environ.get(self.var) == "1" else None


Probe = Callable[[], Optional[PreemptionNotice]]


def default_probes() -> List[Probe]:
    probes: List[Probe] = [EnvProbe()]
    sentinel = os.environ.get("B200ZERO_PREEMPT_FILE")
    if sentinel:
        probes.append(FileProbe(sentinel))
    cloud = os.environ.get("B200ZERO_CLOUD", "").lower()
    if cloud in ("aws", "auto"):
        probes.append(AwsSpotProbe())
    if cloud in ("gcp", "auto"):
        probes.append(GcpPreemptProbe())
    return probes


class SpotInstanceResiliencyManager:
    """Polls for preemption notices and drives the emergency-checkpoint protocol."""

    def __init__(self, check_interval_sec: float = 5, probes: Optional[List[Probe]] = None,
                 grace_seconds: float = 110.0,
                 on_checkpointed: Optional[Callable[[PreemptionNotice, Dict[str, Any]], Any]] = None,
                 install_sigterm_handler: bool = False):
        self.check_interval_sec = check_interval_sec
        self.probes: List[Probe] = list(probes) if probes is not None else default_probes()
        self.grace_seconds = grace_seconds
        self.on_checkpointed = on_checkpointed
        self.is_running = False
        self.last_notice: Optional[PreemptionNotice] = None
        self.last_result: Optional[Dict[str, Any]] = None
        self._sigterm_seen = False
        if install_sigterm_handler:
            signal.signal(signal.SIGTERM, self._on_sigterm)

    def _on_sigterm(self, signum, frame) -> None:  # pragma: no cover - signal context
        self._sigterm_seen = True

    # -- detection ----------------------------------------------------------
    def check_once(self) -> Optional[PreemptionNotice]:
        if self._sigterm_seen:
            return PreemptionNotice("sigterm", "SIGTERM delivered to orchestrator", deadline_s=self.grace_seconds)
        if self._simulate_interruption():
            return PreemptionNotice("simulated", "_simulate_interruption() returned True")
        for probe in self.probes:
            try:
                notice = probe()
            except Exception:
                notice = None
            if notice is not None:
                return notice
        return None

    def _simulate_interruption(self) -> bool:
        """Override (or monkeypatch) to inject a preemption; off by default (ref ``:39-41``)."""
        return False

    async def monitor_preemption_notices(self, deepspeed_launcher_ref=None, job_id: Optional[str] = None):
        """Poll until a notice arrives (or :meth:`stop`), then run the emergency protocol.
        ``deepspeed_launcher_ref`` is a :class:`DeepSpeedLauncher` (its registry owns the
        job's process group) — the reference accepted but never used it (ref ``:17,43``)."""
        self.is_running = True
        try:
            while self.is_running:
                notice = await asyncio.get_running_loop().run_in_executor(None, self.check_once)
                if notice is not None:
                    self.last_notice = notice
                    self.last_result = await self._execute_emergency_checkpoint(deepspeed_launcher_ref, job_id, notice)
                    return self.last_result
                await asyncio.sleep(self.check_interval_sec)
        finally:
            self.is_running = False
        return None

    def stop(self) -> None:
        self.is_running = False

    # -- action -------------------------------------------------------------
    async def _execute_emergency_checkpoint(self, launcher, job_id: Optional[str] = None,
                                            notice: Optional[PreemptionNotice] = None) -> Dict[str, Any]:
        """SIGTERM every running job of ``launcher`` (or just ``job_id``), wait for the
        workers' emergency checkpoints, report what happened."""
        notice = notice or PreemptionNotice("manual")
        registry = getattr(launcher, "registry", None)
        outcome: Dict[str, Any] = {"notice": notice.source, "detail": notice.detail, "jobs": {}}
        if registry is None:
            outcome["error"] = "no job registry attached to launcher"
            return outcome
        targets = [job_id] if job_id else [r.job_id for r in registry.list() if r.status == "running"]
        grace = min(self.grace_seconds, notice.deadline_s)
        loop = asyncio.get_running_loop()
        stops: List[Awaitable] = [loop.run_in_executor(None, registry.stop, jid, signal.SIGTERM, grace)
                                  for jid in targets]
        for jid, rec in zip(targets, await asyncio.gather(*stops, return_exceptions=True)):
            if isinstance(rec, Exception) or rec is None:
                outcome["jobs"][jid] = {"status": "unknown", "error": str(rec)}
            else:
                outcome["jobs"][jid] = {"status": rec.status, "returncode": rec.returncode,
                                        "checkpointed": rec.returncode == 0}
        if self.on_checkpointed is not None:
            res = self.on_checkpointed(notice, outcome)
            if asyncio.iscoroutine(res):
                await res
        return outcome
