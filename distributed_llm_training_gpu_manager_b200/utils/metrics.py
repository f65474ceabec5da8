"""Prometheus text exposition for fleet, monitors and jobs (new: the reference has no
metrics export). Hand-rolled text format keeps this dependency-free and
avoids global-registry state between tests."""
from __future__ import annotations

import math
from typing import Dict


def _line(name: str, value: float, **labels: str) -> str:
    if isinstance(value, float) and not math.isfinite(value):
        value = float("nan")
    lab = ",".join(f'{k}="{v}"' for k, v in labels.items())
    return f"{name}{{{lab}}} {value}" if lab else f"{name} {value}"


def render_prometheus(gpu_manager, monitors: Dict[str, object], registry) -> str:
    out = []
    fleet = gpu_manager.get_fleet_status()
    out += ["# TYPE b200zero_gpus_total gauge", _line("b200zero_gpus_total", fleet.total_gpus),
            "# TYPE b200zero_gpus_available gauge", _line("b200zero_gpus_available", fleet.available_gpus)]
    for d in fleet.devices:
        lab = dict(gpu=str(d.index), uuid=d.uuid)
        out += [_line("b200zero_gpu_temperature_celsius", d.temperature_celsius, **lab),
                _line("b200zero_gpu_utilization_pct", d.gpu_utilization_pct, **lab),
                _line("b200zero_gpu_memory_used_mib", d.memory_used_mib, **lab),
                _line("b200zero_gpu_power_watts", d.power_draw_watts, **lab),
                _line("b200zero_gpu_sm_clock_mhz", d.sm_clock_mhz, **lab),
                _line("b200zero_gpu_throttled", int(bool(d.throttle_reasons)), **lab),
                _line("b200zero_gpu_nvlink_active_links", d.nvlink_active_links, **lab)]
    for jid, mon in monitors.items():
        st = mon.state
        out += [_line("b200zero_monitor_steps_total", st.total_steps, job=jid),
                _line("b200zero_monitor_alerts_total", st.alert_count, job=jid),
                _line("b200zero_monitor_best_loss", float(st.best_loss), job=jid),
                _line("b200zero_monitor_plateau_counter", st.plateau_counter, job=jid)]
    by_status: Dict[str, int] = {}
    for rec in registry.list():
        by_status[rec.status] = by_status.get(rec.status, 0) + 1
    for status, n in by_status.items():
        out.append(_line("b200zero_jobs", n, status=status))
    return "\n".join(out) + "\n"
