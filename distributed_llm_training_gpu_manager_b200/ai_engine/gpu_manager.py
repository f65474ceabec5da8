"""GPU fleet telemetry, health classification and (topology-aware) device selection.

API kept from the reference (``ai_engine/gpu_manager.py:20-431``): ``GPUHealthStatus``,
``GPUProcess``, ``GPUDevice`` (+ ``is_available``), ``GPUFleetStatus``, and
``GPUManager.{parse_xml, parse_csv, get_fleet_status, select_best_gpu, get_mock_fleet}``
including the string-injection seams (``xml_str`` / ``csv_str``) that make the parsers
testable without a GPU.

Re-designed for B200 nodes:

* **NVML first.** ``GPUManager(backend="auto")`` polls NVML through ``pynvml``
  (no fork/exec per query; the reference shells out to ``nvidia-smi`` with a 30 s
  timeout, ref ``:103-117``). It adds what the reference cannot see: enforced power
  limit, clock-event (throttle) reasons, ECC and retired-page counters, per-link
  NVLink state. ``nvidia-smi`` XML → CSV remain as fallbacks, then an empty fleet.
* Thresholds are per-instance (:class:`HealthThresholds`), not class constants.
* ``select_gpus(n)`` places a multi-GPU job: healthy, unthrottled devices, preferring
  an NVLink-connected clique with the most free memory (the reference only ever picks
  one device, ref ``:323-346``). ``select_best_gpu`` keeps its signature.
* A background :class:`FleetSampler` keeps a ring buffer of snapshots for trend
  queries and Prometheus export.
* Fixed: ``current_power_limit`` is honoured in XML (modern drivers; the reference
  reads only ``power_limit`` so XML power alerts never fire — SURVEY.md C10), and the
  ``Element or Element`` truthiness bug.
"""
from __future__ import annotations

import re
import subprocess
import threading
import xml.etree.ElementTree as ET
from collections import deque
from datetime import datetime, timezone
from enum import Enum
from typing import Any, Deque, Dict, List, Optional, Sequence, Tuple

from pydantic import BaseModel, Field


def _now_iso() -> str:
    return datetime.now(timezone.utc).replace(tzinfo=None).isoformat()


class GPUHealthStatus(str, Enum):
    HEALTHY = "healthy"
    WARNING = "warning"
    CRITICAL = "critical"
    UNREACHABLE = "unreachable"


class GPUProcess(BaseModel):
    pid: int
    name: str = ""
    used_memory_mib: int = 0
    gpu_instance_id: Optional[str] = None


class GPUDevice(BaseModel):
    index: int
    name: str = Field(default="Unknown GPU")
    uuid: str = ""
    temperature_celsius: int = 0
    gpu_utilization_pct: float = 0.0
    memory_used_mib: int = 0
    memory_total_mib: int = 0
    memory_free_mib: int = 0
    memory_utilization_pct: float = 0.0
    power_draw_watts: float = 0.0
    power_limit_watts: float = 0.0
    fan_speed_pct: int = 0
    driver_version: str = ""
    cuda_version: str = ""
    compute_mode: str = "Default"
    pci_bus_id: str = ""
    processes: List[GPUProcess] = Field(default_factory=list)
    health: GPUHealthStatus = GPUHealthStatus.HEALTHY
    alerts: List[str] = Field(default_factory=list)
    # --- NVML-only extras (absent from nvidia-smi CSV; defaults keep old payloads valid) ---
    sm_clock_mhz: int = 0
    sm_clock_max_mhz: int = 0
    throttle_reasons: List[str] = Field(default_factory=list)
    ecc_errors_uncorrected: int = 0
    retired_pages_pending: bool = False
    nvlink_active_links: int = 0
    nvlink_total_links: int = 0
    numa_node: Optional[int] = None

    @property
    def is_available(self) -> bool:
        """Schedulable: <80 % memory used, <90 % busy, not critical (ref ``:57-62``)."""
        return (self.memory_utilization_pct < 80 and self.gpu_utilization_pct < 90
                and self.health != GPUHealthStatus.CRITICAL)


class GPUFleetStatus(BaseModel):
    timestamp: str = Field(default_factory=_now_iso)
    total_gpus: int = 0
    healthy_gpus: int = 0
    available_gpus: int = 0
    total_memory_mib: int = 0
    used_memory_mib: int = 0
    avg_utilization_pct: float = 0.0
    avg_temperature_celsius: float = 0.0
    total_power_watts: float = 0.0
    devices: List[GPUDevice] = Field(default_factory=list)
    alerts: List[str] = Field(default_factory=list)
    source: str = "none"  # nvml | xml | csv | mock | none


class HealthThresholds(BaseModel):
    temp_warning: int = 80
    temp_critical: int = 90
    mem_warning: float = 85
    mem_critical: float = 95
    util_warning: float = 95
    power_warning: float = 0.9


_HW_THROTTLE = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "hw_power_brake_slowdown")
_CSV_FIELDS = ("index,name,uuid,temperature.gpu,utilization.gpu,utilization.memory,memory.used,memory.total,"
               "memory.free,power.draw,power.limit,fan.speed")
_NUM = re.compile(r"\d+")
_FLT = re.compile(r"[\d.]+")


def _first_int(text: str) -> int:
    m = _NUM.search(text or "")
    return int(m.group()) if m else 0


def _first_float(text: str) -> float:
    m = _FLT.search(text or "")
    try:
        return float(m.group()) if m else 0.0
    except ValueError:
        return 0.0


def _child_text(node: Optional[ET.Element], tag: str, default: str = "") -> str:
    if node is None:
        return default
    el = node.find(tag)
    return el.text.strip() if (el is not None and el.text) else default


class GPUManager:
    """Fleet monitor. ``backend``: ``auto`` (NVML → XML → CSV), ``nvml``, ``smi``."""

    # class-level defaults kept so code that read them off the class still works (ref :92-98)
    TEMP_WARNING, TEMP_CRITICAL = 80, 90
    MEM_WARNING, MEM_CRITICAL = 85, 95
    UTIL_WARNING = 95
    POWER_WARNING = 0.9

    def __init__(self, nvidia_smi_path: str = "nvidia-smi", backend: str = "auto",
                 thresholds: Optional[HealthThresholds] = None):
        self.nvidia_smi_path = nvidia_smi_path
        self.backend = backend
        self.thresholds = thresholds or HealthThresholds(
            temp_warning=self.TEMP_WARNING, temp_critical=self.TEMP_CRITICAL, mem_warning=self.MEM_WARNING,
            mem_critical=self.MEM_CRITICAL, util_warning=self.UTIL_WARNING, power_warning=self.POWER_WARNING)
        self._nvml = None
        self._nvml_lock = threading.Lock()

    # ------------------------------------------------------------------ nvidia-smi seam
    def _run_nvidia_smi(self, args: List[str]) -> str:
        try:
            res = subprocess.run([self.nvidia_smi_path, *args], capture_output=True, text=True, timeout=30)
        except FileNotFoundError:
            raise RuntimeError(f"nvidia-smi not found at '{self.nvidia_smi_path}'. Ensure NVIDIA drivers are installed.")
        except subprocess.TimeoutExpired:
            raise RuntimeError("nvidia-smi timed out after 30 s")
        if res.returncode != 0:
            raise RuntimeError(f"nvidia-smi failed: {res.stderr}")
        return res.stdout

    def parse_xml(self, xml_str: Optional[str] = None) -> List[GPUDevice]:
        """Parse ``nvidia-smi -q -x`` output (live query when ``xml_str`` is None)."""
        root = ET.fromstring(xml_str if xml_str is not None else self._run_nvidia_smi(["-q", "-x"]))
        driver = _child_text(root, "driver_version")
        cuda = _child_text(root, "cuda_version")
        out: List[GPUDevice] = []
        for idx, gpu in enumerate(root.findall("gpu")):
            mem = gpu.find("fb_memory_usage")
            total, used, free = (_first_int(_child_text(mem, t, "0")) for t in ("total", "used", "free"))
            power = gpu.find("gpu_power_readings")
            if power is None:
                power = gpu.find("power_readings")
            draw = _first_float(_child_text(power, "power_draw", "0"))
            limit = 0.0
            for tag in ("power_limit", "current_power_limit", "enforced_power_limit", "default_power_limit"):
                limit = _first_float(_child_text(power, tag, "0"))
                if limit > 0:
                    break
            util = gpu.find("utilization")
            temp = _first_int(_child_text(gpu.find("temperature"), "gpu_temp", "0"))
            gpu_util = _first_float(_child_text(util, "gpu_util", "0"))
            procs = []
            pnode = gpu.find("processes")
            if pnode is not None:
                for p in pnode.findall("process_info"):
                    procs.append(GPUProcess(pid=_first_int(_child_text(p, "pid", "0")),
                                            name=_child_text(p, "process_name"),
                                            used_memory_mib=_first_int(_child_text(p, "used_memory", "0")),
                                            gpu_instance_id=_child_text(p, "gpu_instance_id") or None))
            clocks, max_clocks = gpu.find("clocks"), gpu.find("max_clocks")
            reasons = []
            for tag in ("clocks_event_reasons", "clocks_throttle_reasons"):
                rnode = gpu.find(tag)
                if rnode is not None:
                    reasons = [c.tag.replace("clocks_event_reason_", "").replace("clocks_throttle_reason_", "")
                               for c in rnode if (c.text or "").strip() == "Active"]
                    break
            mem_pct = used / total * 100 if total > 0 else 0.0
            health, alerts = self._assess_health(temp, gpu_util, mem_pct, draw, limit, reasons)
            out.append(GPUDevice(
                index=idx, name=_child_text(gpu, "product_name", f"GPU {idx}"), uuid=_child_text(gpu, "uuid"),
                temperature_celsius=temp, gpu_utilization_pct=gpu_util, memory_used_mib=used,
                memory_total_mib=total, memory_free_mib=free, memory_utilization_pct=round(mem_pct, 1),
                power_draw_watts=draw, power_limit_watts=limit,
                fan_speed_pct=_first_int(_child_text(gpu, "fan_speed", "0")), driver_version=driver,
                cuda_version=cuda, compute_mode=_child_text(gpu, "compute_mode", "Default"),
                pci_bus_id=_child_text(gpu.find("pci"), "pci_bus_id"), processes=procs, health=health,
                alerts=alerts, sm_clock_mhz=_first_int(_child_text(clocks, "sm_clock", "0")),
                sm_clock_max_mhz=_first_int(_child_text(max_clocks, "sm_clock", "0")), throttle_reasons=reasons))
        return out

    def parse_csv(self, csv_str: Optional[str] = None) -> List[GPUDevice]:
        """Parse ``--query-gpu=... --format=csv,noheader,nounits`` (12 columns)."""
        if csv_str is None:
            csv_str = self._run_nvidia_smi([f"--query-gpu={_CSV_FIELDS}", "--format=csv,noheader,nounits"])
        out: List[GPUDevice] = []
        for line in csv_str.strip().splitlines():
            cols = [c.strip() for c in line.split(",")]
            if len(cols) < 12 or not cols[0]:
                continue

            def num(i: int, cast=float):
                try:
                    return cast(float(cols[i]))
                except ValueError:  # "[N/A]", "[Not Supported]", ...
                    return cast(0)

            total, used, free = num(7, int), num(6, int), num(8, int)
            temp, util, draw, limit = num(3, int), num(4), num(9), num(10)
            mem_pct = used / total * 100 if total > 0 else 0.0
            health, alerts = self._assess_health(temp, util, mem_pct, draw, limit)
            out.append(GPUDevice(index=num(0, int), name=cols[1], uuid=cols[2], temperature_celsius=temp,
                                 gpu_utilization_pct=util, memory_used_mib=used, memory_total_mib=total,
                                 memory_free_mib=free, memory_utilization_pct=round(mem_pct, 1),
                                 power_draw_watts=draw, power_limit_watts=limit, fan_speed_pct=num(11, int),
                                 health=health, alerts=alerts))
        return out

    # ------------------------------------------------------------------ NVML
    def _nvml_handle(self):
        with self._nvml_lock:
            if self._nvml is None:
                import pynvml  # nvidia-ml-py
                pynvml.nvmlInit()
                self._nvml = pynvml
            return self._nvml

    def poll_nvml(self) -> List[GPUDevice]:
        """One NVML sweep over all devices (H4 in SURVEY.md §8)."""
        nv = self._nvml_handle()

        def safe(fn, *a, default=None):
            try:
                return fn(*a)
            except Exception:
                return default

        def as_str(v) -> str:
            return v.decode() if isinstance(v, bytes) else (v or "")

        driver = as_str(safe(nv.nvmlSystemGetDriverVersion, default=""))
        cuda_v = safe(nv.nvmlSystemGetCudaDriverVersion, default=0) or 0
        cuda = f"{cuda_v // 1000}.{(cuda_v % 1000) // 10}" if cuda_v else ""
        reason_bits = {
            "sw_power_cap": getattr(nv, "nvmlClocksEventReasonSwPowerCap", getattr(nv, "nvmlClocksThrottleReasonSwPowerCap", 0x4)),
            "hw_slowdown": getattr(nv, "nvmlClocksThrottleReasonHwSlowdown", 0x8),
            "sw_thermal_slowdown": getattr(nv, "nvmlClocksThrottleReasonSwThermalSlowdown", 0x20),
            "hw_thermal_slowdown": getattr(nv, "nvmlClocksThrottleReasonHwThermalSlowdown", 0x40),
            "hw_power_brake_slowdown": getattr(nv, "nvmlClocksThrottleReasonHwPowerBrakeSlowdown", 0x80),
            "applications_clocks_setting": getattr(nv, "nvmlClocksThrottleReasonApplicationsClocksSetting", 0x2),
        }
        out: List[GPUDevice] = []
        for i in range(nv.nvmlDeviceGetCount()):
            h = nv.nvmlDeviceGetHandleByIndex(i)
            mem = safe(nv.nvmlDeviceGetMemoryInfo, h)
            total = int(mem.total // (1 << 20)) if mem else 0
            used = int(mem.used // (1 << 20)) if mem else 0
            free = int(mem.free // (1 << 20)) if mem else 0
            util = safe(nv.nvmlDeviceGetUtilizationRates, h)
            temp = safe(nv.nvmlDeviceGetTemperature, h, nv.NVML_TEMPERATURE_GPU, default=0) or 0
            draw = (safe(nv.nvmlDeviceGetPowerUsage, h, default=0) or 0) / 1000.0
            limit = (safe(nv.nvmlDeviceGetEnforcedPowerLimit, h, default=0) or 0) / 1000.0
            mask = safe(getattr(nv, "nvmlDeviceGetCurrentClocksEventReasons", None) or
                        nv.nvmlDeviceGetCurrentClocksThrottleReasons, h, default=0) or 0
            reasons = [k for k, bit in reason_bits.items() if mask & bit]
            links_total = links_up = 0
            for link in range(getattr(nv, "NVML_NVLINK_MAX_LINKS", 18)):
                st = safe(nv.nvmlDeviceGetNvLinkState, h, link)
                if st is None:
                    break
                links_total += 1
                links_up += 1 if st else 0
            procs = []
            for p in safe(nv.nvmlDeviceGetComputeRunningProcesses, h, default=[]) or []:
                procs.append(GPUProcess(pid=p.pid, name=as_str(safe(nv.nvmlSystemGetProcessName, p.pid, default="")),
                                        used_memory_mib=int((p.usedGpuMemory or 0) // (1 << 20))))
            ecc = safe(nv.nvmlDeviceGetTotalEccErrors, h, nv.NVML_MEMORY_ERROR_TYPE_UNCORRECTED,
                       nv.NVML_VOLATILE_ECC, default=0) or 0
            pending = bool(safe(nv.nvmlDeviceGetRetiredPagesPendingStatus, h, default=0))
            pci = safe(nv.nvmlDeviceGetPciInfo, h)
            mem_pct = used / total * 100 if total else 0.0
            gpu_util = float(util.gpu) if util else 0.0
            health, alerts = self._assess_health(temp, gpu_util, mem_pct, draw, limit, reasons, ecc, pending)
            out.append(GPUDevice(
                index=i, name=as_str(safe(nv.nvmlDeviceGetName, h, default=f"GPU {i}")),
                uuid=as_str(safe(nv.nvmlDeviceGetUUID, h, default="")), temperature_celsius=int(temp),
                gpu_utilization_pct=gpu_util, memory_used_mib=used, memory_total_mib=total, memory_free_mib=free,
                memory_utilization_pct=round(mem_pct, 1), power_draw_watts=draw, power_limit_watts=limit,
                fan_speed_pct=int(safe(nv.nvmlDeviceGetFanSpeed, h, default=0) or 0), driver_version=driver,
                cuda_version=cuda, pci_bus_id=as_str(getattr(pci, "busId", "")) if pci else "", processes=procs,
                health=health, alerts=alerts,
                sm_clock_mhz=int(safe(nv.nvmlDeviceGetClockInfo, h, nv.NVML_CLOCK_SM, default=0) or 0),
                sm_clock_max_mhz=int(safe(nv.nvmlDeviceGetMaxClockInfo, h, nv.NVML_CLOCK_SM, default=0) or 0),
                throttle_reasons=reasons, ecc_errors_uncorrected=int(ecc), retired_pages_pending=pending,
                nvlink_active_links=links_up, nvlink_total_links=links_total,
                numa_node=safe(getattr(nv, "nvmlDeviceGetNumaNodeId", lambda _h: None), h)))
        return out

    # ------------------------------------------------------------------ aggregation
    def _query_devices(self, xml_str: Optional[str] = None) -> Tuple[List[GPUDevice], str]:
        """NVML → XML → CSV. Raises RuntimeError only when every source fails."""
        errors = []
        if xml_str is None and self.backend in ("auto", "nvml"):
            try:
                return self.poll_nvml(), "nvml"
            except Exception as e:
                errors.append(f"nvml: {e}")
                if self.backend == "nvml":
                    raise RuntimeError("; ".join(errors))
        try:
            return self.parse_xml(xml_str), "xml"
        except Exception as e:
            errors.append(f"xml: {e}")
        try:
            return self.parse_csv(), "csv"
        except Exception as e:
            errors.append(f"csv: {e}")
        raise RuntimeError("no GPU telemetry source available (" + "; ".join(errors) + ")")

    def summarize(self, devices: Sequence[GPUDevice], source: str = "none") -> GPUFleetStatus:
        n = len(devices)
        alerts = [f"GPU {d.index} ({d.name}): {a}" for d in devices for a in d.alerts]
        available = sum(d.is_available for d in devices)
        if n and not available:
            alerts.insert(0, "CRITICAL: No GPUs available for scheduling")
        return GPUFleetStatus(
            total_gpus=n, healthy_gpus=sum(d.health == GPUHealthStatus.HEALTHY for d in devices),
            available_gpus=available, total_memory_mib=sum(d.memory_total_mib for d in devices),
            used_memory_mib=sum(d.memory_used_mib for d in devices),
            avg_utilization_pct=round(sum(d.gpu_utilization_pct for d in devices) / max(n, 1), 1),
            avg_temperature_celsius=round(sum(d.temperature_celsius for d in devices) / max(n, 1), 1),
            total_power_watts=round(sum(d.power_draw_watts for d in devices), 1), devices=list(devices),
            alerts=alerts, source=source)

    def get_fleet_status(self, xml_str: Optional[str] = None) -> GPUFleetStatus:
        """Never raises: an unreachable driver yields an empty fleet with one alert
        (reference contract, ref ``:282-290``)."""
        try:
            devices, source = self._query_devices(xml_str)
        except Exception:
            return GPUFleetStatus(alerts=["Unable to query nvidia-smi. No GPUs detected."])
        return self.summarize(devices, source)

    # ------------------------------------------------------------------ selection
    def select_best_gpu(self, required_memory_mib: int = 0, xml_str: Optional[str] = None) -> Optional[GPUDevice]:
        """Most free memory among schedulable devices; ties → least busy. Raises
        RuntimeError when no telemetry source works (so callers can fall back to mock)."""
        devices, _ = self._query_devices(xml_str)
        fit = [d for d in devices if d.is_available and d.memory_free_mib >= required_memory_mib]
        return min(fit, key=lambda d: (-d.memory_free_mib, d.gpu_utilization_pct), default=None)

    def select_gpus(self, count: int, required_memory_mib: int = 0, devices: Optional[Sequence[GPUDevice]] = None,
                    link_matrix: Optional[Dict[int, Dict[int, int]]] = None) -> Optional[List[GPUDevice]]:
        """Pick ``count`` devices for one job. Eligible = schedulable, enough free memory,
        no hardware throttle, no pending page retirement. Among eligible devices prefer the
        subset with the best worst-case NVLink connectivity (``link_matrix[i][j]`` = number of
        links, from :mod:`parallel.topology`); on an NVSwitch box every pair is equal and the
        choice degenerates to most-free-memory. Returns None if fewer than ``count`` qualify."""
        if devices is None:
            devices, _ = self._query_devices()
        ok = [d for d in devices
              if d.is_available and d.memory_free_mib >= required_memory_mib and not d.retired_pages_pending
              and not any(r in _HW_THROTTLE for r in d.throttle_reasons)]
        if len(ok) < count:
            return None
        ok.sort(key=lambda d: (-d.memory_free_mib, d.gpu_utilization_pct, d.index))
        if not link_matrix or count == 1:
            return sorted(ok[:count], key=lambda d: d.index)
        # greedy clique growth from each seed; score = min pairwise link count
        best: Optional[List[GPUDevice]] = None
        best_score = (-1, -1)
        for seed in ok:
            chosen = [seed]
            pool = [d for d in ok if d is not seed]
            while len(chosen) < count and pool:
                nxt = max(pool, key=lambda d: (min(link_matrix.get(d.index, {}).get(c.index, 0) for c in chosen),
                                               d.memory_free_mib))
                chosen.append(nxt)
                pool.remove(nxt)
            if len(chosen) < count:
                continue
            score = (min(link_matrix.get(a.index, {}).get(b.index, 0) for a in chosen for b in chosen if a is not b),
                     sum(d.memory_free_mib for d in chosen))
            if score > best_score:
                best, best_score = chosen, score
        return sorted(best, key=lambda d: d.index) if best else None

    # ------------------------------------------------------------------ health
    def _assess_health(self, temp: int, util: float, mem_pct: float, power: float, power_limit: float,
                       throttle_reasons: Sequence[str] = (), ecc_uncorrected: int = 0,
                       retired_pending: bool = False) -> Tuple[GPUHealthStatus, List[str]]:
        t = self.thresholds
        alerts: List[str] = []
        level = 0  # 0 healthy, 1 warning, 2 critical

        def raise_to(lvl: int, text: str) -> None:
            nonlocal level
            alerts.append(text)
            level = max(level, lvl)

        if temp >= t.temp_critical:
            raise_to(2, f"CRITICAL: Temperature {temp}°C exceeds {t.temp_critical}°C")
        elif temp >= t.temp_warning:
            raise_to(1, f"WARNING: Temperature {temp}°C exceeds {t.temp_warning}°C")
        if mem_pct >= t.mem_critical:
            raise_to(2, f"CRITICAL: Memory {mem_pct:.0f}% exceeds {t.mem_critical:g}%")
        elif mem_pct >= t.mem_warning:
            raise_to(1, f"WARNING: Memory {mem_pct:.0f}% exceeds {t.mem_warning:g}%")
        if util >= t.util_warning:
            raise_to(1, f"WARNING: Utilization {util:.0f}% at max capacity")
        if power_limit > 0 and power / power_limit >= t.power_warning:
            raise_to(1, f"WARNING: Power {power:.0f}W near limit {power_limit:.0f}W")
        hw = [r for r in throttle_reasons if r in _HW_THROTTLE]
        if hw:
            raise_to(1, f"WARNING: Clocks throttled ({', '.join(hw)})")
        if ecc_uncorrected > 0:
            raise_to(2, f"CRITICAL: {ecc_uncorrected} uncorrected ECC errors")
        if retired_pending:
            raise_to(1, "WARNING: Page retirement pending (reset required)")
        return (GPUHealthStatus.HEALTHY, GPUHealthStatus.WARNING, GPUHealthStatus.CRITICAL)[level], alerts

    # reference-compatible helper names
    _parse_int = staticmethod(_first_int)
    _parse_float = staticmethod(_first_float)

    @staticmethod
    def _text(parent, tag, default=""):
        return _child_text(parent, tag, default) if isinstance(tag, str) else default

    # ------------------------------------------------------------------ mock
    def get_mock_fleet(self) -> GPUFleetStatus:
        """Canned two-GPU fleet for UI/tests without hardware (values: ref ``:400-431``)."""
        spec = [
            dict(index=0, uuid="GPU-mock-0001", temperature_celsius=45, gpu_utilization_pct=23.0,
                 memory_used_mib=12400, power_draw_watts=67.0,
                 processes=[GPUProcess(pid=12345, name="python", used_memory_mib=12400)]),
            dict(index=1, uuid="GPU-mock-0002", temperature_celsius=72, gpu_utilization_pct=89.0,
                 memory_used_mib=71000, power_draw_watts=295.0, health=GPUHealthStatus.WARNING,
                 alerts=["WARNING: Memory 86.7% exceeds 85%"],
                 processes=[GPUProcess(pid=23456, name="python", used_memory_mib=35000),
                            GPUProcess(pid=23457, name="python", used_memory_mib=36000)]),
        ]
        total = 81920
        devices = [GPUDevice(name="NVIDIA A100-SXM4-80GB", memory_total_mib=total,
                             memory_free_mib=total - s["memory_used_mib"],
                             memory_utilization_pct=round(s["memory_used_mib"] / total * 100, 1),
                             power_limit_watts=400.0, fan_speed_pct=0, driver_version="535.129.03",
                             cuda_version="12.2", **s) for s in spec]
        return GPUFleetStatus(total_gpus=2, healthy_gpus=1, available_gpus=1, total_memory_mib=163840,
                              used_memory_mib=83400, avg_utilization_pct=56.0, avg_temperature_celsius=58.5,
                              total_power_watts=362.0, devices=devices, alerts=devices[1].alerts, source="mock")


class FleetSampler:
    """Background poller: keeps the last ``capacity`` fleet snapshots (ring buffer) so the
    API can answer trend queries and the engine can correlate a loss spike with a thermal
    or clock event on the same timeline."""

    def __init__(self, manager: Optional[GPUManager] = None, interval_s: float = 1.0, capacity: int = 600):
        self.manager = manager or GPUManager()
        self.interval_s = interval_s
        self.snapshots: Deque[GPUFleetStatus] = deque(maxlen=capacity)
        self._stop = threading.Event()
        self._thread: Optional[threading.Thread] = None

    def start(self) -> "FleetSampler":
        if self._thread is None or not self._thread.is_alive():
            self._stop.clear()
            self._thread = threading.Thread(target=self._run, daemon=True, name="fleet-sampler")
            self._thread.start()
        return self

    def _run(self) -> None:
        while not self._stop.is_set():
            self.snapshots.append(self.manager.get_fleet_status())
            self._stop.wait(self.interval_s)

    def stop(self) -> None:
        self._stop.set()
        if self._thread is not None:
            self._thread.join(timeout=5)

    def latest(self) -> Optional[GPUFleetStatus]:
        return self.snapshots[-1] if self.snapshots else None

    def clocks_summary(self) -> Dict[str, Any]:
        """Median SM clock under load and the union of throttle reasons seen (bench.py's
        ``clocks`` key)."""
        mhz: List[int] = []
        mx = 0
        reasons: set = set()
        for snap in self.snapshots:
            for d in snap.devices:
                if d.gpu_utilization_pct >= 50 and d.sm_clock_mhz:
                    mhz.append(d.sm_clock_mhz)
                mx = max(mx, d.sm_clock_max_mhz)
                reasons.update(d.throttle_reasons)
        mhz.sort()
        return {"sm_mhz": mhz[len(mhz) // 2] if mhz else 0, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(self.snapshots)}
