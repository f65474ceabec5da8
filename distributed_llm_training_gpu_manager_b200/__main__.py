"""Command line for the orchestrator (the reference has none: its only entry point is the uvicorn app).

    python -m distributed_llm_training_gpu_manager_b200 config  --preset 7b [--set zero_stage=2 ...]
    python -m distributed_llm_training_gpu_manager_b200 launch  --preset 70b --script builtin:llama3-70b [--dry-run] [-- script args]
    python -m distributed_llm_training_gpu_manager_b200 fleet   [--mock]
    python -m distributed_llm_training_gpu_manager_b200 select  --count 4 [--memory-mib 40000]
    python -m distributed_llm_training_gpu_manager_b200 topology [--mock]
    python -m distributed_llm_training_gpu_manager_b200 serve   [--host 0.0.0.0 --port 8000]
    python -m distributed_llm_training_gpu_manager_b200 doctor  [--gpus N]     # is this box ready to train?

Everything prints JSON on stdout, so the commands compose with jq / scripts.
"""
from __future__ import annotations

import argparse
import json
import sys
from typing import Any, Dict, List


def _coerce(v: str) -> Any:
    for conv in (int, float):
        try:
            return conv(v)
        except ValueError:
            pass
    if v.lower() in ("true", "false"):
        return v.lower() == "true"
    return v


def _config(ns):
    from .ai_engine import DeepSpeedConfig, DeepSpeedLauncher
    presets = dict(DeepSpeedLauncher.presets(), **{f"b200-{k}": v for k, v in DeepSpeedLauncher.b200_presets().items()})
    if ns.preset:
        if ns.preset not in presets:
            raise SystemExit(f"unknown preset {ns.preset!r}; available: {sorted(presets)}")
        base = presets[ns.preset].model_dump()
    else:
        base = {}
    for kv in ns.set or []:
        k, _, v = kv.partition("=")
        base[k] = _coerce(v)
    return DeepSpeedConfig(**base)


def cmd_config(ns) -> Dict[str, Any]:
    from .ai_engine import DeepSpeedLauncher
    cfg = _config(ns)
    launcher = DeepSpeedLauncher()
    ds = launcher.generate_config(cfg)
    out = {"deepspeed_config": ds, "effective_batch_size": cfg.effective_batch_size, "warnings": launcher.validate(cfg)}
    if getattr(ns, "explain", False):
        from .engine.config import explain
        out["native_engine_handling"] = explain(ds)       # key by key: honoured / fixed / not applicable, and why
    return out


def cmd_launch(ns) -> Dict[str, Any]:
    from .ai_engine import DeepSpeedLauncher
    res = DeepSpeedLauncher().launch(_config(ns), ns.script, list(ns.script_args or []), dry_run=ns.dry_run)
    return res.model_dump()


def cmd_fleet(ns) -> Dict[str, Any]:
    from .ai_engine import GPUManager
    m = GPUManager()
    return (m.get_mock_fleet() if ns.mock else m.get_fleet_status()).model_dump()


def cmd_select(ns) -> Dict[str, Any]:
    from .ai_engine import GPUManager
    from .parallel.topology import discover_topology
    m = GPUManager()
    devices = m.get_mock_fleet().devices if ns.mock else None
    links = discover_topology(force_mock=ns.mock).link_counts()
    picked = m.select_gpus(ns.count, required_memory_mib=ns.memory_mib, devices=devices, link_matrix=links)
    return {"requested": ns.count, "selected": [d.index for d in picked or []]}


def cmd_topology(ns) -> Dict[str, Any]:
    from .parallel.topology import discover_topology
    t = discover_topology(force_mock=ns.mock)
    return {"node_type": t.node_type, "interconnect": t.interconnect, "links_per_gpu": t.links_per_gpu,
            "matrix": t.matrix_labels(), "bottlenecks": t.bottlenecks()}


def cmd_doctor(ns) -> Dict[str, Any]:
    """Pre-flight check of one node: kernel library, devices, peer access between the GPUs a job would use, VMM /
    multicast support for the NVLS path, host memory for the offload tiers, /dev/shm and the pinned-memory limit.
    Every item is {ok, detail}; ``ready`` is the conjunction of the items a default (fused, no offload) job needs."""
    import os
    import resource
    import shutil

    import torch

    from .ops import build as _build
    from .ops import lib as _lib
    checks: Dict[str, Dict[str, Any]] = {}

    def put(name: str, ok: bool, detail: Any, needed: bool = True) -> None:
        checks[name] = {"ok": bool(ok), "detail": detail, "needed_for_default_job": needed}

    put("kernel_library", _build.LIB_CUDA.exists(), str(_build.LIB_CUDA))
    put("host_library", _build.LIB_HOST.exists(), str(_build.LIB_HOST), needed=False)
    n_dev = torch.cuda.device_count() if torch.cuda.is_available() else 0
    want = ns.gpus or n_dev
    put("cuda_devices", n_dev >= max(want, 1), {"visible": n_dev, "wanted": want})
    if n_dev:
        props = [torch.cuda.get_device_properties(i) for i in range(n_dev)]
        put("sm_100", all(p.major == 10 for p in props),
            [{"index": i, "name": p.name, "cc": f"{p.major}.{p.minor}", "sms": p.multi_processor_count,
              "hbm_gib": round(p.total_memory / 2 ** 30, 1)} for i, p in enumerate(props)])
        L = None
        try:
            L = _lib.load(required=True)
            put("library_loads", True, {"version": int(L.bz_version()), "sm_count": int(L.bz_sm_count())})
        except Exception as exc:  # noqa: BLE001
            put("library_loads", False, str(exc))
        bad = [(i, j) for i in range(min(want, n_dev)) for j in range(min(want, n_dev))
               if i != j and not torch.cuda.can_device_access_peer(i, j)]
        put("peer_access", not bad, {"pairs_without_p2p": bad}, needed=want > 1)
        if L is not None:
            import ctypes as C
            vmm, mc = C.c_int(0), C.c_int(0)
            rc = L.bz_vmm_supported(C.byref(vmm), C.byref(mc))
            put("vmm_allocator", rc == 0 and bool(vmm.value), {"posix_fd_handles": bool(vmm.value)}, needed=False)
            put("nvswitch_multicast", rc == 0 and bool(mc.value), {"multicast_objects": bool(mc.value)}, needed=False)
    try:
        page = os.sysconf("SC_PAGE_SIZE")
        total_gib = os.sysconf("SC_PHYS_PAGES") * page / 2 ** 30
        avail_gib = os.sysconf("SC_AVPHYS_PAGES") * page / 2 ** 30
        put("host_memory", True, {"total_gib": round(total_gib, 1), "available_gib": round(avail_gib, 1),
                                  "note": "optimizer offload needs 12 bytes per parameter of this node's shards, "
                                          "parameter offload 2 more"}, needed=False)
    except (ValueError, OSError):
        pass
    soft, _hard = resource.getrlimit(resource.RLIMIT_MEMLOCK)
    put("memlock_limit", soft == resource.RLIM_INFINITY or soft >= 64 << 30,
        {"soft": "unlimited" if soft == resource.RLIM_INFINITY else soft,
         "note": "pinned staging for offload / checkpoints needs a large locked-memory limit (IPC_LOCK in containers)"},
        needed=False)
    try:
        shm = shutil.disk_usage("/dev/shm")
        put("dev_shm", shm.free >= 1 << 30, {"free_gib": round(shm.free / 2 ** 30, 1)}, needed=False)
    except OSError:
        put("dev_shm", False, "no /dev/shm", needed=False)
    put("cpu_cores", True, {"usable": len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else os.cpu_count(),
                            "note": "host AdamW threads = cores / ranks on the node"}, needed=False)
    ready = all(c["ok"] for c in checks.values() if c["needed_for_default_job"])
    return {"ready": ready, "checks": checks}


def cmd_serve(ns) -> None:
    import uvicorn
    uvicorn.run("distributed_llm_training_gpu_manager_b200.backend.main:app", host=ns.host, port=ns.port)


def main(argv: List[str] = None) -> int:
    ap = argparse.ArgumentParser(prog="distributed_llm_training_gpu_manager_b200", description=__doc__,
                                 formatter_class=argparse.RawDescriptionHelpFormatter)
    sub = ap.add_subparsers(dest="cmd", required=True)
    for name in ("config", "launch"):
        p = sub.add_parser(name)
        p.add_argument("--preset", default=None, help="7b | 13b | 70b | b200-<name>")
        p.add_argument("--set", action="append", metavar="FIELD=VALUE", help="override a DeepSpeedConfig field")
        if name == "config":
            p.add_argument("--explain", action="store_true", help="say, key by key, what the native engine does with it")
        if name == "launch":
            p.add_argument("--script", required=True, help="training script, or builtin:<model> for the packaged worker")
            p.add_argument("--dry-run", action="store_true")
            p.add_argument("script_args", nargs="*")
    for name in ("fleet", "topology"):
        sub.add_parser(name).add_argument("--mock", action="store_true")
    p = sub.add_parser("select")
    p.add_argument("--count", type=int, default=1)
    p.add_argument("--memory-mib", type=int, default=0)
    p.add_argument("--mock", action="store_true")
    sub.add_parser("doctor").add_argument("--gpus", type=int, default=0, help="GPUs the job will use (default: all)")
    p = sub.add_parser("serve")
    p.add_argument("--host", default="0.0.0.0")
    p.add_argument("--port", type=int, default=8000)
    ns = ap.parse_args(argv)
    if ns.cmd == "serve":
        cmd_serve(ns)
        return 0
    out = {"config": cmd_config, "launch": cmd_launch, "fleet": cmd_fleet, "select": cmd_select,
           "topology": cmd_topology, "doctor": cmd_doctor}[ns.cmd](ns)
    from .utils.json_safe import sanitize
    print(json.dumps(sanitize(out), indent=2, default=str))
    return 0


if __name__ == "__main__":
    sys.exit(main())
