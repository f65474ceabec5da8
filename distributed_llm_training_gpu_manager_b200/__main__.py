"""Command line for the orchestrator (the reference has none: its only entry point is the uvicorn app).

    python -m distributed_llm_training_gpu_manager_b200 config  --preset 7b [--set zero_stage=2 ...]
    python -m distributed_llm_training_gpu_manager_b200 launch  --preset 70b --script builtin:llama3-70b [--dry-run] [-- script args]
    python -m distributed_llm_training_gpu_manager_b200 fleet   [--mock]
    python -m distributed_llm_training_gpu_manager_b200 select  --count 4 [--memory-mib 40000]
    python -m distributed_llm_training_gpu_manager_b200 topology [--mock]
    python -m distributed_llm_training_gpu_manager_b200 serve   [--host 0.0.0.0 --port 8000]

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
    return {"deepspeed_config": launcher.generate_config(cfg), "effective_batch_size": cfg.effective_batch_size,
            "warnings": launcher.validate(cfg)}


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
    p = sub.add_parser("serve")
    p.add_argument("--host", default="0.0.0.0")
    p.add_argument("--port", type=int, default=8000)
    ns = ap.parse_args(argv)
    if ns.cmd == "serve":
        cmd_serve(ns)
        return 0
    out = {"config": cmd_config, "launch": cmd_launch, "fleet": cmd_fleet, "select": cmd_select,
           "topology": cmd_topology}[ns.cmd](ns)
    from .utils.json_safe import sanitize
    print(json.dumps(sanitize(out), indent=2, default=str))
    return 0


if __name__ == "__main__":
    sys.exit(main())
