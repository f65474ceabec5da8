"""Native job launcher: ``python -m <pkg>.engine.launch --num_gpus=N script --deepspeed_config=… [args]``.

Drop-in for the ``deepspeed`` CLI invocation the reference builds
(``ai_engine/deepspeed_launcher.py:277-300``): same flags, one worker process per GPU with
``RANK / LOCAL_RANK / WORLD_SIZE / MASTER_ADDR / MASTER_PORT`` set, each running
``script --local_rank=i --deepspeed_config=… args``. ``script`` may be ``builtin:<model>`` to run
the packaged worker (``engine/worker.py``).

It is also the **supervisor** the reference lacks: it keeps the child handles,
forwards SIGTERM (emergency checkpoint), and when any rank dies abnormally it tears the others
down and — with ``--max_restarts`` — relaunches the whole job, which resumes from the newest
verified checkpoint (rank failure → detect → roll back → resume).
"""
from __future__ import annotations

import argparse
import os
import signal
import subprocess
import sys
import time
from typing import List

PKG = (__package__ or __name__).split(".")[0]


def parse(argv: List[str]) -> argparse.Namespace:
    ap = argparse.ArgumentParser(prog="b200zero-launch", allow_abbrev=False)
    ap.add_argument("--num_gpus", type=int, default=1)
    ap.add_argument("--num_nodes", type=int, default=1)
    ap.add_argument("--node_rank", type=int, default=int(os.environ.get("NODE_RANK", "0")))
    ap.add_argument("--master_addr", default=os.environ.get("MASTER_ADDR", "127.0.0.1"))
    ap.add_argument("--master_port", type=int, default=int(os.environ.get("MASTER_PORT", "29500")))
    ap.add_argument("--max_restarts", type=int, default=int(os.environ.get("B200ZERO_MAX_RESTARTS", "0")))
    ap.add_argument("--backend", default=os.environ.get("B200ZERO_BACKEND", "auto"))
    ap.add_argument("--placement", default="topology", choices=["topology", "linear"])
    ap.add_argument("script")
    ap.add_argument("script_args", nargs=argparse.REMAINDER)
    return ap.parse_args(argv)


def _devices_for(n: int, placement: str) -> List[int]:
    if placement == "topology":
        try:
            from ..parallel.topology import discover_topology
            order = discover_topology().rank_order()
            if len(order) >= n:
                return order[:n]
        except Exception:
            pass
    return list(range(n))


def _spawn(ns: argparse.Namespace, restart: int) -> List[subprocess.Popen]:
    world = ns.num_gpus * ns.num_nodes
    addr = "127.0.0.1" if ns.master_addr == "localhost" else ns.master_addr
    devices = _devices_for(ns.num_gpus, ns.placement)
    procs = []
    for local in range(ns.num_gpus):
        rank = ns.node_rank * ns.num_gpus + local
        env = dict(os.environ, RANK=str(rank), LOCAL_RANK=str(local), WORLD_SIZE=str(world), MASTER_ADDR=addr,
                   MASTER_PORT=str(ns.master_port + restart), LOCAL_WORLD_SIZE=str(ns.num_gpus),
                   B200ZERO_RESTART_COUNT=str(restart), B200ZERO_BACKEND=ns.backend,
                   B200ZERO_DEVICE=str(devices[local]))
        if "OMP_NUM_THREADS" not in env:  # N ranks x all cores oversubscribes the host badly
            env["OMP_NUM_THREADS"] = str(max(1, (os.cpu_count() or 1) // max(ns.num_gpus, 1)))
        if ns.script.startswith("builtin:"):
            cmd = [sys.executable, "-m", f"{PKG}.engine.worker", f"--model={ns.script.split(':', 1)[1]}"]
        else:
            cmd = [sys.executable, ns.script]
        cmd += [f"--local_rank={local}"] + list(ns.script_args)
        procs.append(subprocess.Popen(cmd, env=env))
    return procs


def _supervise(procs: List[subprocess.Popen], term_flag: List[bool]) -> int:
    """Wait for all ranks; on the first abnormal exit kill the rest. Returns the job's exit code."""
    forwarded = False
    while True:
        if term_flag[0] and not forwarded:
            for p in procs:
                if p.poll() is None:
                    p.send_signal(signal.SIGTERM)
            forwarded = True
        codes = [p.poll() for p in procs]
        bad = [c for c in codes if c not in (None, 0)]
        if bad:
            for p in procs:
                if p.poll() is None:
                    p.terminate()
            deadline = time.time() + 10
            for p in procs:
                try:
                    p.wait(timeout=max(0.1, deadline - time.time()))
                except subprocess.TimeoutExpired:
                    p.kill()
            return bad[0]
        if all(c == 0 for c in codes):
            return 0
        time.sleep(0.05)


def main(argv: List[str] = None) -> int:
    ns = parse(sys.argv[1:] if argv is None else argv)
    term = [False]
    signal.signal(signal.SIGTERM, lambda *_: term.__setitem__(0, True))
    restart = 0
    while True:
        procs = _spawn(ns, restart)
        rc = _supervise(procs, term)
        if rc == 0 or term[0] or restart >= ns.max_restarts:
            return rc
        restart += 1
        print(f"[launch] job failed with exit code {rc}; restart {restart}/{ns.max_restarts} "
              f"(workers resume from the newest verified checkpoint)", flush=True)


if __name__ == "__main__":
    sys.exit(main())
