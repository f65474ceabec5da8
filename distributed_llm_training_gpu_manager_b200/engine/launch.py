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

A job can also stop without anybody dying: a lost peer leaves the other ranks waiting inside a
collective (the fused kernels poll peer flags; NCCL kernels poll too). ``--stall_timeout S`` makes
every engine touch a per-rank heartbeat file after each optimizer step (``B200ZERO_HEARTBEAT_DIR``,
``ZeroEngine._optimizer_step``); when no rank has completed a step for ``S`` seconds the job is torn
down and relaunched like after a crash, and — ``--stall_fallback nccl``, the default — the relaunched
workers run on NCCL collectives (``B200ZERO_COMM=nccl``).
"""
from __future__ import annotations

import argparse
import os
import shutil
import signal
import subprocess
import sys
import tempfile
import time
from typing import List, Optional

STALLED = 124          # exit code of a job that was torn down because it stopped making progress

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
    ap.add_argument("--bind_cores", default=os.environ.get("B200ZERO_BIND_CORES", "none"), choices=["none", "even", "numa"],
                    help="CPU affinity of the rank processes: even = the usable cores cut into one contiguous slice per "
                         "rank; numa = the cores of the GPU's NUMA node shared between the ranks on that node. Keeps one "
                         "rank's launch thread, host optimizer and data loader off another rank's cores")
    ap.add_argument("--stall_timeout", type=float, default=float(os.environ.get("B200ZERO_STALL_TIMEOUT", "0")),
                    help="seconds without a completed optimizer step on any rank before the job counts as stalled "
                         "(0 = no stall detection); the first step gets this long plus --stall_startup")
    ap.add_argument("--stall_startup", type=float, default=float(os.environ.get("B200ZERO_STALL_STARTUP", "600")),
                    help="extra time allowed before the first heartbeat (model build, rendezvous, checkpoint load)")
    ap.add_argument("--stall_fallback", default=os.environ.get("B200ZERO_STALL_FALLBACK", "nccl"),
                    choices=["none", "nccl"], help="after a stall, relaunch on NCCL collectives (default) or unchanged")
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


def _parse_cpulist(text: str) -> List[int]:
    out: List[int] = []
    for part in text.strip().split(","):
        if not part:
            continue
        a, _, b = part.partition("-")
        out.extend(range(int(a), int(b or a) + 1))
    return out


def core_sets(n_ranks: int, mode: str, numa_of_rank: Optional[List[Optional[int]]] = None,
              usable: Optional[List[int]] = None, node_cpus=None) -> List[Optional[List[int]]]:
    """Cores for each local rank (None = leave the affinity alone). ``node_cpus(node) -> [cpu, ...]`` reads
    /sys/devices/system/node/node<N>/cpulist by default."""
    if mode == "none" or n_ranks < 1:
        return [None] * n_ranks
    usable = sorted(usable if usable is not None else os.sched_getaffinity(0))
    if mode == "numa" and numa_of_rank and all(n is not None for n in numa_of_rank[:n_ranks]):
        if node_cpus is None:
            def node_cpus(node):
                try:
                    with open(f"/sys/devices/system/node/node{node}/cpulist") as f:
                        return _parse_cpulist(f.read())
                except OSError:
                    return []
        out: List[Optional[List[int]]] = [None] * n_ranks
        for node in sorted(set(numa_of_rank[:n_ranks])):
            ranks = [r for r in range(n_ranks) if numa_of_rank[r] == node]
            cpus = [c for c in node_cpus(node) if c in set(usable)]
            per = len(cpus) // len(ranks)
            if per < 1:
                return core_sets(n_ranks, "even", usable=usable)
            for i, r in enumerate(ranks):
                out[r] = cpus[i * per:(i + 1) * per]
        return out
    per = len(usable) // n_ranks
    if per < 1:
        return [None] * n_ranks
    return [usable[r * per:(r + 1) * per] for r in range(n_ranks)]


def _spawn(ns: argparse.Namespace, restart: int, extra_env: Optional[dict] = None) -> List[subprocess.Popen]:
    world = ns.num_gpus * ns.num_nodes
    addr = "127.0.0.1" if ns.master_addr == "localhost" else ns.master_addr
    devices = _devices_for(ns.num_gpus, ns.placement)
    numa = None
    if ns.bind_cores == "numa":
        try:
            from ..parallel.topology import discover_topology
            topo = discover_topology()
            numa = [topo.numa[d] if d < len(topo.numa) else None for d in devices]
        except Exception:
            numa = None
    cores = core_sets(ns.num_gpus, ns.bind_cores, numa)
    procs = []
    for local in range(ns.num_gpus):
        rank = ns.node_rank * ns.num_gpus + local
        env = dict(os.environ, RANK=str(rank), LOCAL_RANK=str(local), WORLD_SIZE=str(world), MASTER_ADDR=addr,
                   MASTER_PORT=str(ns.master_port + restart), LOCAL_WORLD_SIZE=str(ns.num_gpus),
                   B200ZERO_RESTART_COUNT=str(restart), B200ZERO_BACKEND=ns.backend,
                   B200ZERO_DEVICE=str(devices[local]), **(extra_env or {}))
        if "OMP_NUM_THREADS" not in env:  # N ranks x all cores oversubscribes the host badly
            env["OMP_NUM_THREADS"] = str(max(1, (os.cpu_count() or 1) // max(ns.num_gpus, 1)))
        if ns.script.startswith("builtin:"):
            cmd = [sys.executable, "-m", f"{PKG}.engine.worker", f"--model={ns.script.split(':', 1)[1]}"]
        else:
            cmd = [sys.executable, ns.script]
        cmd += [f"--local_rank={local}"] + list(ns.script_args)
        mine = cores[local]
        if mine and "OMP_NUM_THREADS" not in os.environ:
            env["OMP_NUM_THREADS"] = str(len(mine))
        procs.append(subprocess.Popen(cmd, env=env, preexec_fn=(lambda c=mine: os.sched_setaffinity(0, c)) if mine else None))
    return procs


def _teardown(procs: List[subprocess.Popen], hard: bool = False) -> None:
    for p in procs:
        if p.poll() is None:
            p.kill() if hard else p.terminate()
    deadline = time.time() + 10
    for p in procs:
        try:
            p.wait(timeout=max(0.1, deadline - time.time()))
        except subprocess.TimeoutExpired:
            p.kill()


def newest_heartbeat(hb_dir: str) -> Optional[float]:
    """mtime of the most recent per-rank heartbeat file, None before the first one."""
    try:
        times = [os.path.getmtime(os.path.join(hb_dir, f)) for f in os.listdir(hb_dir)]
    except OSError:
        return None
    return max(times) if times else None


def _supervise(procs: List[subprocess.Popen], term_flag: List[bool], hb_dir: Optional[str] = None,
               stall_timeout: float = 0.0, stall_startup: float = 600.0) -> int:
    """Wait for all ranks; on the first abnormal exit kill the rest. Returns the job's exit code
    (``STALLED`` when the job was stopped because no rank completed a step for ``stall_timeout`` seconds)."""
    forwarded = False
    t0 = time.time()
    while True:
        if stall_timeout > 0 and hb_dir and not term_flag[0]:
            last = newest_heartbeat(hb_dir)
            now = time.time()
            quiet = now - last if last is not None else now - t0 - stall_startup
            if quiet > stall_timeout:
                print(f"[launch] no rank has completed a step for {quiet:.0f} s "
                      f"({'since the first heartbeat' if last is not None else 'and none ever did'}): "
                      f"stopping the job", flush=True)
                _teardown(procs, hard=True)      # a rank that waits inside a kernel does not react to SIGTERM
                return STALLED
        if term_flag[0] and not forwarded:
            for p in procs:
                if p.poll() is None:
                    p.send_signal(signal.SIGTERM)
            forwarded = True
        codes = [p.poll() for p in procs]
        bad = [c for c in codes if c not in (None, 0)]
        if bad:
            _teardown(procs)
            return bad[0]
        if all(c == 0 for c in codes):
            return 0
        time.sleep(0.05)


def main(argv: List[str] = None) -> int:
    ns = parse(sys.argv[1:] if argv is None else argv)
    term = [False]
    signal.signal(signal.SIGTERM, lambda *_: term.__setitem__(0, True))
    restart = 0
    extra_env: dict = {}
    while True:
        # heartbeat files: in the directory the caller named (the job registry reads progress from it), else in a
        # temporary one when stall detection is on
        given = os.environ.get("B200ZERO_HEARTBEAT_DIR")
        hb_dir = given or (tempfile.mkdtemp(prefix="b200zero_hb_") if ns.stall_timeout > 0 else None)
        if hb_dir:
            os.makedirs(hb_dir, exist_ok=True)
            for f in os.listdir(hb_dir):             # a relaunch starts with a clean slate: old marks are not progress
                if f.startswith("rank") and f[4:].isdigit():
                    try:
                        os.unlink(os.path.join(hb_dir, f))
                    except OSError:
                        pass
            extra_env["B200ZERO_HEARTBEAT_DIR"] = hb_dir
        procs = _spawn(ns, restart, extra_env)
        try:
            rc = _supervise(procs, term, hb_dir, ns.stall_timeout, ns.stall_startup)
        finally:
            if hb_dir and not given:
                shutil.rmtree(hb_dir, ignore_errors=True)
        if rc == 0 or term[0] or restart >= ns.max_restarts:
            return rc
        restart += 1
        if rc == STALLED and ns.stall_fallback == "nccl":
            extra_env["B200ZERO_COMM"] = "nccl"
        print(f"[launch] job {'stalled' if rc == STALLED else f'failed with exit code {rc}'}; restart "
              f"{restart}/{ns.max_restarts} (workers resume from the newest verified checkpoint"
              f"{', on NCCL collectives' if extra_env.get('B200ZERO_COMM') == 'nccl' else ''})", flush=True)


if __name__ == "__main__":
    sys.exit(main())
