"""Helpers for multi-process CPU (gloo) tests: spawn N ranks of a function on 127.0.0.1."""
import os
import socket
import traceback

import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _entry(rank, world, port, fn, args, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    torch.set_num_threads(2)
    try:
        dist.init_process_group("gloo", rank=rank, world_size=world)
        out = fn(rank, world, *args)
        ret[rank] = ("ok", out)
    except BaseException:
        ret[rank] = ("err", traceback.format_exc())
    finally:
        if dist.is_initialized():
            dist.destroy_process_group()


def run_ranks(fn, world: int = 2, args=()):
    """Run ``fn(rank, world, *args)`` on ``world`` gloo ranks; returns the list of results."""
    ctx = mp.get_context("spawn")
    with ctx.Manager() as mgr:
        ret = mgr.dict()
        port = free_port()
        procs = [ctx.Process(target=_entry, args=(r, world, port, fn, args, ret)) for r in range(world)]
        for p in procs:
            p.start()
        for p in procs:
            p.join(600)
        for p in procs:
            if p.is_alive():
                p.kill()
        results = dict(ret)
    for r in range(world):
        assert r in results, f"rank {r} produced no result (crashed or timed out)"
        status, payload = results[r]
        assert status == "ok", f"rank {r} failed:\n{payload}"
    return [results[r][1] for r in range(world)]


def _entry_nccl(rank, world, port, fn, args, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    try:
        torch.cuda.set_device(rank)
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
        out = fn(rank, world, *args)
        ret[rank] = ("ok", out)
    except BaseException:
        ret[rank] = ("err", traceback.format_exc())
    finally:
        if dist.is_initialized():
            dist.destroy_process_group()


def run_ranks_nccl(fn, world: int, args=(), timeout: float = 900):
    """One process per GPU over NCCL on this node."""
    ctx = mp.get_context("spawn")
    with ctx.Manager() as mgr:
        ret = mgr.dict()
        port = free_port()
        procs = [ctx.Process(target=_entry_nccl, args=(r, world, port, fn, args, ret)) for r in range(world)]
        for p in procs:
            p.start()
        for p in procs:
            p.join(timeout)
        for p in procs:
            if p.is_alive():
                p.kill()
        results = dict(ret)
    for r in range(world):
        assert r in results, f"rank {r} produced no result (crashed or timed out)"
        status, payload = results[r]
        assert status == "ok", f"rank {r} failed:\n{payload}"
    return [results[r][1] for r in range(world)]
