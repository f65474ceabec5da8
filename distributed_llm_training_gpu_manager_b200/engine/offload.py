"""ZeRO-Offload optimizer step (H1 / X9): fp32 master weights and Adam moments live
in pinned host memory (``offload_optimizer.device = "cpu"``, the reference's default —
ref ``ai_engine/deepspeed_launcher.py:39,197-203``); each step streams the gradient shard to the
host in chunks on a side stream, runs the C++ AdamW (``csrc/host/host_adam.cpp``) chunk by chunk,
and streams the new bf16 weights back, with device→host, host compute and host→device of
consecutive chunks overlapped (double-buffered pinned staging, ``buffer_count`` ≥ 2)."""
from __future__ import annotations

import ctypes as C
from typing import Optional

import torch

from ..ops import lib as _lib

_CHUNK = 1 << 26  # 64 Mi elements per chunk (128 MB bf16 down, 128 MB bf16 up)


def host_threads(world_size: int) -> int:
    """Worker threads for this rank's host optimizer: the cores this process may run on, shared fairly between the
    ranks of the node (every rank runs its host AdamW at the same moment of the step)."""
    import os
    try:
        cores = len(os.sched_getaffinity(0))
    except (AttributeError, OSError):
        cores = os.cpu_count() or 1
    local = int(os.environ.get("LOCAL_WORLD_SIZE", 0)) or max(1, world_size)
    return max(1, int(os.environ.get("B200ZERO_HOST_ADAM_THREADS", 0)) or cores // local)


def _host_lib():
    L = _lib.load_host(required=True)
    if not getattr(L, "_bz_bound", False):
        L.bz_host_adamw.restype = C.c_int
        L.bz_host_adamw.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_longlong,
                                    C.c_float, C.c_float, C.c_float, C.c_float, C.c_float, C.c_int, C.c_float, C.c_int]
        L._bz_bound = True
    return L


def host_adamw(master: torch.Tensor, m: torch.Tensor, v: torch.Tensor, grad: torch.Tensor, out_bf16: Optional[torch.Tensor],
               *, lr: float, beta1: float, beta2: float, eps: float, weight_decay: float, step: int,
               grad_scale: float = 1.0, threads: int = 0) -> None:
    """CPU tensors only. ``grad``: fp32 or bf16; ``out_bf16``: bf16 tensor or None."""
    L = _host_lib()
    assert not master.is_cuda and master.dtype == torch.float32
    rc = L.bz_host_adamw(master.data_ptr(), m.data_ptr(), v.data_ptr(), grad.data_ptr(),
                         int(grad.dtype == torch.bfloat16), out_bf16.data_ptr() if out_bf16 is not None else None,
                         master.numel(), lr, beta1, beta2, eps, weight_decay, step, grad_scale, threads)
    if rc != 0:
        raise RuntimeError(f"bz_host_adamw failed rc={rc}")


def host_adam_step(engine, lr: float, step_no: int) -> None:
    """Called by ``ZeroEngine._optimizer_step`` when optimizer state is offloaded."""
    cfg = engine.cfg
    dev = engine.device
    n = engine.layout.shard_numel
    # clip coefficient and skip flag were produced on the device; the host update needs them now
    host_vals = torch.stack([engine._coef[0], engine._flag[0].float()]).cpu()
    coef, skip = float(host_vals[0]), bool(host_vals[1] > 0)
    if skip:
        return
    # sub_group_size (ref deepspeed_launcher.py:194) caps the tile the optimizer works on; below that cap the shard is
    # cut further so that the copy down, the host update and the copy up of neighbouring tiles overlap
    CH = max(1 << 16, min(_CHUNK, int(cfg.sub_group_size)))
    if not hasattr(engine, "_off_stream"):
        engine._off_stream = torch.cuda.Stream(device=dev)
        nbuf = 2
        gdt = engine.shard_grads.dtype
        engine._off_g = [torch.empty(min(CH, n), dtype=gdt).pin_memory() for _ in range(nbuf)]
        engine._off_p = [torch.empty(min(CH, n), dtype=torch.bfloat16).pin_memory() for _ in range(nbuf)]
    side = engine._off_stream
    nthreads = host_threads(engine.world_size)
    side.wait_stream(torch.cuda.current_stream())
    chunks = [(o, min(o + CH, n)) for o in range(0, n, CH)]
    d2h_evt = [None] * len(chunks)
    h2d_evt = [None] * 2

    def issue_d2h(i):
        lo, hi = chunks[i]
        b = i % 2
        with torch.cuda.stream(side):
            if h2d_evt[b] is not None:
                side.wait_event(h2d_evt[b])
            engine._off_g[b][: hi - lo].copy_(engine.shard_grads[lo:hi], non_blocking=True)
            d2h_evt[i] = torch.cuda.Event()
            d2h_evt[i].record(side)

    issue_d2h(0)
    to_bf16 = engine.compute_dtype == torch.bfloat16
    for i, (lo, hi) in enumerate(chunks):
        if i + 1 < len(chunks):
            issue_d2h(i + 1)
        d2h_evt[i].synchronize()
        b = i % 2
        to_host_shard = getattr(engine, "param_offload", False)     # parameters live on the host too: no copy back
        if to_host_shard and to_bf16:
            outp = engine.shard_params[lo:hi]
        else:
            outp = engine._off_p[b][: hi - lo] if to_bf16 else None
        host_adamw(engine.master[lo:hi], engine.exp_avg[lo:hi], engine.exp_avg_sq[lo:hi], engine._off_g[b][: hi - lo],
                   outp, lr=lr, beta1=cfg.betas[0], beta2=cfg.betas[1], eps=cfg.eps, weight_decay=cfg.weight_decay,
                   step=step_no, grad_scale=coef, threads=nthreads)
        with torch.cuda.stream(side):
            if to_host_shard:
                if not to_bf16:
                    engine.shard_params[lo:hi].copy_(engine.master[lo:hi].to(engine.compute_dtype))
            elif to_bf16:
                engine.shard_params[lo:hi].copy_(outp, non_blocking=True)
            else:
                engine.shard_params[lo:hi].copy_(engine.master[lo:hi].to(engine.compute_dtype), non_blocking=True)
            h2d_evt[b] = torch.cuda.Event()
            h2d_evt[b].record(side)
    torch.cuda.current_stream().wait_stream(side)
