"""NVMe tier of the optimizer-state offload ("ZeRO-Infinity", ``offload_optimizer.device = "nvme"`` —
the reference exposes the enum value, ref ``ai_engine/deepspeed_launcher.py:29-32``, and emits the
offload block ``:197-203``; the capability itself lives in the DeepSpeed wheel's libaio op).

The fp32 master weights and both Adam moments of this rank's shard live in three files under
``<nvme_path>/rank<r>/``. Host memory holds only ``buffer_count`` staging slots of one chunk each.
An optimizer step streams the files through those slots with the C++ async I/O engine
(``csrc/host/aio.cpp``):

    read(chunk i+1)  ||  gradient D2H(chunk i+1)  ||  AdamW(chunk i) on the host  ||  write(chunk i-1)
                                                  ||  new bf16 rows H2D(chunk i-1)

Cold paths (initialisation, checkpoint save/load, ``full_param``) see the same files as memory-mapped
tensors (``engine.master`` etc.), coherent with the buffered pread/pwrite of the hot path through the
page cache, so nothing else in the engine knows the state is on disk.
"""
from __future__ import annotations

import ctypes as C
import os
import tempfile
from pathlib import Path
from typing import Dict, List, Optional

import torch

from ..ops import lib as _lib

STATE_NAMES = ("master", "exp_avg", "exp_avg_sq")
DEFAULT_CHUNK = 1 << 24          # 16 Mi elements: 64 MiB per state per slot


def _aio_lib():
    L = _lib.load_host(required=True)
    if not getattr(L, "_bz_aio_bound", False):
        L.bz_aio_create.restype = C.c_void_p
        L.bz_aio_create.argtypes = [C.c_int, C.c_longlong]
        L.bz_aio_destroy.restype = None
        L.bz_aio_destroy.argtypes = [C.c_void_p]
        L.bz_aio_open.restype = C.c_int
        L.bz_aio_open.argtypes = [C.c_char_p, C.c_int, C.c_longlong, C.c_int, C.POINTER(C.c_int)]
        L.bz_aio_close.restype = C.c_int
        L.bz_aio_close.argtypes = [C.c_int]
        L.bz_aio_fsync.restype = C.c_int
        L.bz_aio_fsync.argtypes = [C.c_int]
        for fn in (L.bz_aio_read, L.bz_aio_write):
            fn.restype = C.c_longlong
            fn.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_longlong, C.c_longlong]
        L.bz_aio_wait.restype = C.c_int
        L.bz_aio_wait.argtypes = [C.c_void_p, C.c_longlong]
        L.bz_aio_wait_all.restype = C.c_int
        L.bz_aio_wait_all.argtypes = [C.c_void_p]
        for fn in (L.bz_aio_bytes_read, L.bz_aio_bytes_written):
            fn.restype = C.c_longlong
            fn.argtypes = [C.c_void_p]
        L._bz_aio_bound = True
    return L


class AsyncFileIO:
    """Thin handle on the C++ engine: ``read``/``write`` return tickets, ``wait`` blocks on one."""

    def __init__(self, threads: int = 8, block_bytes: int = 8 << 20):
        self.L = _aio_lib()
        self.h = self.L.bz_aio_create(threads, block_bytes)
        if not self.h:
            raise RuntimeError("bz_aio_create failed")
        self._fds: List[int] = []

    def open(self, path: os.PathLike, size_bytes: int = 0, create: bool = True, direct: bool = False) -> int:
        used = C.c_int(0)
        fd = self.L.bz_aio_open(str(path).encode(), int(create), int(size_bytes), int(direct), C.byref(used))
        if fd < 0:
            raise OSError(-fd, f"bz_aio_open({path}): {os.strerror(-fd)}")
        self._fds.append(fd)
        self.direct_used = bool(used.value)
        return fd

    def read(self, fd: int, buf: torch.Tensor, byte_offset: int) -> int:
        assert buf.is_contiguous() and not buf.is_cuda
        return self.L.bz_aio_read(self.h, fd, buf.data_ptr(), buf.numel() * buf.element_size(), byte_offset)

    def write(self, fd: int, buf: torch.Tensor, byte_offset: int) -> int:
        assert buf.is_contiguous() and not buf.is_cuda
        return self.L.bz_aio_write(self.h, fd, buf.data_ptr(), buf.numel() * buf.element_size(), byte_offset)

    def wait(self, ticket: int) -> None:
        rc = self.L.bz_aio_wait(self.h, ticket)
        if rc != 0:
            raise OSError(rc if rc > 0 else 0, "short read (EOF)" if rc < 0 else os.strerror(rc))

    def wait_all(self) -> None:
        rc = self.L.bz_aio_wait_all(self.h)
        if rc != 0:
            raise OSError(rc if rc > 0 else 0, "short read (EOF)" if rc < 0 else os.strerror(rc))

    def fsync(self, fd: int) -> None:
        rc = self.L.bz_aio_fsync(fd)
        if rc != 0:
            raise OSError(rc, os.strerror(rc))

    @property
    def bytes_read(self) -> int:
        return int(self.L.bz_aio_bytes_read(self.h))

    @property
    def bytes_written(self) -> int:
        return int(self.L.bz_aio_bytes_written(self.h))

    def close(self) -> None:
        if self.h:
            self.L.bz_aio_wait_all(self.h)
            for fd in self._fds:
                self.L.bz_aio_close(fd)
            self._fds.clear()
            self.L.bz_aio_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def default_nvme_path() -> str:
    return os.environ.get("B200ZERO_NVME_PATH") or os.path.join(tempfile.gettempdir(), "b200zero_nvme")


class NvmeStateStore:
    """File-backed fp32 optimizer state of one rank + the staging slots the step streams through."""

    def __init__(self, root: os.PathLike, rank: int, numel: int, *, chunk: int = DEFAULT_CHUNK, buffer_count: int = 4,
                 threads: int = 8, pin: bool = False, grad_dtype: torch.dtype = torch.float32):
        self.dir = Path(root) / f"rank{rank}"
        self.dir.mkdir(parents=True, exist_ok=True)
        self.numel = int(numel)
        self.chunk = int(min(max(chunk, 1024), max(numel, 1024)))
        self.slots = max(int(buffer_count), 3)      # read-ahead, compute, write-behind
        self.io = AsyncFileIO(threads=threads)
        self.paths = {n: self.dir / f"{n}.f32" for n in STATE_NAMES}
        self.fds = {n: self.io.open(self.paths[n], size_bytes=max(self.numel, 1) * 4) for n in STATE_NAMES}
        # memory-mapped views for the cold paths (shared mapping: writes reach the file)
        self.views: Dict[str, torch.Tensor] = {
            n: torch.from_file(str(self.paths[n]), shared=True, size=max(self.numel, 1), dtype=torch.float32)[: self.numel]
            for n in STATE_NAMES}

        def mk(dtype):
            t = torch.empty(self.chunk, dtype=dtype)
            return t.pin_memory() if pin else t
        self.stage = [{n: mk(torch.float32) for n in STATE_NAMES} for _ in range(self.slots)]
        self.grad_stage = [mk(grad_dtype) for _ in range(self.slots)]
        self.out_stage = [mk(torch.bfloat16) for _ in range(self.slots)]
        self._write_tickets: List[List[int]] = [[] for _ in range(self.slots)]
        self._read_tickets: Dict[int, List[int]] = {}

    # -- cold-path coherence -----------------------------------------------------------------
    def view(self, name: str) -> torch.Tensor:
        return self.views[name]

    def host_bytes(self) -> int:
        per_slot = self.chunk * (3 * 4 + self.grad_stage[0].element_size() + 2)
        return per_slot * self.slots

    # -- hot path ------------------------------------------------------------------------------
    def chunks(self):
        return [(o, min(o + self.chunk, self.numel)) for o in range(0, self.numel, self.chunk)]

    def issue_read(self, i: int, lo: int, hi: int) -> None:
        s = i % self.slots
        for t in self._write_tickets[s]:            # the slot's previous contents must be on disk first
            self.io.wait(t)
        self._write_tickets[s] = []
        self._read_tickets[i] = [self.io.read(self.fds[n], self.stage[s][n][: hi - lo], lo * 4) for n in STATE_NAMES]

    def acquire(self, i: int, lo: int, hi: int):
        for t in self._read_tickets.pop(i):
            self.io.wait(t)
        s = i % self.slots
        return tuple(self.stage[s][n][: hi - lo] for n in STATE_NAMES)

    def release(self, i: int, lo: int, hi: int) -> None:
        s = i % self.slots
        self._write_tickets[s] = [self.io.write(self.fds[n], self.stage[s][n][: hi - lo], lo * 4) for n in STATE_NAMES]

    def drain(self) -> None:
        self.io.wait_all()
        self._write_tickets = [[] for _ in range(self.slots)]
        self._read_tickets.clear()

    def close(self) -> None:
        self.views.clear()
        self.io.close()


def nvme_adam_step(engine, lr: float, step_no: int) -> None:
    """AdamW over the file-backed shard. Works for CUDA engines (pinned staging, side-stream copies)
    and for CPU engines (the gloo test tier)."""
    from .offload import host_adamw
    cfg, store = engine.cfg, engine._nvme
    on_cuda = engine.on_cuda
    if on_cuda:
        host_vals = torch.stack([engine._coef[0], engine._flag[0].float()]).cpu()
    else:
        host_vals = torch.stack([engine._coef[0], engine._flag[0].float()])
    coef, skip = float(host_vals[0]), bool(host_vals[1] > 0)
    if skip:
        return
    chunks = store.chunks()
    if not chunks:
        return
    to_bf16 = engine.compute_dtype == torch.bfloat16
    side = None
    if on_cuda:
        if not hasattr(engine, "_off_stream"):
            engine._off_stream = torch.cuda.Stream(device=engine.device)
        side = engine._off_stream
        side.wait_stream(torch.cuda.current_stream())
    d2h_evt: Dict[int, Optional[torch.cuda.Event]] = {}
    h2d_evt: List[Optional[torch.cuda.Event]] = [None] * store.slots

    def prefetch(i: int) -> None:
        lo, hi = chunks[i]
        s = i % store.slots
        if on_cuda:
            with torch.cuda.stream(side):
                if h2d_evt[s] is not None:          # the slot's previous bf16 rows must have left the host
                    side.wait_event(h2d_evt[s])
                store.grad_stage[s][: hi - lo].copy_(engine.shard_grads[lo:hi], non_blocking=True)
                d2h_evt[i] = torch.cuda.Event()
                d2h_evt[i].record(side)
        store.issue_read(i, lo, hi)

    prefetch(0)
    for i, (lo, hi) in enumerate(chunks):
        if i + 1 < len(chunks):
            prefetch(i + 1)
        s = i % store.slots
        p, m, v = store.acquire(i, lo, hi)
        if on_cuda:
            d2h_evt.pop(i).synchronize()
            g = store.grad_stage[s][: hi - lo]
        else:
            g = engine.shard_grads[lo:hi]
        if on_cuda and h2d_evt[s] is not None:
            h2d_evt[s].synchronize()                # out_stage[s] is about to be overwritten by the host
        outp = store.out_stage[s][: hi - lo] if to_bf16 else None
        host_adamw(p, m, v, g, outp, lr=lr, beta1=cfg.betas[0], beta2=cfg.betas[1], eps=cfg.eps,
                   weight_decay=cfg.weight_decay, step=step_no, grad_scale=coef)
        if on_cuda:
            with torch.cuda.stream(side):
                src = outp if to_bf16 else p.to(engine.compute_dtype)
                engine.shard_params[lo:hi].copy_(src, non_blocking=to_bf16)
                h2d_evt[s] = torch.cuda.Event()
                h2d_evt[s].record(side)
        else:
            engine.shard_params[lo:hi].copy_(outp if to_bf16 else p.to(engine.compute_dtype))
        store.release(i, lo, hi)
    store.drain()
    if on_cuda:
        torch.cuda.current_stream().wait_stream(side)
