"""Symmetric (peer-mapped) device memory for one-process-per-GPU jobs (H2).

Each rank allocates a buffer with ``cudaMalloc`` through the kernel library, exports a CUDA IPC
handle, all-gathers the 64-byte handles over the existing ``torch.distributed`` group (NCCL is
only the bootstrap here), and opens every peer's handle — after which kernels on this GPU can
issue plain ``ld.global`` / ``st.global`` on peer addresses over NVLink 5 / NVSwitch. The
reference has no communication layer at all.

``SymmetricBuffer.tensor()`` exposes the local allocation to PyTorch zero-copy through
``__cuda_array_interface__``; the engine places gradient buckets and parameter buffers there so
the fused kernels can reach them from any rank.
"""
from __future__ import annotations

import ctypes as C
from typing import List, Optional

import torch
import torch.distributed as dist

from ..ops import lib as _lib


class _DevMem:
    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 3,
                                         "strides": None}


class SymmetricBuffer:
    """``nbytes`` of device memory on every rank of ``group``, mutually mapped."""

    def __init__(self, nbytes: int, group=None, device: Optional[torch.device] = None):
        self.L = _lib.load(required=True)
        self.nbytes = int((nbytes + 255) // 256 * 256)
        self.group = group
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.device = device or torch.device("cuda", torch.cuda.current_device())
        out = C.c_void_p()
        with torch.cuda.device(self.device):
            _check(self.L, self.L.bz_symm_alloc(self.nbytes, C.byref(out)), "symm_alloc")
        self.local_ptr = int(out.value)
        self._holder = _DevMem(self.local_ptr, self.nbytes)
        self._bytes = torch.as_tensor(self._holder, device=self.device)
        self.peer_ptrs: List[int] = [self.local_ptr] * self.world
        self._opened: List[int] = []
        if self.world > 1:
            handle = C.create_string_buffer(64)
            _check(self.L, self.L.bz_ipc_export(self.local_ptr, handle), "ipc_export")
            handles: List[Optional[bytes]] = [None] * self.world
            dist.all_gather_object(handles, handle.raw, group=group)
            for r, h in enumerate(handles):
                if r == self.rank:
                    continue
                p = C.c_void_p()
                with torch.cuda.device(self.device):
                    _check(self.L, self.L.bz_ipc_import(h, C.byref(p)), f"ipc_import(rank {r})")
                self.peer_ptrs[r] = int(p.value)
                self._opened.append(int(p.value))
            dist.barrier(group=group)

    def tensor(self, dtype: torch.dtype, numel: Optional[int] = None, byte_offset: int = 0) -> torch.Tensor:
        es = torch.empty((), dtype=dtype).element_size()
        if numel is None:
            numel = (self.nbytes - byte_offset) // es
        return self._bytes[byte_offset:byte_offset + numel * es].view(dtype)

    def ptr_array(self, byte_offset: int = 0):
        """``void*[world]`` (ctypes) of every rank's mapping, shifted by ``byte_offset``."""
        arr = (C.c_void_p * self.world)()
        for r in range(self.world):
            arr[r] = self.peer_ptrs[r] + byte_offset
        return arr

    def close(self) -> None:
        if self.world > 1 and dist.is_initialized():
            try:
                torch.cuda.synchronize(self.device)
                dist.barrier(group=self.group)
            except Exception:
                pass
        for p in self._opened:
            self.L.bz_ipc_close(p)
        self._opened = []
        if self.local_ptr:
            self._bytes = None
            self.L.bz_symm_free(self.local_ptr)
            self.local_ptr = 0


def _check(L, rc: int, what: str) -> None:
    if rc != 0:
        raise RuntimeError(f"b200zero {what} failed (rc={rc}): {L.bz_last_error().decode(errors='replace')}")


class TorchSymmetricBuffer:
    """Same interface as :class:`SymmetricBuffer`, but allocated and rendezvoused through
    ``torch.distributed._symmetric_memory`` (CUDA VMM + fabric/fd handle exchange), which also binds
    the allocation into an NVLS multicast object when the fabric supports it. Only allocation and
    address exchange are borrowed; every kernel that touches the memory is ours. ``multicast_ptr`` is
    0 when NVLS is unavailable."""

    def __init__(self, nbytes: int, group=None, device: Optional[torch.device] = None):
        import torch.distributed._symmetric_memory as symm_mem
        self.nbytes = int((nbytes + 255) // 256 * 256)
        self.device = device or torch.device("cuda", torch.cuda.current_device())
        self.group = group if group is not None else dist.group.WORLD
        self.rank = dist.get_rank(self.group)
        self.world = dist.get_world_size(self.group)
        self._bytes = symm_mem.empty(self.nbytes, dtype=torch.uint8, device=self.device)
        self._hdl = symm_mem.rendezvous(self._bytes, self.group)
        self._bytes.zero_()
        self.peer_ptrs: List[int] = [int(p) for p in self._hdl.buffer_ptrs]
        self.local_ptr = self.peer_ptrs[self.rank]
        try:
            self.multicast_ptr = int(self._hdl.multicast_ptr or 0)
        except Exception:
            self.multicast_ptr = 0
        torch.cuda.synchronize(self.device)
        dist.barrier(group=self.group)

    def tensor(self, dtype: torch.dtype, numel: Optional[int] = None, byte_offset: int = 0) -> torch.Tensor:
        es = torch.empty((), dtype=dtype).element_size()
        if numel is None:
            numel = (self.nbytes - byte_offset) // es
        return self._bytes[byte_offset:byte_offset + numel * es].view(dtype)

    def ptr_array(self, byte_offset: int = 0):
        arr = (C.c_void_p * self.world)()
        for r in range(self.world):
            arr[r] = self.peer_ptrs[r] + byte_offset
        return arr

    def close(self) -> None:
        self._hdl = None
        self._bytes = None


def nvls_buffer_or_none(nbytes: int, device: torch.device):
    """A multicast-capable symmetric buffer, or None when NVLS cannot be used (collective call: every
    rank gets the same answer)."""
    ok, buf = 1, None
    try:
        buf = TorchSymmetricBuffer(nbytes, device=device)
        if not buf.multicast_ptr:
            ok = 0
    except Exception:
        ok = 0
    t = torch.tensor([ok], device=device, dtype=torch.int32)
    dist.all_reduce(t, op=dist.ReduceOp.MIN)
    if not bool(t.item()):
        if buf is not None:
            buf.close()
        return None
    return buf
