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
    import os
    ok, buf = 1, None
    try:
        # the library's own VMM + multicast allocator (csrc/vmm.cu); B200ZERO_SYMM_ALLOC=torch borrows torch's instead
        cls = TorchSymmetricBuffer if os.environ.get("B200ZERO_SYMM_ALLOC", "vmm") == "torch" else VmmSymmetricBuffer
        buf = cls(nbytes, device=device)
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


# ---------------------------------------------------------------------- own VMM / multicast backend (csrc/vmm.cu)
def exchange_fds(my_fds: List[int], rank: int, world: int, group=None, tag: str = "") -> List[List[int]]:
    """All-to-all of open file descriptors between the ranks of one node: ``result[r]`` are rank ``r``'s
    descriptors, duplicated into this process (``result[rank]`` is ``my_fds`` itself). Descriptors cannot travel
    through torch.distributed; they go over unix-domain sockets as SCM_RIGHTS ancillary data, one listening socket
    per rank in the abstract namespace (no files to clean up)."""
    import os
    import socket
    import threading
    import uuid
    box = [uuid.uuid4().hex if rank == 0 else None]
    dist.broadcast_object_list(box, src=dist.get_global_rank(group, 0) if group is not None else 0, group=group)
    name = lambda r: f"\0b200zero-{box[0]}-{tag}-{r}"      # noqa: E731
    srv = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
    srv.bind(name(rank))
    srv.listen(world)
    got: List[Optional[List[int]]] = [None] * world
    got[rank] = list(my_fds)
    err: List[BaseException] = []

    def accept_all():
        try:
            for _ in range(world - 1):
                conn, _addr = srv.accept()
                with conn:
                    msg, fds, _flags, _a = socket.recv_fds(conn, 16, len(my_fds))
                    got[int(msg.decode())] = list(fds)
        except BaseException as e:  # noqa: BLE001
            err.append(e)

    th = threading.Thread(target=accept_all, daemon=True)
    th.start()
    dist.barrier(group=group)                              # every rank is listening
    for r in range(world):
        if r == rank:
            continue
        with socket.socket(socket.AF_UNIX, socket.SOCK_STREAM) as c:
            c.connect(name(r))
            socket.send_fds(c, [str(rank).encode()], list(my_fds))
    th.join(timeout=120)
    srv.close()
    if err or any(g is None for g in got):
        for g in got:
            for fd in (g or []):
                if g is not got[rank]:
                    os.close(fd)
        raise RuntimeError(f"file-descriptor exchange failed: {err or 'timeout'}")
    return got  # type: ignore[return-value]


class VmmSymmetricBuffer:
    """Same interface as :class:`SymmetricBuffer` on the library's own VMM allocator (``csrc/vmm.cu``): the buffer is
    one ``cuMemCreate`` allocation per rank, exported as a file descriptor and mapped by every peer; with
    ``multicast=True`` and a fabric that supports it, all ranks' allocations are additionally bound into one NVSwitch
    multicast object and ``multicast_ptr`` is its address on this rank (0 otherwise)."""

    def __init__(self, nbytes: int, group=None, device: Optional[torch.device] = None, multicast: bool = True):
        import os
        self.L = L = _lib.load(required=True)
        u64, vp = C.c_ulonglong, C.c_void_p
        self.group = group
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.device = device or torch.device("cuda", torch.cuda.current_device())
        self.multicast_ptr = 0
        self._mc_handle = 0
        self._peer_maps: List[tuple] = []
        with torch.cuda.device(self.device):
            has_vmm, has_mc = C.c_int(0), C.c_int(0)
            _check(L, L.bz_vmm_supported(C.byref(has_vmm), C.byref(has_mc)), "vmm_supported")
            if not has_vmm.value:
                raise RuntimeError("this device / driver cannot export VMM allocations as file descriptors")
            size = u64(0)
            _check(L, L.bz_vmm_round_size(u64(int(nbytes)), self.world, C.byref(size)), "vmm_round_size")
            self.nbytes = int(size.value)
            ptr, handle, fd = vp(), u64(0), C.c_int(-1)
            _check(L, L.bz_vmm_alloc(u64(self.nbytes), C.byref(ptr), C.byref(handle), C.byref(fd)), "vmm_alloc")
        self.local_ptr, self._handle, self._fd = int(ptr.value), int(handle.value), int(fd.value)
        self._holder = _DevMem(self.local_ptr, self.nbytes)
        self._bytes = torch.as_tensor(self._holder, device=self.device)
        self.peer_ptrs: List[int] = [self.local_ptr] * self.world
        if self.world > 1:
            fds = exchange_fds([self._fd], self.rank, self.world, group, tag="mem")
            with torch.cuda.device(self.device):
                for r in range(self.world):
                    if r == self.rank:
                        continue
                    p, h = vp(), u64(0)
                    _check(L, L.bz_vmm_import(fds[r][0], u64(self.nbytes), C.byref(p), C.byref(h)), f"vmm_import(rank {r})")
                    os.close(fds[r][0])
                    self.peer_ptrs[r] = int(p.value)
                    self._peer_maps.append((int(p.value), int(h.value)))
            want_mc = torch.tensor([int(bool(multicast and has_mc.value))], device=self.device, dtype=torch.int32)
            dist.all_reduce(want_mc, op=dist.ReduceOp.MIN, group=group)
            if bool(want_mc.item()):
                self._setup_multicast()
            dist.barrier(group=group)

    def _setup_multicast(self) -> None:
        import os
        L, u64, vp = self.L, C.c_ulonglong, C.c_void_p
        mch, mfd = u64(0), C.c_int(-1)
        with torch.cuda.device(self.device):
            if self.rank == 0:
                _check(L, L.bz_mc_create(self.world, u64(self.nbytes), C.byref(mch), C.byref(mfd)), "mc_create")
            fds = exchange_fds([mfd.value] if self.rank == 0 else [self._fd], self.rank, self.world, self.group, tag="mc")
            if self.rank != 0:
                _check(L, L.bz_mc_import(fds[0][0], C.byref(mch)), "mc_import")
            for r in range(self.world):
                if r != self.rank:
                    os.close(fds[r][0])
            if self.rank == 0:
                os.close(mfd.value)
            _check(L, L.bz_mc_add_device(mch), "mc_add_device")
            dist.barrier(group=self.group)                 # every device is part of the object before anyone binds
            mptr = vp()
            _check(L, L.bz_mc_bind_and_map(mch, u64(self._handle), u64(self.nbytes), C.byref(mptr)), "mc_bind_and_map")
        self._mc_handle, self.multicast_ptr = int(mch.value), int(mptr.value)

    tensor = SymmetricBuffer.tensor
    ptr_array = SymmetricBuffer.ptr_array

    def close(self) -> None:
        import os
        if self.world > 1 and dist.is_initialized():
            try:
                torch.cuda.synchronize(self.device)
                dist.barrier(group=self.group)
            except Exception:
                pass
        L, u64 = self.L, C.c_ulonglong
        with torch.cuda.device(self.device):
            if self._mc_handle:
                L.bz_mc_unmap(C.c_void_p(self.multicast_ptr), u64(self.nbytes), u64(self._mc_handle))
                self._mc_handle, self.multicast_ptr = 0, 0
            for p, h in self._peer_maps:
                L.bz_vmm_unmap(C.c_void_p(p), u64(self.nbytes), u64(h))
            self._peer_maps = []
            if self.local_ptr:
                self._bytes = None
                L.bz_vmm_unmap(C.c_void_p(self.local_ptr), u64(self.nbytes), u64(self._handle))
                os.close(self._fd)
                self.local_ptr = 0
