"""ctypes binding of ``libb200zero.so`` (C ABI declared in ``include/b200zero.h``).

Design: the kernel library has no Python/torch dependency — it takes raw device
pointers and a stream handle. This module loads it, declares argument types and
offers small helpers (``ptr``, ``stream``, ``check``). On a machine with a GPU a
missing library is a hard error (never a silent eager fallback); on a CPU-only
machine ``available()`` is False and the ops fall back to their fp32 PyTorch
oracles so the control plane and the gloo test tier still run.
"""
from __future__ import annotations

import ctypes as C
import os
import threading
from pathlib import Path
from typing import Optional

import torch

_OPS_DIR = Path(__file__).resolve().parent
_LIB_PATH = _OPS_DIR / "libb200zero.so"
_HOST_LIB_PATH = _OPS_DIR / "libb200zero_host.so"

_lock = threading.Lock()
_lib: Optional[C.CDLL] = None
_host_lib: Optional[C.CDLL] = None
_launches = 0  # kernels launched through this binding (bench.py reports it)

vp, i32, i64, f32, u64 = C.c_void_p, C.c_int, C.c_longlong, C.c_float, C.c_ulonglong

# name -> (restype, argtypes); kept next to the header so drift is obvious.
_SIGNATURES = {
    "bz_last_error": (C.c_char_p, []),
    "bz_version": (i32, []),
    "bz_sm_count": (i32, []),
    "bz_gemm_bf16": (i32, [vp, vp, vp, i32, i32, i32, i32, i32, i32, i32, i32, i32, i32, i32, vp]),
    "bz_rmsnorm_fwd": (i32, [vp, vp, vp, vp, vp, vp, i32, i32, f32, vp]),
    "bz_rmsnorm_bwd_workspace_floats": (i32, [i32, i32]),
    "bz_rmsnorm_bwd": (i32, [vp, vp, vp, vp, vp, vp, vp, i32, i32, vp, i32, i32, f32, vp]),
    "bz_rope": (i32, [vp, vp, vp, i32, i32, i32, i32, i32, i32, vp]),
    "bz_swiglu_fwd": (i32, [vp, vp, i64, i32, vp]),
    "bz_swiglu_bwd": (i32, [vp, vp, vp, i64, i32, vp]),
    "bz_embedding_fwd": (i32, [vp, i32, vp, vp, i64, i32, i32, vp]),
    "bz_embedding_bwd": (i32, [vp, i32, vp, vp, i32, i64, i32, i32, vp]),
    "bz_cross_entropy": (i32, [vp, vp, i32, vp, vp, i32, i32, i64, f32, i32, vp]),
    "bz_add_bf16": (i32, [vp, vp, vp, i64, vp]),
    "bz_adamw": (i32, [vp, vp, vp, vp, i32, vp, i64, f32, f32, f32, f32, f32, i32, f32, vp, vp, vp]),
    "bz_sumsq_workspace_floats": (i32, []),
    "bz_sumsq": (i32, [vp, i32, i64, vp, vp, i32, vp, vp]),
    "bz_clip_coef": (i32, [vp, f32, f32, vp, vp, vp, vp]),
    "bz_cast_f32_to_bf16": (i32, [vp, vp, i64, vp]),
    "bz_cast_bf16_to_f32": (i32, [vp, vp, i64, vp]),
    "bz_comm_max_world": (i32, []),
    "bz_comm_signal_pad_bytes": (i32, []),
    "bz_comm_segment_bytes": (i32, []),
    "bz_fused_comm_workspace_bytes": (i32, []),
    "bz_symm_alloc": (i32, [u64, C.POINTER(vp)]),
    "bz_symm_free": (i32, [vp]),
    "bz_ipc_export": (i32, [vp, C.c_char_p]),
    "bz_ipc_import": (i32, [C.c_char_p, C.POINTER(vp)]),
    "bz_ipc_close": (i32, [vp]),
    "bz_enable_peer_access": (i32, [i32]),
    "bz_fused_reduce_scatter": (i32, [C.POINTER(vp), C.POINTER(vp), vp, i32, vp, i32, f32, vp, vp, vp, vp, C.c_uint,
                                      i32, i32, i32, i32, i32, vp]),
    "bz_fused_adam_push": (i32, [C.POINTER(vp), C.POINTER(vp), vp, i32, vp, vp, vp, vp, i32, vp, f32, f32, f32, f32,
                                 f32, i32, vp, vp, vp, C.c_uint, i32, i32, i32, i32, i32, i32, i32, vp]),
    "bz_fused_reduce_scatter_mc": (i32, [C.POINTER(vp), C.POINTER(vp), vp, vp, i32, vp, i32, f32, vp, vp, vp, vp,
                                         C.c_uint, i32, i32, i32, i32, i32, vp]),
    "bz_pull_allgather": (i32, [C.POINTER(vp), vp, i32, vp, i32, i32, i32, vp]),
    "bz_pull_allgather_v2": (i32, [C.POINTER(vp), vp, i32, vp, i32, i32, i32, vp]),
    "bz_peer_allreduce4": (i32, [C.POINTER(vp), vp, C.c_uint, i32, i32, i32, vp]),
}


class NativeLibraryMissing(RuntimeError):
    pass


def cuda_present() -> bool:
    return torch.cuda.is_available()


def register(name: str, restype, argtypes) -> None:
    """Allow other modules to declare additional entry points."""
    _SIGNATURES[name] = (restype, argtypes)
    if _lib is not None and hasattr(_lib, name):
        fn = getattr(_lib, name)
        fn.restype, fn.argtypes = restype, argtypes


def _bind(lib: C.CDLL) -> None:
    for name, (res, args) in _SIGNATURES.items():
        try:
            fn = getattr(lib, name)
        except AttributeError:
            continue
        fn.restype, fn.argtypes = res, args


def load(required: Optional[bool] = None) -> Optional[C.CDLL]:
    """Load the CUDA kernel library. ``required`` defaults to ``cuda_present()``."""
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is not None:
            return _lib
        if required is None:
            required = cuda_present()
        if not _LIB_PATH.exists():
            if os.environ.get("B200ZERO_AUTOBUILD", "1") == "1" and required:
                from . import build as _build
                _build.build_cuda()
            if not _LIB_PATH.exists():
                if required:
                    raise NativeLibraryMissing(
                        f"{_LIB_PATH} is missing on a CUDA machine. Run `python -m "
                        f"distributed_llm_training_gpu_manager_b200.ops.build` (or __graft_entry__.build()).")
                return None
        try:
            lib = C.CDLL(str(_LIB_PATH), mode=C.RTLD_GLOBAL)
        except OSError as e:  # e.g. libcudart missing on a CPU box
            if required:
                raise NativeLibraryMissing(f"cannot load {_LIB_PATH}: {e}") from e
            return None
        _bind(lib)
        _lib = lib
        return _lib


def load_host(required: bool = False) -> Optional[C.CDLL]:
    global _host_lib
    if _host_lib is not None:
        return _host_lib
    with _lock:
        if _host_lib is not None:
            return _host_lib
        if not _HOST_LIB_PATH.exists():
            try:
                from . import build as _build
                _build.build_host()
            except Exception:
                if required:
                    raise
        if not _HOST_LIB_PATH.exists():
            if required:
                raise NativeLibraryMissing(f"{_HOST_LIB_PATH} is missing")
            return None
        _host_lib = C.CDLL(str(_HOST_LIB_PATH))
        return _host_lib


def available() -> bool:
    """True when the sm_100a kernels can actually run (GPU present + library loaded)."""
    return cuda_present() and load() is not None


def ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


def stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def check(rc: int, what: str = "") -> None:
    global _launches
    if rc != 0:
        msg = _lib.bz_last_error().decode(errors="replace") if _lib is not None else "?"
        raise RuntimeError(f"b200zero {what} failed (rc={rc}): {msg}")
    _launches += 1


def launches() -> int:
    return _launches


def add_launches(n: int) -> None:
    global _launches
    _launches += n
