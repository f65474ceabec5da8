"""All-gather→GEMM (K2) front-end: ``y = x @ gather(W)ᵀ`` where ``W`` is row-partitioned over the
data-parallel ranks and lives in peer-mapped parameter shards (``csrc/ag_gemm.cu``). The gathered
weight is written through to ``w_full`` and stays valid after the call."""
from __future__ import annotations

import ctypes as C
from typing import Optional, Sequence

import torch

from . import lib as _lib

_lib.register("bz_ag_gemm_bf16", _lib.i32, [_lib.vp, C.POINTER(_lib.vp), _lib.vp, _lib.vp, _lib.vp, _lib.i32, _lib.i32,
                                            _lib.i32, _lib.i32, _lib.i32, _lib.i32, _lib.i32, _lib.i32, _lib.vp])
_lib.register("bz_ag_gemm_max_tiles", _lib.i32, [])

_scratch = {}


def _chunk_scratch(device) -> torch.Tensor:
    key = str(device)
    if key not in _scratch:
        _scratch[key] = torch.zeros(4096, dtype=torch.int32, device=device)
    return _scratch[key]


def ag_linear_fwd(x: torch.Tensor, shard_ptrs, w_full: torch.Tensor, rank: int, ws: int,
                  out: Optional[torch.Tensor] = None, cta_group: int = 0) -> torch.Tensor:
    """``shard_ptrs``: ctypes ``void*[ws]`` — rank p's [N/ws, K] rows of W (peer-mapped device
    pointers; entry ``rank`` is local). ``w_full``: local [N, K] destination of the gather."""
    M, K = x.shape
    N = w_full.shape[0]
    assert w_full.shape == (N, K) and w_full.is_contiguous() and x.stride(1) == 1
    out = out if out is not None else torch.empty((M, N), dtype=torch.bfloat16, device=x.device)
    L = _lib.load(required=True)
    rc = L.bz_ag_gemm_bf16(x.data_ptr(), shard_ptrs, w_full.data_ptr(), out.data_ptr(),
                           _chunk_scratch(x.device).data_ptr(), M, N, K, x.stride(0), out.stride(0), rank, ws,
                           cta_group, _lib.stream())
    _lib.check(rc, "ag_gemm")
    return out


_lib.register("bz_ag_gemm_dgrad_bf16", _lib.i32, [_lib.vp, C.POINTER(_lib.vp), _lib.vp, _lib.vp, _lib.vp, _lib.i32,
                                                  _lib.i32, _lib.i32, _lib.i32, _lib.i32, _lib.i32, _lib.i32, _lib.i32,
                                                  _lib.vp])


def ag_linear_dgrad(dy: torch.Tensor, shard_ptrs, w_full: torch.Tensor, rank: int, ws: int,
                    out: Optional[torch.Tensor] = None, cta_group: int = 0) -> torch.Tensor:
    """``dx = dy @ gather(W)`` for a row-partitioned ``W`` [N, K] (the partitioned dimension is the GEMM's
    reduction dimension; ``csrc/ag_gemm_dgrad.cu``). ``w_full`` receives the gathered weight."""
    M, N = dy.shape
    K = w_full.shape[1]
    assert w_full.shape == (N, K) and w_full.is_contiguous() and dy.stride(1) == 1
    out = out if out is not None else torch.empty((M, K), dtype=torch.bfloat16, device=dy.device)
    L = _lib.load(required=True)
    rc = L.bz_ag_gemm_dgrad_bf16(dy.data_ptr(), shard_ptrs, w_full.data_ptr(), out.data_ptr(),
                                 _chunk_scratch(dy.device).data_ptr(), M, N, K, dy.stride(0), out.stride(0), rank, ws,
                                 cta_group, _lib.stream())
    _lib.check(rc, "ag_gemm_dgrad")
    return out


def ptr_array(ptrs: Sequence[int]):
    return (C.c_void_p * len(ptrs))(*ptrs)
