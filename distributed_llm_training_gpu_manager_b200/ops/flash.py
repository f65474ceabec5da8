"""ctypes front-end of the tcgen05 flash-attention kernels (``csrc/attention.cu``)."""
from __future__ import annotations

from typing import Optional, Tuple

import torch

from . import lib as _lib

_lib.register("bz_flash_attn_fwd", _lib.i32, [_lib.vp, _lib.vp, _lib.vp, _lib.i32, _lib.i32, _lib.i32, _lib.i32,
                                              _lib.i32, _lib.f32, _lib.vp])


def flash_fwd(qkv: torch.Tensor, B: int, S: int, Hq: int, Hk: int, hd: int,
              want_lse: bool = True) -> Tuple[torch.Tensor, Optional[torch.Tensor]]:
    """Causal GQA attention forward. Returns (out [T, Hq*hd] bf16, lse [Hq, T] fp32)."""
    T = B * S
    assert qkv.is_contiguous() and qkv.shape == (T, (Hq + 2 * Hk) * hd) and qkv.dtype == torch.bfloat16
    out = torch.empty((T, Hq * hd), dtype=torch.bfloat16, device=qkv.device)
    lse = torch.empty((Hq, T), dtype=torch.float32, device=qkv.device) if want_lse else None
    L = _lib.load(required=True)
    _lib.check(L.bz_flash_attn_fwd(qkv.data_ptr(), out.data_ptr(), _lib.ptr(lse), B, S, Hq, Hk, hd, hd ** -0.5,
                                   _lib.stream()), "flash_attn_fwd")
    return out, lse


def has_bwd() -> bool:
    L = _lib.load()
    return L is not None and hasattr(L, "bz_flash_attn_bwd")
