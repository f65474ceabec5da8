"""ctypes front-end of the tcgen05 flash-attention kernels (``csrc/attention.cu``)."""
from __future__ import annotations

from typing import Optional, Tuple

import torch

from . import lib as _lib

_lib.register("bz_flash_attn_fwd", _lib.i32, [_lib.vp, _lib.vp, _lib.vp, _lib.i32, _lib.i32, _lib.i32, _lib.i32,
                                              _lib.i32, _lib.f32, _lib.vp])


_lib.register("bz_flash_attn_fwd_v3", _lib.i32, [_lib.vp, _lib.vp, _lib.vp, _lib.vp, _lib.i32, _lib.i32, _lib.i32,
                                                 _lib.i32, _lib.i32, _lib.f32, _lib.i32, _lib.vp])
_err_flag = {}


def flash_fwd(qkv: torch.Tensor, B: int, S: int, Hq: int, Hk: int, hd: int,
              want_lse: bool = True, version: Optional[int] = None) -> Tuple[torch.Tensor, Optional[torch.Tensor]]:
    """Causal GQA attention forward. Returns (out [T, Hq*hd] bf16, lse [Hq, T] fp32).
    ``version`` (default ``$B200ZERO_FLASH_FWD`` or 3): 3 = the double-buffered-P kernel of
    ``csrc/attention_v3.cu``; 32 / 34 / 38 = the same with every 2nd / 4th / 8th exponential evaluated on the FMA
    pipe instead of the SFU. Measured on B200 (B=2, S=4096, 32q/8kv, d=128): 836 TFLOP/s, 38: 827, 34: 832,
    32: 767; the first-generation kernel (removed) was at 686 (profiles/kernel_checks_r2.txt)."""
    import os
    T = B * S
    assert qkv.is_contiguous() and qkv.shape == (T, (Hq + 2 * Hk) * hd) and qkv.dtype == torch.bfloat16
    out = torch.empty((T, Hq * hd), dtype=torch.bfloat16, device=qkv.device)
    lse = torch.empty((Hq, T), dtype=torch.float32, device=qkv.device) if want_lse else None
    L = _lib.load(required=True)
    if version is None:
        version = int(os.environ.get("B200ZERO_FLASH_FWD", "3"))
    if version in (3, 32, 34, 38):
        key = str(qkv.device)
        if key not in _err_flag:
            _err_flag[key] = torch.zeros(1, dtype=torch.int32, device=qkv.device)
        _lib.check(L.bz_flash_attn_fwd_v3(qkv.data_ptr(), out.data_ptr(), _lib.ptr(lse), _err_flag[key].data_ptr(), B, S,
                                          Hq, Hk, hd, hd ** -0.5, version % 10 if version > 3 else 0, _lib.stream()),
                   "flash_attn_fwd_v3")
        return out, lse
    raise ValueError(f"flash_fwd: unknown kernel version {version} (3, 32, 34 or 38; the first generation was removed)")


_lib.register("bz_flash_attn_bwd_workspace_floats", _lib.i64, [_lib.i32, _lib.i32, _lib.i32])
_lib.register("bz_flash_attn_bwd", _lib.i32, [_lib.vp, _lib.vp, _lib.vp, _lib.vp, _lib.vp, _lib.vp, _lib.i32, _lib.i32,
                                              _lib.i32, _lib.i32, _lib.i32, _lib.f32, _lib.vp])
_ws = {}


_lib.register("bz_flash_attn_bwd_v3", _lib.i32, [_lib.vp, _lib.vp, _lib.vp, _lib.vp, _lib.vp, _lib.vp, _lib.vp, _lib.i32,
                                                 _lib.i32, _lib.i32, _lib.i32, _lib.i32, _lib.f32, _lib.vp])


def flash_bwd(qkv: torch.Tensor, out: torch.Tensor, lse: torch.Tensor, dout: torch.Tensor, B: int, S: int, Hq: int,
              Hk: int, hd: int, dqkv: Optional[torch.Tensor] = None, version: Optional[int] = None) -> torch.Tensor:
    """Gradient w.r.t. the fused QKV buffer (same layout as ``qkv``). ``version`` (default ``$B200ZERO_FLASH_BWD`` or
    3): the pipelined schedule of ``csrc/attention_bwd_v3.cu``. Measured on B200 at the shape above: 711 TFLOP/s (the
    removed first-generation schedule: 508)."""
    import os
    T = B * S
    L = _lib.load(required=True)
    if dqkv is None:
        dqkv = torch.empty_like(qkv)
    need = L.bz_flash_attn_bwd_workspace_floats(B, S, Hq)
    key = str(qkv.device)
    if key not in _ws or _ws[key].numel() < need:
        _ws[key] = torch.empty(need, dtype=torch.float32, device=qkv.device)
    assert dout.is_contiguous() and out.is_contiguous() and dqkv.is_contiguous()
    if version is None:
        version = int(os.environ.get("B200ZERO_FLASH_BWD", "3"))
    if version == 3:
        if key not in _err_flag:
            _err_flag[key] = torch.zeros(1, dtype=torch.int32, device=qkv.device)
        _lib.check(L.bz_flash_attn_bwd_v3(qkv.data_ptr(), out.data_ptr(), lse.data_ptr(), dout.data_ptr(), dqkv.data_ptr(),
                                          _ws[key].data_ptr(), _err_flag[key].data_ptr(), B, S, Hq, Hk, hd, hd ** -0.5,
                                          _lib.stream()), "flash_attn_bwd_v3")
        _lib.add_launches(2)
        return dqkv
    raise ValueError(f"flash_bwd: unknown kernel version {version} (3; the first generation was removed)")


def has_bwd() -> bool:
    L = _lib.load()
    return L is not None and hasattr(L, "bz_flash_attn_bwd")
