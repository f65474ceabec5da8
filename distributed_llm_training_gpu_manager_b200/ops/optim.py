"""Optimizer-side kernels: fused AdamW on flat partitions, ‖g‖² + isfinite, on-device clip
coefficient, dtype casts (``csrc/optim.cu``). CPU bodies are fp32 PyTorch restatements
(and the oracle for tests)."""
from __future__ import annotations

import math
from typing import Optional

import torch

from . import lib as _lib
from .elementwise import _native, _workspace


def adamw_ref(master, m, v, grad, lr, beta1, beta2, eps, wd, step, grad_scale=1.0):
    g = grad.float() * grad_scale
    m.mul_(beta1).add_(g, alpha=1 - beta1)
    v.mul_(beta2).addcmul_(g, g, value=1 - beta2)
    bc1, bc2 = 1 - beta1 ** step, 1 - beta2 ** step
    denom = v.sqrt() / math.sqrt(bc2) + eps
    master.mul_(1 - lr * wd).addcdiv_(m, denom, value=-lr / bc1)


def adamw_step(master: torch.Tensor, m: torch.Tensor, v: torch.Tensor, grad: torch.Tensor, *, lr: float,
               beta1: float = 0.9, beta2: float = 0.999, eps: float = 1e-8, weight_decay: float = 0.01, step: int,
               grad_scale: float = 1.0, grad_scale_dev: Optional[torch.Tensor] = None,
               skip_flag: Optional[torch.Tensor] = None, param_out: Optional[torch.Tensor] = None) -> None:
    """In-place AdamW on flat fp32 ``master/m/v``; writes the low-precision copy to
    ``param_out``. ``grad_scale_dev`` (1-elem fp32) and ``skip_flag`` (1-elem int32) are read on
    the device, so clipping and the non-finite skip need no host sync."""
    n = master.numel()
    if not _native(master, m, v, grad, param_out):
        if skip_flag is not None and int(skip_flag.item()) != 0:
            return
        gs = grad_scale * (float(grad_scale_dev.item()) if grad_scale_dev is not None else 1.0)
        adamw_ref(master, m, v, grad, lr, beta1, beta2, eps, weight_decay, step, gs)
        if param_out is not None:
            param_out.copy_(master.to(param_out.dtype))
        return
    assert param_out is None or param_out.dtype == torch.bfloat16
    L = _lib.load()
    _lib.check(L.bz_adamw(master.data_ptr(), m.data_ptr(), v.data_ptr(), grad.data_ptr(),
                          int(grad.dtype == torch.float32), _lib.ptr(param_out), n, lr, beta1, beta2, eps,
                          weight_decay, step, grad_scale, _lib.ptr(grad_scale_dev), _lib.ptr(skip_flag),
                          _lib.stream()), "adamw")


def sumsq(g: torch.Tensor, out: torch.Tensor, accumulate: bool = False,
          nonfinite_flag: Optional[torch.Tensor] = None) -> torch.Tensor:
    """out[0] (+)= Σ g²; raises ``nonfinite_flag[0]`` (int32) if the sum is NaN/Inf."""
    if not _native(g, out, nonfinite_flag):
        s = g.float().pow(2).sum()
        if nonfinite_flag is not None and not bool(torch.isfinite(s)):
            nonfinite_flag.fill_(1)
        out[0] = (out[0] + s) if accumulate else s
        return out
    L = _lib.load()
    ws = _workspace(g.device, L.bz_sumsq_workspace_floats(), "sumsq")
    _lib.check(L.bz_sumsq(g.data_ptr(), int(g.dtype == torch.float32), g.numel(), ws.data_ptr(), out.data_ptr(),
                          int(accumulate), _lib.ptr(nonfinite_flag), _lib.stream()), "sumsq")
    _lib.add_launches(1)
    return out


def clip_coef(sumsq_t: torch.Tensor, max_norm: float, extra_scale: float, coef_out: torch.Tensor,
              norm_out: Optional[torch.Tensor] = None, nonfinite_flag: Optional[torch.Tensor] = None) -> torch.Tensor:
    if not _native(sumsq_t, coef_out):
        nrm = sumsq_t[0].float().sqrt()
        if norm_out is not None:
            norm_out[0] = nrm
        if not bool(torch.isfinite(nrm)):
            if nonfinite_flag is not None:
                nonfinite_flag.fill_(1)
            coef_out[0] = 0.0
            return coef_out
        c = min(1.0, max_norm / (float(nrm) + 1e-6)) if max_norm > 0 else 1.0
        coef_out[0] = c * extra_scale
        return coef_out
    L = _lib.load()
    _lib.check(L.bz_clip_coef(sumsq_t.data_ptr(), max_norm, extra_scale, coef_out.data_ptr(), _lib.ptr(norm_out),
                              _lib.ptr(nonfinite_flag), _lib.stream()), "clip_coef")
    return coef_out


def cast_to_bf16(src: torch.Tensor, dst: torch.Tensor) -> torch.Tensor:
    if not _native(src, dst):
        dst.copy_(src.to(dst.dtype))
        return dst
    L = _lib.load()
    _lib.check(L.bz_cast_f32_to_bf16(src.data_ptr(), dst.data_ptr(), src.numel(), _lib.stream()), "cast_bf16")
    return dst


def cast_to_f32(src: torch.Tensor, dst: torch.Tensor) -> torch.Tensor:
    if not _native(src, dst):
        dst.copy_(src.float())
        return dst
    L = _lib.load()
    _lib.check(L.bz_cast_bf16_to_f32(src.data_ptr(), dst.data_ptr(), src.numel(), _lib.stream()), "cast_f32")
    return dst
