"""Causal GQA attention on a fused QKV buffer.

``qkv``: [T, (Hq + 2·Hkv)·hd] bf16 with T = B·S tokens (RoPE already applied to q and k).
Returns ``out`` [T, Hq·hd] and a context for the backward pass, which produces ``dqkv`` in the
same fused layout (so the QKV dgrad/wgrad GEMMs consume it directly).

Backends:
* ``native`` — the tcgen05 flash-attention kernels in ``csrc/attention.cu`` (K6).
* ``sdpa``   — PyTorch SDPA; the fp32-accumulating oracle for tests and the CPU path.
The reference has no attention code (it ships no model).
"""
from __future__ import annotations

import logging
import os
from typing import Any, Optional, Tuple

import torch
import torch.nn.functional as F

from . import lib as _lib

_BACKEND = os.environ.get("B200ZERO_ATTN", "auto")
_log = logging.getLogger(__name__)
# how many forward calls each backend served in this process (bench.py reports it, so a library fallback on a GPU
# box is visible in the result line instead of hiding behind "auto")
backend_calls = {"native": 0, "sdpa": 0}
_warned = set()


def native_available() -> bool:
    if not _lib.available():
        return False
    L = _lib.load()
    return hasattr(L, "bz_flash_attn_fwd") and hasattr(L, "bz_flash_attn_bwd")


def _split(qkv: torch.Tensor, B: int, S: int, Hq: int, Hk: int, hd: int):
    q = qkv[:, : Hq * hd].view(B, S, Hq, hd).transpose(1, 2)
    k = qkv[:, Hq * hd: (Hq + Hk) * hd].view(B, S, Hk, hd).transpose(1, 2)
    v = qkv[:, (Hq + Hk) * hd:].view(B, S, Hk, hd).transpose(1, 2)
    return q, k, v


def attention_ref(qkv: torch.Tensor, B: int, S: int, Hq: int, Hk: int, hd: int) -> torch.Tensor:
    """fp32 oracle."""
    q, k, v = _split(qkv.float(), B, S, Hq, Hk, hd)
    rep = Hq // Hk
    k, v = k.repeat_interleave(rep, 1), v.repeat_interleave(rep, 1)
    s = (q @ k.transpose(-1, -2)) / (hd ** 0.5)
    mask = torch.ones(S, S, dtype=torch.bool, device=qkv.device).tril()
    s = s.masked_fill(~mask, float("-inf"))
    o = torch.softmax(s, -1) @ v
    return o.transpose(1, 2).reshape(B * S, Hq * hd)


def attention_fwd(qkv: torch.Tensor, B: int, S: int, Hq: int, Hk: int, hd: int,
                  backend: Optional[str] = None) -> Tuple[torch.Tensor, Any]:
    backend = backend or _BACKEND
    if backend == "auto":
        backend = "native" if (qkv.is_cuda and native_available() and hd == 128 and S % 128 == 0) else "sdpa"
        if backend == "sdpa" and qkv.is_cuda and (S, hd) not in _warned:
            _warned.add((S, hd))
            why = ("kernel library has no flash-attention entry points" if not native_available() else
                   f"shape outside the tcgen05 kernel's range (head_dim={hd}, seq={S}; needs head_dim 128, seq % 128 == 0)")
            _log.warning("attention: using torch SDPA on a GPU instead of the native kernel: %s", why)
    backend_calls[backend] = backend_calls.get(backend, 0) + 1
    if backend == "native":
        from .flash import flash_fwd
        out, lse = flash_fwd(qkv, B, S, Hq, Hk, hd)
        return out, ("native", qkv, out, lse, (B, S, Hq, Hk, hd))
    q, k, v = _split(qkv, B, S, Hq, Hk, hd)
    q, k, v = (t.detach().requires_grad_() for t in (q, k, v))
    with torch.enable_grad():
        o = F.scaled_dot_product_attention(q, k, v, is_causal=True, enable_gqa=(Hq != Hk))
    out = o.transpose(1, 2).reshape(B * S, Hq * hd)
    return out.detach(), ("sdpa", (q, k, v, o), (B, S, Hq, Hk, hd))


def attention_bwd(ctx: Any, dout: torch.Tensor, dqkv: Optional[torch.Tensor] = None) -> torch.Tensor:
    if ctx[0] == "native":
        from .flash import flash_bwd
        _, qkv, out, lse, dims = ctx
        return flash_bwd(qkv, out, lse, dout, *dims, dqkv=dqkv)
    _, (q, k, v, o), (B, S, Hq, Hk, hd) = ctx
    do = dout.view(B, S, Hq, hd).transpose(1, 2)
    dq, dk, dv = torch.autograd.grad(o, (q, k, v), do)
    T = B * S
    if dqkv is None:
        dqkv = torch.empty((T, (Hq + 2 * Hk) * hd), dtype=dout.dtype, device=dout.device)
    dqkv[:, : Hq * hd].view(B, S, Hq, hd).copy_(dq.transpose(1, 2))
    dqkv[:, Hq * hd: (Hq + Hk) * hd].view(B, S, Hk, hd).copy_(dk.transpose(1, 2))
    dqkv[:, (Hq + Hk) * hd:].view(B, S, Hk, hd).copy_(dv.transpose(1, 2))
    return dqkv
