"""Python front-ends for the bandwidth-bound sm_100a kernels (``csrc/elementwise.cu``).

Every function has two bodies: the native one (ctypes → C ABI, used whenever the tensors
live on a GPU — a missing library is then a hard error) and an fp32 PyTorch restatement
used on CPU. The restatements double as the numerical oracles in ``tests/``.
"""
from __future__ import annotations

import functools
from typing import Optional, Tuple

import torch

from . import lib as _lib

_ws_cache = {}


def _native(*ts) -> bool:
    ts = [t for t in ts if t is not None]
    return bool(ts) and all(t.is_cuda for t in ts) and _lib.available()


def _workspace(device, nfloats: int, tag: str) -> torch.Tensor:
    key = (str(device), tag)
    buf = _ws_cache.get(key)
    if buf is None or buf.numel() < nfloats:
        buf = torch.empty(max(nfloats, 1), dtype=torch.float32, device=device)
        _ws_cache[key] = buf
    return buf


# ------------------------------------------------------------------ RMSNorm
_lib.register("bz_rmsnorm_fwd_v2", _lib.i32, [_lib.vp, _lib.vp, _lib.vp, _lib.vp, _lib.vp, _lib.vp, _lib.i32, _lib.i32,
                                              _lib.f32, _lib.vp])
_lib.register("bz_rmsnorm_bwd_v2", _lib.i32, [_lib.vp, _lib.vp, _lib.vp, _lib.vp, _lib.vp, _lib.vp, _lib.vp, _lib.i32,
                                              _lib.i32, _lib.vp, _lib.i32, _lib.i32, _lib.f32, _lib.vp])


_lib.register("bz_swiglu_fwd_v2", _lib.i32, [_lib.vp, _lib.vp, _lib.i64, _lib.i32, _lib.i32, _lib.vp])
_lib.register("bz_swiglu_bwd_v2", _lib.i32, [_lib.vp, _lib.vp, _lib.vp, _lib.i64, _lib.i32, _lib.i32, _lib.vp])


def _swiglu_version() -> int:
    """1 = csrc/elementwise.cu; 2 = the division-free kernels of csrc/elementwise_v2.cu (bit-identical to 1);
    3 (default) = the same with rcp.approx in the sigmoid. Measured on B200 at [8192, 2·14336]: forward
    303 / 148 / 119 µs, backward 209 / 211 / 193 µs (profiles/kernel_checks_r2.txt)."""
    import os
    return int(os.environ.get("B200ZERO_SWIGLU", "3"))


def _rmsnorm_version() -> int:
    """2 (default) = the row-ahead-prefetch kernels of csrc/elementwise_v2.cu, bit-identical to 1
    (csrc/elementwise.cu); measured on B200 at [8192, 4096]: backward 114 -> 93 µs, forward unchanged."""
    import os
    return int(os.environ.get("B200ZERO_RMSNORM", "2"))


def rmsnorm_ref(x, w, eps, delta=None):
    xo = x.float() + (delta.float() if delta is not None else 0)
    if delta is not None:
        xo = xo.to(x.dtype).float()
    rstd = torch.rsqrt(xo.pow(2).mean(-1, keepdim=True) + eps)
    return (xo * rstd * w.float()).to(x.dtype), rstd.squeeze(-1), (xo.to(x.dtype) if delta is not None else None)


def rmsnorm_fwd(x: torch.Tensor, w: torch.Tensor, eps: float = 1e-5, delta: Optional[torch.Tensor] = None,
                want_rstd: bool = True) -> Tuple[torch.Tensor, Optional[torch.Tensor], Optional[torch.Tensor]]:
    """Returns (y, rstd, x_out). ``x_out = x + delta`` when ``delta`` is given (fused residual join)."""
    T, D = x.shape
    if not _native(x, w, delta):
        y, rstd, xo = rmsnorm_ref(x, w, eps, delta)
        return y, (rstd if want_rstd else None), xo
    y = torch.empty_like(x)
    xo = torch.empty_like(x) if delta is not None else None
    rstd = torch.empty(T, dtype=torch.float32, device=x.device) if want_rstd else None
    L = _lib.load()
    fn = L.bz_rmsnorm_fwd_v2 if _rmsnorm_version() == 2 else L.bz_rmsnorm_fwd
    _lib.check(fn(x.data_ptr(), _lib.ptr(delta), w.data_ptr(), _lib.ptr(xo), y.data_ptr(),
                  _lib.ptr(rstd), T, D, eps, _lib.stream()), "rmsnorm_fwd")
    return y, rstd, xo


def rmsnorm_bwd_ref(x, w, dy, eps, dres=None):
    xf, wf, dyf = x.float(), w.float(), dy.float()
    rstd = torch.rsqrt(xf.pow(2).mean(-1, keepdim=True) + eps)
    xhat = xf * rstd
    g = dyf * wf
    dx = rstd * (g - xhat * (g * xhat).mean(-1, keepdim=True))
    if dres is not None:
        dx = dx + dres.float()
    return dx.to(x.dtype), (dyf * xhat).sum(0)


def rmsnorm_bwd(x: torch.Tensor, w: torch.Tensor, dy: torch.Tensor, eps: float = 1e-5,
                rstd: Optional[torch.Tensor] = None, dres: Optional[torch.Tensor] = None,
                dw_out: Optional[torch.Tensor] = None, dw_accumulate: bool = False,
                dx_out: Optional[torch.Tensor] = None) -> Tuple[torch.Tensor, Optional[torch.Tensor]]:
    """dx = dres + ∂rmsnorm/∂x·dy ; dw (+)= Σ_rows dy·x̂ (into ``dw_out``: bf16 or fp32, may be a
    slice of the flat gradient bucket)."""
    T, D = x.shape
    if not _native(x, w, dy, dres, dw_out):
        dx, dw = rmsnorm_bwd_ref(x, w, dy, eps, dres)
        if dx_out is not None:
            dx_out.copy_(dx)
            dx = dx_out
        if dw_out is not None:
            dw_out.copy_((dw + (dw_out.float() if dw_accumulate else 0)).to(dw_out.dtype))
            return dx, dw_out
        return dx, dw
    L = _lib.load()
    dx = dx_out if dx_out is not None else torch.empty_like(x)
    if dw_out is None:
        dw_out = torch.empty(D, dtype=torch.float32, device=x.device)
        dw_accumulate = False
    ws = _workspace(x.device, L.bz_rmsnorm_bwd_workspace_floats(T, D), "rms_bwd")
    fn = L.bz_rmsnorm_bwd_v2 if _rmsnorm_version() == 2 else L.bz_rmsnorm_bwd
    _lib.check(fn(x.data_ptr(), w.data_ptr(), _lib.ptr(rstd), dy.data_ptr(), _lib.ptr(dres),
                  dx.data_ptr(), dw_out.data_ptr(), int(dw_out.dtype == torch.float32),
                  int(dw_accumulate), ws.data_ptr(), T, D, eps, _lib.stream()), "rmsnorm_bwd")
    _lib.add_launches(1)  # two kernels behind one entry point
    return dx, dw_out


# ------------------------------------------------------------------ RoPE
@functools.lru_cache(maxsize=16)
def _rope_tables_cpu(S: int, hd: int, theta: float):
    inv = 1.0 / (theta ** (torch.arange(0, hd, 2, dtype=torch.float64) / hd))
    ang = torch.outer(torch.arange(S, dtype=torch.float64), inv)
    return ang.cos().float().contiguous(), ang.sin().float().contiguous()


_rope_dev_cache = {}


def rope_tables(S: int, hd: int, theta: float, device) -> Tuple[torch.Tensor, torch.Tensor]:
    key = (S, hd, float(theta), str(device))
    if key not in _rope_dev_cache:
        c, s = _rope_tables_cpu(S, hd, float(theta))
        _rope_dev_cache[key] = (c.to(device), s.to(device))
    return _rope_dev_cache[key]


def rope_ref(buf: torch.Tensor, S: int, n_rot_heads: int, hd: int, theta: float, inverse: bool = False):
    T = buf.shape[0]
    cos, sin = rope_tables(S, hd, theta, buf.device)
    pos = torch.arange(T, device=buf.device) % S
    c, s = cos[pos][:, None, :], sin[pos][:, None, :]
    if inverse:
        s = -s
    part = buf[:, : n_rot_heads * hd].float().reshape(T, n_rot_heads, hd)
    a, b = part[..., : hd // 2], part[..., hd // 2:]
    out = torch.cat([a * c - b * s, b * c + a * s], -1).reshape(T, n_rot_heads * hd)
    buf[:, : n_rot_heads * hd] = out.to(buf.dtype)
    return buf


def rope_(buf: torch.Tensor, S: int, n_rot_heads: int, hd: int, theta: float = 500000.0,
          inverse: bool = False) -> torch.Tensor:
    """In-place rotary embedding on the first ``n_rot_heads`` heads of each row of a fused
    [T, (Hq+2·Hkv)·hd] QKV buffer (pass Hq+Hkv to rotate q and k, leaving v untouched)."""
    if not _native(buf):
        return rope_ref(buf, S, n_rot_heads, hd, theta, inverse)
    cos, sin = rope_tables(S, hd, theta, buf.device)
    L = _lib.load()
    _lib.check(L.bz_rope(buf.data_ptr(), cos.data_ptr(), sin.data_ptr(), buf.shape[0], S, n_rot_heads, hd,
                         buf.stride(0), int(inverse), _lib.stream()), "rope")
    return buf


# ------------------------------------------------------------------ SwiGLU
def swiglu_fwd(gu: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    T, F2 = gu.shape
    F = F2 // 2
    if not _native(gu):
        g, u = gu[:, :F].float(), gu[:, F:].float()
        r = (torch.nn.functional.silu(g) * u).to(gu.dtype)
        if out is not None:
            out.copy_(r)
            return out
        return r
    out = out if out is not None else torch.empty((T, F), dtype=gu.dtype, device=gu.device)
    L = _lib.load()
    ver = _swiglu_version()
    if ver >= 2:
        _lib.check(L.bz_swiglu_fwd_v2(gu.data_ptr(), out.data_ptr(), T, F, int(ver == 3), _lib.stream()), "swiglu_fwd")
    else:
        _lib.check(L.bz_swiglu_fwd(gu.data_ptr(), out.data_ptr(), T, F, _lib.stream()), "swiglu_fwd")
    return out


def swiglu_bwd(gu: torch.Tensor, dact: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    T, F2 = gu.shape
    F = F2 // 2
    if not _native(gu, dact):
        g, u, d = gu[:, :F].float(), gu[:, F:].float(), dact.float()
        sig = torch.sigmoid(g)
        r = torch.cat([d * u * sig * (1 + g * (1 - sig)), d * g * sig], 1).to(gu.dtype)
        if out is not None:
            out.copy_(r)
            return out
        return r
    out = out if out is not None else torch.empty_like(gu)
    L = _lib.load()
    ver = _swiglu_version()
    if ver >= 2:
        _lib.check(L.bz_swiglu_bwd_v2(gu.data_ptr(), dact.data_ptr(), out.data_ptr(), T, F, int(ver == 3), _lib.stream()),
                   "swiglu_bwd")
    else:
        _lib.check(L.bz_swiglu_bwd(gu.data_ptr(), dact.data_ptr(), out.data_ptr(), T, F, _lib.stream()), "swiglu_bwd")
    return out


# ------------------------------------------------------------------ embedding
def embedding_fwd(ids: torch.Tensor, table: torch.Tensor) -> torch.Tensor:
    ids = ids.reshape(-1)
    V, D = table.shape
    if not _native(ids, table):
        return table[ids.clamp(0, V - 1)]
    out = torch.empty((ids.numel(), D), dtype=table.dtype, device=table.device)
    L = _lib.load()
    _lib.check(L.bz_embedding_fwd(ids.data_ptr(), int(ids.dtype == torch.int64), table.data_ptr(), out.data_ptr(),
                                  ids.numel(), D, V, _lib.stream()), "embedding_fwd")
    return out


def embedding_bwd(ids: torch.Tensor, dout: torch.Tensor, dtable: torch.Tensor) -> torch.Tensor:
    """dtable[ids[t]] += dout[t] (dtable: bf16 or fp32 [V, D], typically a bucket slice)."""
    ids = ids.reshape(-1)
    V, D = dtable.shape
    if not _native(ids, dout, dtable):
        acc = torch.zeros((V, D), dtype=torch.float32, device=dtable.device)
        acc.index_add_(0, ids.clamp(0, V - 1), dout.float())
        dtable.copy_((dtable.float() + acc).to(dtable.dtype))
        return dtable
    L = _lib.load()
    _lib.check(L.bz_embedding_bwd(ids.data_ptr(), int(ids.dtype == torch.int64), dout.data_ptr(), dtable.data_ptr(),
                                  int(dtable.dtype == torch.float32), ids.numel(), D, V, _lib.stream()),
               "embedding_bwd")
    return dtable


# ------------------------------------------------------------------ cross entropy
def cross_entropy_(logits: torch.Tensor, targets: torch.Tensor, grad_scale: float = 1.0,
                   write_grad: bool = True) -> torch.Tensor:
    """Per-token loss (fp32 [T]); when ``write_grad`` the logits buffer is overwritten in
    place with ``(softmax − onehot)·grad_scale`` (ignored targets (<0) get zero rows)."""
    T, V = logits.shape
    targets = targets.reshape(-1)
    if not _native(logits, targets):
        z = logits.float()
        lse = torch.logsumexp(z, -1)
        valid = (targets >= 0) & (targets < V)
        tgt = targets.clamp(0, V - 1)
        loss = torch.where(valid, lse - z.gather(1, tgt[:, None]).squeeze(1), torch.zeros_like(lse))
        if write_grad:
            p = torch.softmax(z, -1)
            p[torch.arange(T, device=z.device), tgt] -= 1.0
            p = p * grad_scale * valid[:, None].float()
            logits.copy_(p.to(logits.dtype))
        return loss
    loss = torch.empty(T, dtype=torch.float32, device=logits.device)
    L = _lib.load()
    _lib.check(L.bz_cross_entropy(logits.data_ptr(), targets.data_ptr(), int(targets.dtype == torch.int64),
                                  loss.data_ptr(), None, T, V, logits.stride(0), grad_scale, int(write_grad),
                                  _lib.stream()), "cross_entropy")
    return loss


def add(a: torch.Tensor, b: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    out = out if out is not None else torch.empty_like(a)
    if not _native(a, b, out) or a.numel() % 8:
        out.copy_((a.float() + b.float()).to(a.dtype))
        return out
    L = _lib.load()
    _lib.check(L.bz_add_bf16(a.data_ptr(), b.data_ptr(), out.data_ptr(), a.numel(), _lib.stream()), "add")
    return out
