"""bf16 GEMM front-end for the tcgen05 kernel (``csrc/gemm.cu``, C ABI ``bz_gemm_bf16``).

The three training GEMMs map onto one kernel through operand "major-ness" flags, so
no transpose is ever materialised (see the header comment of ``csrc/gemm.cu``):

* ``linear_fwd(x, w)``        y[T,N]  = x[T,K]  @ w[N,K]^T
* ``linear_dgrad(dy, w)``     dx[T,K] = dy[T,N] @ w[N,K]
* ``linear_wgrad(dy, x)``     dw[N,K] = dy[T,N]^T @ x[T,K]   (optionally accumulated in place
  into a slice of the flat gradient bucket — SURVEY.md §8 row K4)

On CPU (no GPU present) the same functions run an fp32 PyTorch oracle so the gloo
test tier exercises identical call sites. The reference has no GEMM code (it never
imports torch: SURVEY.md §0); cuBLAS via a user script is what these replace (F1).
"""
from __future__ import annotations

from typing import Optional

import torch

from . import lib as _lib

OUT_BF16, OUT_BF16_ACC, OUT_F32, OUT_F32_ACC = 0, 1, 2, 3


def _native(*ts: torch.Tensor) -> bool:
    return all(t.is_cuda for t in ts) and _lib.available()


def gemm(a: torch.Tensor, b: torch.Tensor, *, a_mn_major: bool = False, b_mn_major: bool = False,
         out: Optional[torch.Tensor] = None, accumulate: bool = False, out_dtype: Optional[torch.dtype] = None,
         cta_group: int = 0, max_ctas: int = 0) -> torch.Tensor:
    """C[M,N] (+)= A·B. ``a``: [M,K] (or [K,M] if ``a_mn_major``); ``b``: [N,K] (or [K,N] if ``b_mn_major``)."""
    assert a.dim() == 2 and b.dim() == 2
    M, K = (a.shape[1], a.shape[0]) if a_mn_major else (a.shape[0], a.shape[1])
    N, Kb = (b.shape[1], b.shape[0]) if b_mn_major else (b.shape[0], b.shape[1])
    assert K == Kb, f"reduction dims differ: {K} vs {Kb}"
    if out is None:
        assert not accumulate, "accumulate=True needs an existing output"
        out_dtype = out_dtype or a.dtype
        out = torch.empty((M, N), dtype=out_dtype, device=a.device)
    else:
        assert out.shape == (M, N) and out.stride(1) == 1
        out_dtype = out.dtype
    if not _native(a, b, out):
        af = (a.t() if a_mn_major else a).float()
        bf = (b if b_mn_major else b.t()).float()
        r = af @ bf
        if accumulate:
            r = r + out.float()
        out.copy_(r.to(out.dtype))
        return out
    assert a.dtype == torch.bfloat16 and b.dtype == torch.bfloat16, "tcgen05 path is bf16-only"
    assert a.stride(1) == 1 and b.stride(1) == 1, "operands must be row-major (unit inner stride)"
    if out_dtype == torch.bfloat16:
        mode = OUT_BF16_ACC if accumulate else OUT_BF16
    elif out_dtype == torch.float32:
        mode = OUT_F32_ACC if accumulate else OUT_F32
    else:
        raise TypeError(f"unsupported output dtype {out_dtype}")
    L = _lib.load()
    rc = L.bz_gemm_bf16(a.data_ptr(), b.data_ptr(), out.data_ptr(), M, N, K, a.stride(0), b.stride(0),
                        out.stride(0), int(a_mn_major), int(b_mn_major), mode, cta_group, max_ctas, _lib.stream())
    _lib.check(rc, "gemm_bf16")
    return out


def linear_fwd(x: torch.Tensor, w: torch.Tensor, out: Optional[torch.Tensor] = None, **kw) -> torch.Tensor:
    return gemm(x, w, out=out, **kw)


def linear_dgrad(dy: torch.Tensor, w: torch.Tensor, out: Optional[torch.Tensor] = None, **kw) -> torch.Tensor:
    return gemm(dy, w, b_mn_major=True, out=out, **kw)


def linear_wgrad(dy: torch.Tensor, x: torch.Tensor, out: Optional[torch.Tensor] = None, accumulate: bool = False,
                 out_dtype: Optional[torch.dtype] = None, **kw) -> torch.Tensor:
    return gemm(dy, x, a_mn_major=True, b_mn_major=True, out=out, accumulate=accumulate, out_dtype=out_dtype, **kw)
