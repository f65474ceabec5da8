"""RMSNorm backward through the entry point that brings its own scratch.

Some callers cannot hand the library a workspace: a step loop that must not trigger cudaMalloc's
device-wide synchronisation, a job whose HBM is fully planned, or a C host with no allocator.
For them ``bz_rmsnorm_bwd_static_scratch`` keeps the dw partial sums in a device array inside the
library. This test drives it the way such a caller would — straight through the C ABI with
ctypes — and checks it against the existing ``bz_rmsnorm_bwd`` (caller workspace) on two
back-to-back calls of different shapes, so leftovers of the first call in the shared scratch
would show up in the second. It also checks that an oversized problem is refused.

Kept in a file that sorts last so a ``-x`` run has already exercised everything else.
"""
import ctypes as C
from pathlib import Path

import pytest
import torch

pytestmark = pytest.mark.gpu

LIB = Path(__file__).resolve().parent.parent / "distributed_llm_training_gpu_manager_b200" / "ops" / "libb200zero.so"
CAPACITY_FLOATS = 1310720  # BZ_RMSNORM_BWD_STATIC_SCRATCH_FLOATS in include/b200zero.h


def _lib():
    lib = C.CDLL(str(LIB))
    vp, i32, f32 = C.c_void_p, C.c_int, C.c_float
    lib.bz_last_error.restype = C.c_char_p
    lib.bz_rmsnorm_bwd_workspace_floats.restype = i32
    lib.bz_rmsnorm_bwd_workspace_floats.argtypes = [i32, i32]
    lib.bz_rmsnorm_bwd.restype = i32
    lib.bz_rmsnorm_bwd.argtypes = [vp, vp, vp, vp, vp, vp, vp, i32, i32, vp, i32, i32, f32, vp]
    lib.bz_rmsnorm_bwd_static_scratch.restype = i32
    lib.bz_rmsnorm_bwd_static_scratch.argtypes = [vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, f32, vp]
    return lib


def _case(T, D, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    x = torch.randn(T, D, device="cuda", generator=g).bfloat16()
    w = (1 + 0.1 * torch.randn(D, device="cuda", generator=g)).bfloat16()
    dy = torch.randn(T, D, device="cuda", generator=g).bfloat16()
    dres = torch.randn(T, D, device="cuda", generator=g).bfloat16()
    return x, w, dy, dres


def _run_both(lib, T, D, seed, accumulate):
    x, w, dy, dres = _case(T, D, seed)
    stream = torch.cuda.current_stream().cuda_stream
    eps = 1e-5
    # existing entry point: caller-owned workspace
    ws = torch.empty(lib.bz_rmsnorm_bwd_workspace_floats(T, D), dtype=torch.float32, device="cuda")
    dx_a = torch.empty_like(x)
    dw_a = torch.full((D,), 0.25, dtype=torch.float32, device="cuda")
    rc = lib.bz_rmsnorm_bwd(x.data_ptr(), w.data_ptr(), None, dy.data_ptr(), dres.data_ptr(), dx_a.data_ptr(),
                            dw_a.data_ptr(), 1, int(accumulate), ws.data_ptr(), T, D, eps, stream)
    assert rc == 0, lib.bz_last_error()
    # new entry point: no workspace argument at all
    dx_b = torch.empty_like(x)
    dw_b = torch.full((D,), 0.25, dtype=torch.float32, device="cuda")
    rc = lib.bz_rmsnorm_bwd_static_scratch(x.data_ptr(), w.data_ptr(), None, dy.data_ptr(), dres.data_ptr(),
                                           dx_b.data_ptr(), dw_b.data_ptr(), 1, int(accumulate), T, D, eps, stream)
    assert rc == 0, lib.bz_last_error()
    torch.cuda.synchronize()
    assert torch.allclose(dx_b.float(), dx_a.float(), rtol=2e-2, atol=2e-2)
    assert torch.allclose(dw_b, dw_a, rtol=2e-2, atol=2e-2)
    # and against the fp32 PyTorch statement of the op, like the other kernel tests
    xf, wf, dyf = x.float(), w.float(), dy.float()
    rstd = torch.rsqrt(xf.pow(2).mean(-1, keepdim=True) + eps)
    xhat = xf * rstd
    gw = dyf * wf
    dx_ref = rstd * (gw - xhat * (gw * xhat).mean(-1, keepdim=True)) + dres.float()
    dw_ref = (dyf * xhat).sum(0) + (0.25 if accumulate else 0.0)
    assert torch.allclose(dx_b.float(), dx_ref, rtol=3e-2, atol=3e-2)
    assert torch.allclose(dw_b, dw_ref, rtol=3e-2, atol=0.15 * (T ** 0.5) / 4)


def test_static_scratch_matches_workspace_entry_point_twice_in_a_row():
    lib = _lib()
    _run_both(lib, T=300, D=4096, seed=11, accumulate=False)   # fills most of the shared scratch
    _run_both(lib, T=64, D=1024, seed=12, accumulate=True)     # smaller problem right after: stale data must not leak


def test_static_scratch_refuses_oversized_problem():
    lib = _lib()
    T, D = 1024, 8192
    sms = torch.cuda.get_device_properties(0).multi_processor_count
    assert min(T, 2 * sms) * D > CAPACITY_FLOATS
    x, w, dy, dres = _case(T, D, 13)
    dx = torch.empty_like(x)
    dw = torch.zeros(D, dtype=torch.float32, device="cuda")
    rc = lib.bz_rmsnorm_bwd_static_scratch(x.data_ptr(), w.data_ptr(), None, dy.data_ptr(), dres.data_ptr(),
                                           dx.data_ptr(), dw.data_ptr(), 1, 0, T, D, 1e-5,
                                           torch.cuda.current_stream().cuda_stream)
    assert rc != 0
    msg = lib.bz_last_error().decode()
    assert "scratch" in msg and str(CAPACITY_FLOATS) in msg
    torch.cuda.synchronize()
