"""Bandwidth-bound kernels vs their fp32 PyTorch oracles. The same tests run on CPU (oracle
path, checks the wrappers' plumbing) and on B200 under ``-m gpu`` (native kernels)."""
import pytest
import torch

from distributed_llm_training_gpu_manager_b200.ops import elementwise as E
from distributed_llm_training_gpu_manager_b200.ops import optim as O

DEVS = ["cpu", pytest.param("cuda", marks=pytest.mark.gpu)]


def close(a, b, rtol=2e-2, atol=2e-2):
    assert torch.allclose(a.float().cpu(), b.float().cpu(), rtol=rtol, atol=atol), \
        f"max diff {(a.float().cpu() - b.float().cpu()).abs().max().item()}"


@pytest.mark.parametrize("dev", DEVS)
@pytest.mark.parametrize("T,D", [(37, 128), (64, 4096), (9, 8192), (300, 1600)])
def test_rmsnorm_fwd_bwd(dev, T, D):
    torch.manual_seed(0)
    x = torch.randn(T, D, device=dev).bfloat16()
    dl = torch.randn(T, D, device=dev).bfloat16()
    w = (1 + 0.1 * torch.randn(D, device=dev)).bfloat16()
    dy = torch.randn(T, D, device=dev).bfloat16()
    y, rstd, xo = E.rmsnorm_fwd(x, w, 1e-5, delta=dl)
    yr, rr, xor_ = E.rmsnorm_ref(x, w, 1e-5, dl)
    close(y, yr); close(rstd, rr, 1e-3, 1e-3); close(xo, xor_)
    y2, _, none = E.rmsnorm_fwd(x, w, 1e-5)
    assert none is None
    close(y2, E.rmsnorm_ref(x, w, 1e-5)[0])
    # backward vs autograd on the fp32 restatement
    xf = x.float().requires_grad_(); wf = w.float().requires_grad_()
    out = xf * torch.rsqrt(xf.pow(2).mean(-1, keepdim=True) + 1e-5) * wf
    out.backward(dy.float())
    dres = torch.randn(T, D, device=dev).bfloat16()
    dx, dw = E.rmsnorm_bwd(x, w, dy, 1e-5, rstd=None, dres=dres)
    close(dx, xf.grad + dres.float(), 3e-2, 3e-2)
    close(dw, wf.grad, 3e-2, 0.15 * (T ** 0.5) / 4)
    acc = torch.ones(D, device=dev, dtype=torch.float32)
    E.rmsnorm_bwd(x, w, dy, 1e-5, dw_out=acc, dw_accumulate=True)
    close(acc, wf.grad + 1, 3e-2, 0.15 * (T ** 0.5) / 4)


@pytest.mark.parametrize("dev", DEVS)
def test_rope_roundtrip_and_oracle(dev):
    torch.manual_seed(1)
    T, S, Hq, Hk, hd = 96, 48, 4, 2, 128
    qkv = torch.randn(T, (Hq + 2 * Hk) * hd, device=dev).bfloat16()
    orig = qkv.clone()
    ref = E.rope_ref(orig.clone(), S, Hq + Hk, hd, 500000.0)
    E.rope_(qkv, S, Hq + Hk, hd, 500000.0)
    close(qkv, ref, 1e-2, 1e-2)
    assert torch.equal(qkv[:, (Hq + Hk) * hd:], orig[:, (Hq + Hk) * hd:])  # V untouched
    E.rope_(qkv, S, Hq + Hk, hd, 500000.0, inverse=True)
    close(qkv, orig, 3e-2, 3e-2)


@pytest.mark.parametrize("dev", DEVS)
def test_swiglu(dev):
    torch.manual_seed(2)
    T, F = 50, 256
    gu = torch.randn(T, 2 * F, device=dev).bfloat16()
    d = torch.randn(T, F, device=dev).bfloat16()
    guf = gu.float().requires_grad_()
    ref = torch.nn.functional.silu(guf[:, :F]) * guf[:, F:]
    ref.backward(d.float())
    close(E.swiglu_fwd(gu), ref)
    close(E.swiglu_bwd(gu, d), guf.grad, 3e-2, 3e-2)


@pytest.mark.parametrize("dev", DEVS)
@pytest.mark.parametrize("dt", [torch.bfloat16, torch.float32])
def test_embedding(dev, dt):
    torch.manual_seed(3)
    V, D, T = 1000, 128, 777
    table = torch.randn(V, D, device=dev).bfloat16()
    ids = torch.randint(0, V, (T,), device=dev)
    close(E.embedding_fwd(ids, table), table[ids], 0, 0)
    dout = torch.randn(T, D, device=dev).bfloat16()
    dtab = torch.zeros(V, D, device=dev, dtype=dt)
    E.embedding_bwd(ids, dout, dtab)
    ref = torch.zeros(V, D, device=dev).index_add_(0, ids, dout.float())
    close(dtab, ref, 3e-2, 6e-2)


@pytest.mark.parametrize("dev", DEVS)
def test_cross_entropy(dev):
    torch.manual_seed(4)
    T, V = 65, 2048
    z = (3 * torch.randn(T, V, device=dev)).bfloat16()
    y = torch.randint(0, V, (T,), device=dev)
    y[5] = -100
    zf = z.float().requires_grad_()
    ref = torch.nn.functional.cross_entropy(zf, y, ignore_index=-100, reduction="none")
    (ref.sum() * 0.5).backward()
    work = z.clone()
    loss = E.cross_entropy_(work, y, grad_scale=0.5)
    close(loss, ref, 1e-2, 2e-2)
    close(work, zf.grad, 5e-2, 2e-3)


@pytest.mark.parametrize("dev", DEVS)
@pytest.mark.parametrize("gdt", [torch.bfloat16, torch.float32])
def test_adamw_matches_torch(dev, gdt):
    torch.manual_seed(5)
    n = 4099  # exercises the scalar tail
    p0 = torch.randn(n, device=dev)
    ref_p = torch.nn.Parameter(p0.clone())
    opt = torch.optim.AdamW([ref_p], lr=1e-2, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.01)
    master, m, v = p0.clone(), torch.zeros(n, device=dev), torch.zeros(n, device=dev)
    pout = torch.empty(n, device=dev, dtype=torch.bfloat16)
    for step in range(1, 4):
        g = torch.randn(n, device=dev).to(gdt)
        ref_p.grad = g.float().clone()
        opt.step()
        O.adamw_step(master, m, v, g, lr=1e-2, weight_decay=0.01, step=step, param_out=pout)
    close(master, ref_p.data, 1e-4, 1e-5)
    close(pout, ref_p.data, 1e-2, 1e-2)
    # skip flag: nothing may change
    before = master.clone()
    flag = torch.ones(1, dtype=torch.int32, device=dev)
    O.adamw_step(master, m, v, torch.randn(n, device=dev).to(gdt), lr=1e-2, step=4, skip_flag=flag)
    assert torch.equal(before, master)


@pytest.mark.parametrize("dev", DEVS)
def test_sumsq_clip_and_nonfinite(dev):
    torch.manual_seed(6)
    g = torch.randn(100003, device=dev).bfloat16()
    out = torch.zeros(1, device=dev)
    flag = torch.zeros(1, dtype=torch.int32, device=dev)
    O.sumsq(g, out, nonfinite_flag=flag)
    close(out, g.float().pow(2).sum()[None], 1e-3, 1e-1)
    assert int(flag.item()) == 0
    O.sumsq(g.float(), out, accumulate=True)
    close(out, 2 * g.float().pow(2).sum()[None], 1e-3, 2e-1)
    coef, nrm = torch.zeros(1, device=dev), torch.zeros(1, device=dev)
    O.clip_coef(out, 1.0, 0.5, coef, nrm, flag)
    close(coef, torch.tensor([0.5 / float(out.sqrt())]), 1e-3, 1e-6)
    g[77] = float("nan")
    O.sumsq(g, out, nonfinite_flag=flag)
    assert int(flag.item()) == 1


@pytest.mark.parametrize("dev", DEVS)
def test_casts(dev):
    x = torch.randn(1001, device=dev)
    b = torch.empty(1001, device=dev, dtype=torch.bfloat16)
    O.cast_to_bf16(x, b)
    assert torch.equal(b, x.bfloat16())
    f = torch.empty(1001, device=dev)
    O.cast_to_f32(b, f)
    assert torch.equal(f, b.float())


_RMS_V2_CHECK = r"""
import os, sys, torch
sys.path.insert(0, %r)
from distributed_llm_training_gpu_manager_b200.ops import elementwise as E
torch.manual_seed(0)
def run(ver, T, D, with_delta, with_dres):
    os.environ["B200ZERO_RMSNORM"] = str(ver)
    g = torch.Generator(device="cuda").manual_seed(T * 7 + D)
    x = torch.randn(T, D, device="cuda", generator=g).bfloat16()
    w = (1 + 0.1 * torch.randn(D, device="cuda", generator=g)).bfloat16()
    dl = torch.randn(T, D, device="cuda", generator=g).bfloat16() if with_delta else None
    dy = torch.randn(T, D, device="cuda", generator=g).bfloat16()
    dr = torch.randn(T, D, device="cuda", generator=g).bfloat16() if with_dres else None
    y, rstd, xo = E.rmsnorm_fwd(x, w, 1e-5, delta=dl)
    xin = xo if xo is not None else x
    dw = torch.zeros(D, device="cuda", dtype=torch.float32)
    dx, dw = E.rmsnorm_bwd(xin, w, dy, 1e-5, rstd=rstd, dres=dr, dw_out=dw)
    dx2, _ = E.rmsnorm_bwd(xin, w, dy, 1e-5, rstd=None, dres=dr)          # statistics recomputed in the kernel
    torch.cuda.synchronize()
    return y, rstd, xo, dx, dw, dx2
for T, D in ((64, 128), (300, 1600), (1000, 4096), (513, 8192), (8192, 4096)):
    for with_delta, with_dres in ((False, False), (True, True)):
        a = run(1, T, D, with_delta, with_dres)
        b = run(2, T, D, with_delta, with_dres)
        for name, u, v in zip(("y", "rstd", "xo", "dx", "dw", "dx_norstd"), a, b):
            if u is None:
                assert v is None
                continue
            assert torch.equal(u, v), (name, T, D, with_delta, (u.float() - v.float()).abs().max().item())
# timing on the Llama-3-8B shape
T, D = 8192, 4096
x = torch.randn(T, D, device="cuda").bfloat16(); w = torch.ones(D, device="cuda").bfloat16()
dy = torch.randn(T, D, device="cuda").bfloat16(); dr = torch.randn(T, D, device="cuda").bfloat16()
dw = torch.zeros(D, device="cuda", dtype=torch.float32)
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
res = {}
for ver in (1, 2):
    os.environ["B200ZERO_RMSNORM"] = str(ver)
    _, rstd, _ = E.rmsnorm_fwd(x, w, 1e-5)
    for name, fn in (("fwd", lambda: E.rmsnorm_fwd(x, w, 1e-5)),
                     ("bwd", lambda: E.rmsnorm_bwd(x, w, dy, 1e-5, rstd=rstd, dres=dr, dw_out=dw))):
        ts = []
        for _ in range(8):
            flush.zero_()
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record(); fn(); e.record(); torch.cuda.synchronize()
            ts.append(s.elapsed_time(e))
        res[(name, ver)] = round(sorted(ts)[len(ts) // 2] * 1e3, 1)
# SwiGLU v2: division-free indexing, correctly rounded reciprocal -> bit-identical to v1
for T, F_ in ((64, 512), (300, 6400), (1000, 14336), (8192, 14336)):
    gu = torch.randn(T, 2 * F_, device="cuda").bfloat16()
    da = torch.randn(T, F_, device="cuda").bfloat16()
    outs = {}
    for ver in (1, 2):
        os.environ["B200ZERO_SWIGLU"] = str(ver)
        outs[ver] = (E.swiglu_fwd(gu), E.swiglu_bwd(gu, da))
    torch.cuda.synchronize()
    assert torch.equal(outs[1][0], outs[2][0]), ("swiglu fwd", T, F_)
    assert torch.equal(outs[1][1], outs[2][1]), ("swiglu bwd", T, F_)
    os.environ["B200ZERO_SWIGLU"] = "3"                      # rcp.approx: within a bf16 ulp of v1
    f3, b3 = E.swiglu_fwd(gu), E.swiglu_bwd(gu, da)
    assert torch.allclose(f3.float(), outs[1][0].float(), rtol=8e-3, atol=1e-3), ("swiglu fwd fast", T, F_)
    assert torch.allclose(b3.float(), outs[1][1].float(), rtol=8e-3, atol=1e-3), ("swiglu bwd fast", T, F_)
T, F_ = 8192, 14336
gu = torch.randn(T, 2 * F_, device="cuda").bfloat16(); da = torch.randn(T, F_, device="cuda").bfloat16()
for ver in (1, 2, 3):
    os.environ["B200ZERO_SWIGLU"] = str(ver)
    for name, fn in (("swiglu_fwd", lambda: E.swiglu_fwd(gu)), ("swiglu_bwd", lambda: E.swiglu_bwd(gu, da))):
        ts = []
        for _ in range(8):
            flush.zero_()
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record(); fn(); e.record(); torch.cuda.synchronize()
            ts.append(s.elapsed_time(e))
        res[(name, ver)] = round(sorted(ts)[len(ts) // 2] * 1e3, 1)
print("RMSNORM_V2_OK us:", res)
"""


@pytest.mark.gpu
def test_rmsnorm_v2_is_bit_identical_to_v1():
    """RMSNorm v2: same arithmetic in the same order, only the loads are issued a row ahead; SwiGLU v2: no integer
    division, correctly rounded reciprocal. Outputs must be bit-identical to v1.
    Runs in a child process so that a device fault cannot take the rest of the session down."""
    import os
    import subprocess
    import sys
    from pathlib import Path
    root = str(Path(__file__).resolve().parent.parent)
    res = subprocess.run([sys.executable, "-c", _RMS_V2_CHECK % root], capture_output=True, text=True, timeout=150,
                         env=dict(os.environ))
    try:
        (Path(root) / "gpurun_out").mkdir(exist_ok=True)
        (Path(root) / "gpurun_out" / "rmsnorm_v2_check.log").write_text(res.stdout[-4000:] + "\n--- stderr ---\n" + res.stderr[-4000:])
    except OSError:
        pass
    assert "RMSNORM_V2_OK" in res.stdout, res.stdout[-1500:] + res.stderr[-3000:]
