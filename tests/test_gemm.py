"""tcgen05 GEMM (csrc/gemm.cu) against an fp32 PyTorch oracle of the same op."""
import pytest
import torch

from distributed_llm_training_gpu_manager_b200.ops import gemm as G

pytestmark = pytest.mark.gpu


def _ref(a, b, a_mn, b_mn):
    af = (a.t() if a_mn else a).float()
    bf = (b if b_mn else b.t()).float()
    return af @ bf


def _mk(M, N, K, a_mn, b_mn, seed=0):
    g = torch.Generator(device="cuda").manual_seed(seed)
    a = torch.randn((K, M) if a_mn else (M, K), device="cuda", generator=g).to(torch.bfloat16)
    b = torch.randn((K, N) if b_mn else (N, K), device="cuda", generator=g).to(torch.bfloat16)
    return a, b


def _check(c, ref, K):
    err = (c.float() - ref).abs().max().item()
    scale = ref.abs().max().item() + 1e-6
    # bf16 output rounding (2^-8 relative) dominates; fp32 accumulation error is ~sqrt(K)*eps
    assert err / scale < 1.5e-2, f"max err {err} vs scale {scale}"
    # and the mean error must be far smaller (catches swapped tiles that happen to be in range)
    assert (c.float() - ref).abs().mean().item() / (ref.abs().mean().item() + 1e-6) < 6e-3


@pytest.mark.parametrize("cg", [1, 2])
@pytest.mark.parametrize("a_mn,b_mn", [(False, False), (False, True), (True, True), (True, False)])
@pytest.mark.parametrize("shape", [(256, 256, 64), (512, 768, 256), (1024, 512, 4096)])
def test_gemm_majors(cg, a_mn, b_mn, shape):
    M, N, K = shape
    a, b = _mk(M, N, K, a_mn, b_mn)
    c = G.gemm(a, b, a_mn_major=a_mn, b_mn_major=b_mn, cta_group=cg)
    torch.cuda.synchronize()
    _check(c, _ref(a, b, a_mn, b_mn), K)


@pytest.mark.parametrize("cg", [1, 2])
@pytest.mark.parametrize("shape", [(100, 72, 40), (300, 520, 136), (129, 257, 65 * 8), (2000, 1000, 1000)])
def test_gemm_tails(cg, shape):
    M, N, K = shape
    a, b = _mk(M, N, K, False, False, seed=1)
    c = G.gemm(a, b, cta_group=cg)
    torch.cuda.synchronize()
    _check(c, _ref(a, b, False, False), K)


@pytest.mark.parametrize("cg", [1, 2])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
def test_gemm_accumulate(cg, dtype):
    M, N, K = 384, 512, 320
    a, b = _mk(M, N, K, True, True, seed=2)
    base = torch.randn(M, N, device="cuda").to(dtype)
    out = base.clone()
    G.gemm(a, b, a_mn_major=True, b_mn_major=True, out=out, accumulate=True, cta_group=cg)
    torch.cuda.synchronize()
    _check(out, _ref(a, b, True, True) + base.float(), K)
    out2 = torch.empty(M, N, device="cuda", dtype=dtype)
    G.gemm(a, b, a_mn_major=True, b_mn_major=True, out=out2, cta_group=cg)
    _check(out2, _ref(a, b, True, True), K)


def test_gemm_llama_shapes_and_persistence():
    # more tiles than SMs: exercises the persistent scheduler, both TMEM buffers and phase wrap
    for (M, N, K) in [(4096, 4096, 4096), (2048, 14336, 4096), (2048, 4096, 14336)]:
        a, b = _mk(M, N, K, False, False, seed=3)
        c = G.gemm(a, b)
        ref = _ref(a, b, False, False)
        _check(c, ref, K)
        c1 = G.gemm(a, b, cta_group=1, max_ctas=60)
        _check(c1, ref, K)


def test_linear_triplet_matches_autograd():
    T, K, N = 512, 1024, 768
    x = torch.randn(T, K, device="cuda").bfloat16()
    w = (torch.randn(N, K, device="cuda") * 0.05).bfloat16()
    dy = torch.randn(T, N, device="cuda").bfloat16()
    xf, wf = x.float().requires_grad_(), w.float().requires_grad_()
    (xf @ wf.t()).backward(dy.float())
    _check(G.linear_fwd(x, w), (x.float() @ w.float().t()), K)
    _check(G.linear_dgrad(dy, w), xf.grad, N)
    _check(G.linear_wgrad(dy, x), wf.grad, T)
