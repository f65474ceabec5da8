"""GPT-2 element-wise ops (LayerNorm, bias + GELU-tanh): the fp32 restatements and their autograd wrappers
against torch autograd on CPU; the sm_100a kernels against the restatements on a GPU."""
import os
import subprocess
import sys
from pathlib import Path

import pytest
import torch
import torch.nn.functional as F

from distributed_llm_training_gpu_manager_b200.ops import gpt2_ops as G2

ROOT = Path(__file__).resolve().parent.parent


def test_layernorm_function_matches_torch_autograd():
    torch.manual_seed(0)
    x = torch.randn(3, 7, 48, dtype=torch.float64).float().requires_grad_()
    w = (1 + 0.1 * torch.randn(48)).requires_grad_()
    b = (0.1 * torch.randn(48)).requires_grad_()
    dy = torch.randn(3, 7, 48)
    y = G2.layer_norm(x, w, b, 1e-5)
    gx, gw, gb = torch.autograd.grad(y, (x, w, b), dy)
    y0 = F.layer_norm(x, (48,), w, b, 1e-5)
    hx, hw, hb = torch.autograd.grad(y0, (x, w, b), dy)
    assert torch.allclose(y, y0, atol=1e-5)
    assert torch.allclose(gx, hx, atol=1e-5) and torch.allclose(gw, hw, atol=1e-4) and torch.allclose(gb, hb, atol=1e-5)


def test_bias_gelu_function_matches_torch_autograd():
    torch.manual_seed(1)
    x = torch.randn(5, 9, 64).requires_grad_()
    b = (0.3 * torch.randn(64)).requires_grad_()
    dy = torch.randn(5, 9, 64)
    y = G2.bias_gelu(x, b)
    gx, gb = torch.autograd.grad(y, (x, b), dy)
    y0 = F.gelu(x + b, approximate="tanh")
    hx, hb = torch.autograd.grad(y0, (x, b), dy)
    assert torch.allclose(y, y0, atol=1e-6)
    assert torch.allclose(gx, hx, atol=1e-5) and torch.allclose(gb, hb, atol=1e-4)


def test_gpt2_block_same_with_own_ops(monkeypatch):
    from distributed_llm_training_gpu_manager_b200.models import build_torch_model, get_config
    mc = get_config("gpt2-tiny")
    torch.manual_seed(2)
    model = build_torch_model(mc)
    ids = torch.randint(0, mc.vocab_size, (2, 16))

    def run():
        model.zero_grad()
        loss = model(ids, ids)
        loss.backward()
        return loss.item(), model.blocks[0].ln_2.weight.grad.clone(), model.blocks[0].mlp_fc.bias.grad.clone(), \
            model.blocks[0].mlp_fc.weight.grad.clone()
    monkeypatch.setenv("B200ZERO_GPT2_NATIVE_OPS", "0")
    a = run()
    monkeypatch.setenv("B200ZERO_GPT2_NATIVE_OPS", "1")
    b = run()
    assert a[0] == pytest.approx(b[0], rel=1e-5)
    for u, v in zip(a[1:], b[1:]):
        assert torch.allclose(u, v, atol=1e-5, rtol=1e-4)


_GPU_CHECK = r"""
import sys, torch
sys.path.insert(0, %r)
from distributed_llm_training_gpu_manager_b200.ops import gpt2_ops as G2
torch.manual_seed(0)
dev = "cuda"
for T, D in ((64, 128), (1000, 1600), (4096, 6400), (33, 4104)):
    x = torch.randn(T, D, device=dev).bfloat16()
    w = (1 + 0.1 * torch.randn(D, device=dev)).bfloat16()
    b = (0.1 * torch.randn(D, device=dev)).bfloat16()
    dy = torch.randn(T, D, device=dev).bfloat16()
    y, mean, rstd = G2.layernorm_fwd(x, w, b, 1e-5)
    y0, mean0, rstd0 = G2.layernorm_ref(x, w, b, 1e-5)
    assert torch.allclose(y.float(), y0.float(), atol=3e-2, rtol=2e-2), ("ln fwd", T, D)
    assert torch.allclose(mean, mean0, atol=1e-4) and torch.allclose(rstd, rstd0, rtol=1e-4), ("ln stats", T, D)
    dx, dw, db = G2.layernorm_bwd(x, w, mean, rstd, dy)
    dx0, dw0, db0 = G2.layernorm_bwd_ref(x, w, mean0, rstd0, dy)
    assert torch.allclose(dx.float(), dx0.float(), atol=3e-2, rtol=2e-2), ("ln dx", T, D)
    assert torch.allclose(dw, dw0, atol=2e-2 * T ** 0.5, rtol=1e-2) and torch.allclose(db, db0, atol=2e-2 * T ** 0.5, rtol=1e-2), ("ln dw/db", T, D)
    g = G2.bias_gelu_fwd(x, b)
    assert torch.allclose(g.float(), G2.bias_gelu_ref(x, b).float(), atol=2e-2, rtol=2e-2), ("gelu fwd", T, D)
    gx, gb = G2.bias_gelu_bwd(x, b, dy)
    gx0, gb0 = G2.bias_gelu_bwd_ref(x, b, dy)
    assert torch.allclose(gx.float(), gx0.float(), atol=2e-2, rtol=2e-2), ("gelu dx", T, D)
    assert torch.allclose(gb, gb0, atol=3e-2 * T ** 0.5, rtol=2e-2), ("gelu db", T, D)
torch.cuda.synchronize()
print("GPT2_OPS_OK")
"""


@pytest.mark.gpu
def test_gpt2_kernels_match_reference_on_gpu():
    """Runs in a child process so a fault in a not-yet-validated kernel cannot poison this process's CUDA
    context for the tests collected after it."""
    res = subprocess.run([sys.executable, "-c", _GPU_CHECK % str(ROOT)], capture_output=True, text=True, timeout=150,
                         env=dict(os.environ))
    assert "GPT2_OPS_OK" in res.stdout, res.stdout[-1500:] + res.stderr[-3000:]


def test_linear_function_matches_torch_autograd():
    torch.manual_seed(4)
    x = torch.randn(3, 5, 24).requires_grad_()
    w = torch.randn(40, 24).requires_grad_()
    b = torch.randn(40).requires_grad_()
    dy = torch.randn(3, 5, 40)
    g = torch.autograd.grad(G2.linear(x, w, b), (x, w, b), dy)
    h = torch.autograd.grad(F.linear(x, w, b), (x, w, b), dy)
    for u, v in zip(g, h):
        assert torch.allclose(u, v, atol=1e-4, rtol=1e-4)
    g2 = torch.autograd.grad(G2.linear(x, w), (x, w), dy)
    h2 = torch.autograd.grad(F.linear(x, w), (x, w), dy)
    assert all(torch.allclose(u, v, atol=1e-4, rtol=1e-4) for u, v in zip(g2, h2))
