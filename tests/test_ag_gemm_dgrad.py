"""All-gather→GEMM, dgrad flavour (``dx = dy · gather(W)``; the gathered dimension is the reduction dimension).
Like the forward kernel it has no cross-rank waits, so ranks are emulated on one GPU. Written after this round's
GPU time was spent; it first ran (and passed) on hardware in round 2. The check runs in a child process."""
import os
import subprocess
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent

_CHECK = r"""
import sys, torch
sys.path.insert(0, %r)
from distributed_llm_training_gpu_manager_b200.ops.ag_gemm import ag_linear_dgrad, ptr_array
from distributed_llm_training_gpu_manager_b200.ops import gemm as G
torch.manual_seed(0)
for cg in (1, 2):
    for ws, M, N, K in ((1, 256, 512, 256), (4, 512, 1024, 512), (8, 1024, 6144, 1024), (8, 300, 2048, 320),
                        (4, 512, 4 * 272, 256), (8, 2048, 14336, 4096)):
        dy = torch.randn(M, N, device="cuda").bfloat16()
        w = (torch.randn(N, K, device="cuda") * 0.1).bfloat16()
        shards = [w[r * (N // ws):(r + 1) * (N // ws)].clone() for r in range(ws)]
        ref = dy.float() @ w.float()
        for rank in sorted({0, ws - 1, ws // 2}):
            w_full = torch.full((N, K), float("nan"), device="cuda", dtype=torch.bfloat16)
            dx = ag_linear_dgrad(dy, ptr_array([s.data_ptr() for s in shards]), w_full, rank, ws, cta_group=cg)
            torch.cuda.synchronize()
            assert torch.equal(w_full, w), ("gathered weight must be bit-identical", cg, ws, M, N, K, rank)
            err = (dx.float() - ref).abs().max().item() / (ref.abs().max().item() + 1e-6)
            assert err < 1.5e-2, ("dx", cg, ws, M, N, K, rank, err)
# timing against the plain NN GEMM on a Llama-3-8B shape (down-projection dgrad)
ws, M, N, K = 8, 8192, 4096, 14336
dy = torch.randn(M, N, device="cuda").bfloat16()
w = (torch.randn(N, K, device="cuda") * 0.05).bfloat16()
shards = [w[r * (N // ws):(r + 1) * (N // ws)].clone() for r in range(ws)]
w_full = torch.zeros(N, K, device="cuda", dtype=torch.bfloat16)
ptrs = ptr_array([s.data_ptr() for s in shards])
def t(fn):
    for _ in range(3): fn()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); s.record()
    for _ in range(10): fn()
    e.record(); torch.cuda.synchronize()
    return s.elapsed_time(e) / 10
a = t(lambda: ag_linear_dgrad(dy, ptrs, w_full, 3, ws))
b = t(lambda: G.gemm(dy, w, b_mn_major=True))
print("AG_DGRAD_OK", {"ag_dgrad_ms": round(a, 4), "plain_ms": round(b, 4), "tflops": round(2.0 * M * N * K / a / 1e9, 1)})
"""


@pytest.mark.gpu
def test_ag_gemm_dgrad_emulated_ranks():
    res = subprocess.run([sys.executable, "-c", _CHECK % str(ROOT)], capture_output=True, text=True, timeout=150,
                         env=dict(os.environ))
    try:
        (ROOT / "gpurun_out").mkdir(exist_ok=True)
        (ROOT / "gpurun_out" / "ag_dgrad_check.log").write_text(res.stdout[-4000:] + "\n--- stderr ---\n" + res.stderr[-4000:])
    except OSError:
        pass
    assert "AG_DGRAD_OK" in res.stdout, res.stdout[-1500:] + res.stderr[-3000:]
