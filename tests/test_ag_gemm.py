"""All-gather→GEMM kernel. The kernel has no cross-rank waits (it only reads static peer shards),
so W ranks can be emulated exactly on one GPU: W shard buffers on the same device, one launch per
emulated rank."""
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("cg", [1, 2])
@pytest.mark.parametrize("ws,M,N,K", [(1, 256, 512, 256), (4, 512, 1024, 512), (8, 1024, 6144, 1024),
                                       (8, 300, 2048, 320), (4, 512, 4 * 272, 256)])
def test_ag_gemm_matches_reference(cg, ws, M, N, K):
    from distributed_llm_training_gpu_manager_b200.ops.ag_gemm import ag_linear_fwd, ptr_array
    torch.manual_seed(0)
    x = torch.randn(M, K, device="cuda").bfloat16()
    w = (torch.randn(N, K, device="cuda") * 0.1).bfloat16()
    shards = [w[r * (N // ws):(r + 1) * (N // ws)].clone() for r in range(ws)]
    ref = x.float() @ w.float().t()
    for rank in range(ws):
        w_full = torch.full((N, K), float("nan"), device="cuda", dtype=torch.bfloat16)
        y = ag_linear_fwd(x, ptr_array([s.data_ptr() for s in shards]), w_full, rank, ws, cta_group=cg)
        torch.cuda.synchronize()
        assert torch.equal(w_full, w), "gathered weight must be bit-identical"
        err = (y.float() - ref).abs().max().item()
        assert err / (ref.abs().max().item() + 1e-6) < 1.5e-2, (rank, err)


def test_ag_gemm_llama_shape_persistent():
    from distributed_llm_training_gpu_manager_b200.ops.ag_gemm import ag_linear_fwd, ptr_array
    torch.manual_seed(1)
    ws, M, N, K = 8, 4096, 6144, 4096
    x = torch.randn(M, K, device="cuda").bfloat16()
    w = (torch.randn(N, K, device="cuda") * 0.05).bfloat16()
    shards = [w[r * (N // ws):(r + 1) * (N // ws)].clone() for r in range(ws)]
    w_full = torch.zeros(N, K, device="cuda", dtype=torch.bfloat16)
    y = ag_linear_fwd(x, ptr_array([s.data_ptr() for s in shards]), w_full, 3, ws)
    ref = x.float() @ w.float().t()
    assert torch.equal(w_full, w)
    assert (y.float() - ref).abs().max().item() / ref.abs().max().item() < 1.5e-2
