"""Micro-benchmarks (1 GPU): flash attention fwd vs SDPA; AG-GEMM (emulated 8 ranks, local
'peers') vs plain GEMM + copy. Writes gpurun_out/kernels_bench.json."""
import json, os, sys
from pathlib import Path
import torch
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from distributed_llm_training_gpu_manager_b200.ops import gemm as G
from distributed_llm_training_gpu_manager_b200.ops.flash import flash_fwd
from distributed_llm_training_gpu_manager_b200.ops.ag_gemm import ag_linear_fwd, ptr_array
from distributed_llm_training_gpu_manager_b200.ops import attention as A

def timeit(fn, iters=10, warm=3):
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device="cuda")
    for _ in range(warm): fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        flush.zero_()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record(); fn(); e.record(); torch.cuda.synchronize()
        ts.append(s.elapsed_time(e))
    ts.sort(); return ts[len(ts) // 2]

res = {}
B, S, Hq, Hk, hd = 2, 4096, 32, 8, 128
qkv = torch.randn(B * S, (Hq + 2 * Hk) * hd, device="cuda").bfloat16()
fl = 4.0 * S * S * hd * Hq * B / 2
ms = timeit(lambda: flash_fwd(qkv, B, S, Hq, Hk, hd))
res["flash_fwd"] = {"ms": ms, "tflops": fl / ms / 1e9}
def sdpa():
    q, k, v = A._split(qkv, B, S, Hq, Hk, hd)
    return torch.nn.functional.scaled_dot_product_attention(q, k, v, is_causal=True, enable_gqa=True)
ms = timeit(sdpa)
res["sdpa_fwd"] = {"ms": ms, "tflops": fl / ms / 1e9}
from distributed_llm_training_gpu_manager_b200.ops.flash import flash_bwd
out, lse = flash_fwd(qkv, B, S, Hq, Hk, hd)
dout = torch.randn_like(out)
ms = timeit(lambda: flash_bwd(qkv, out, lse, dout, B, S, Hq, Hk, hd))
res["flash_bwd"] = {"ms": ms, "tflops": 2.5 * fl / ms / 1e9}
q_, k_, v_ = (t.detach().requires_grad_() for t in A._split(qkv, B, S, Hq, Hk, hd))
o_ = torch.nn.functional.scaled_dot_product_attention(q_, k_, v_, is_causal=True, enable_gqa=True)
do_ = dout.view(B, S, Hq, hd).transpose(1, 2)
ms = timeit(lambda: torch.autograd.grad(o_, (q_, k_, v_), do_, retain_graph=True))
res["sdpa_bwd"] = {"ms": ms, "tflops": 2.5 * fl / ms / 1e9}
print(res, flush=True)
ws, M = 8, 8192
for name, N, K in [("qkv", 6144, 4096), ("gate_up", 28672, 4096), ("down", 4096, 14336), ("o", 4096, 4096)]:
    x = torch.randn(M, K, device="cuda").bfloat16()
    w = (torch.randn(N, K, device="cuda") * 0.05).bfloat16()
    shards = [w[r * (N // ws):(r + 1) * (N // ws)].clone() for r in range(ws)]
    pa = ptr_array([s.data_ptr() for s in shards])
    wf = torch.empty(N, K, device="cuda", dtype=torch.bfloat16)
    f = 2.0 * M * N * K
    ms_ag = timeit(lambda: ag_linear_fwd(x, pa, wf, 0, ws))
    ms_plain = timeit(lambda: G.gemm(x, w))
    def unfused():
        for r in range(ws): wf[r * (N // ws):(r + 1) * (N // ws)].copy_(shards[r])
        return G.gemm(x, wf)
    ms_unf = timeit(unfused)
    res["ag_" + name] = {"ag_gemm_ms": ms_ag, "plain_gemm_ms": ms_plain, "copy_then_gemm_ms": ms_unf,
                         "ag_tflops": f / ms_ag / 1e9, "plain_tflops": f / ms_plain / 1e9}
    print(name, res["ag_" + name], flush=True)
os.makedirs("gpurun_out", exist_ok=True)
json.dump(res, open("gpurun_out/kernels_bench.json", "w"), indent=1)
