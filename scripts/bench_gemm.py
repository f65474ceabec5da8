"""Micro-benchmark: tcgen05 GEMM vs torch.matmul (cuBLAS) on Llama-3-8B shapes.
Writes gpurun_out/gemm_bench.json. CUDA-event timing, L2 flushed between iterations."""
import json, os, sys
from pathlib import Path
import torch
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from distributed_llm_training_gpu_manager_b200.ops import gemm as G

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
    ts.sort()
    return ts[len(ts) // 2]

def main():
    T = 8192
    shapes = [("qkv_fwd", T, 6144, 4096, False, False), ("o_fwd", T, 4096, 4096, False, False),
              ("gateup_fwd", T, 28672, 4096, False, False), ("down_fwd", T, 4096, 14336, False, False),
              ("gateup_dgrad", T, 4096, 28672, False, True), ("down_dgrad", T, 14336, 4096, False, True),
              ("gateup_wgrad", 28672, 4096, T, True, True), ("down_wgrad", 4096, 14336, T, True, True),
              ("square8k", 8192, 8192, 8192, False, False)]
    res = []
    for name, M, N, K, amn, bmn in shapes:
        a = torch.randn((K, M) if amn else (M, K), device="cuda").bfloat16()
        b = torch.randn((K, N) if bmn else (N, K), device="cuda").bfloat16()
        fl = 2.0 * M * N * K
        row = {"name": name, "M": M, "N": N, "K": K}
        for cg in (1, 2):
            try:
                ms = timeit(lambda: G.gemm(a, b, a_mn_major=amn, b_mn_major=bmn, cta_group=cg))
                row[f"cg{cg}_ms"] = ms; row[f"cg{cg}_tflops"] = fl / ms / 1e9
            except Exception as ex:
                row[f"cg{cg}_err"] = str(ex)[:200]
        at = a.t() if amn else a
        bt = b if bmn else b.t()
        ms = timeit(lambda: torch.matmul(at, bt))
        row["cublas_ms"] = ms; row["cublas_tflops"] = fl / ms / 1e9
        print(row, flush=True)
        res.append(row)
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump(res, open("gpurun_out/gemm_bench.json", "w"), indent=1)

if __name__ == "__main__":
    main()
