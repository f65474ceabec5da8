"""Roofline report for the fused compute+collective kernels on REAL peers (run under torchrun, one rank
per GPU). Each kernel is timed in isolation with CUDA events after warm-up, max over ranks; the target
time is the slower of (FLOPs / measured GEMM peak) and (bytes that must cross NVLink / 770 GB/s per
direction per GPU); achieved fraction = target / measured. Writes gpurun_out/fused_roofline_n{N}.json.

    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 scripts/bench_fused_multigpu.py
"""
import json
import os
import sys
from pathlib import Path

import torch
import torch.distributed as dist

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from distributed_llm_training_gpu_manager_b200.ops import gemm as G, lib as bz  # noqa: E402
from distributed_llm_training_gpu_manager_b200.ops.ag_gemm import ag_linear_fwd  # noqa: E402
from distributed_llm_training_gpu_manager_b200.parallel.fused import segment_table  # noqa: E402
from distributed_llm_training_gpu_manager_b200.parallel.symm import SymmetricBuffer, nvls_buffer_or_none  # noqa: E402

NVLINK_GBS = 770.0


def main():
    rank, ws, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", rank=rank, world_size=ws, device_id=dev)
    peaks = json.loads((ROOT / "MEASURED_PEAKS.json").read_text()) if (ROOT / "MEASURED_PEAKS.json").exists() else {}
    gemm_peak = peaks.get("bf16_tflops", 1590.0) * 1e12
    hbm = peaks.get("hbm_gbs", 6650.0) * 1e9
    L = bz.load(required=True)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

    def timeit(fn, iters=8, warm=3):
        for _ in range(warm):
            fn()
        ts = []
        for _ in range(iters):
            flush.zero_()
            torch.cuda.synchronize()
            dist.barrier()
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record(); fn(); e.record()
            torch.cuda.synchronize()
            ts.append(s.elapsed_time(e))
        ts.sort()
        t = torch.tensor([ts[len(ts) // 2]], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)       # max over ranks
        return float(t)

    res = {"world_size": ws, "nvlink_gbs_ref": NVLINK_GBS, "gemm_peak_tflops_ref": gemm_peak / 1e12}
    T, D, F_ = 8192, 4096, 14336
    # ---------------- K2: all-gather -> GEMM, the four forward GEMMs of a Llama-3-8B block ----------------
    x_d = torch.randn(T, D, device=dev).bfloat16()
    x_f = torch.randn(T, F_, device=dev).bfloat16()
    for name, N, K, x in (("qkv", 6144, D, x_d), ("o", D, D, x_d), ("gate_up", 2 * F_, D, x_d), ("down", D, F_, x_f)):
        shard_buf = SymmetricBuffer((N // ws) * K * 2, device=dev)
        shard = shard_buf.tensor(torch.bfloat16, (N // ws) * K).view(N // ws, K)
        shard.copy_((torch.randn(N // ws, K, device=dev) * 0.02).bfloat16())
        torch.cuda.synchronize(); dist.barrier()
        w_full = torch.empty(N, K, device=dev, dtype=torch.bfloat16)
        ptrs = shard_buf.ptr_array()
        ms_ag = timeit(lambda: ag_linear_fwd(x, ptrs, w_full, rank, ws))
        ms_plain = timeit(lambda: G.gemm(x, w_full))
        gathered = [torch.empty_like(shard) for _ in range(ws)]

        def nccl_then_gemm():
            dist.all_gather_into_tensor(w_full.view(-1), shard.reshape(-1))
            G.gemm(x, w_full)
        ms_nccl = timeit(nccl_then_gemm)
        flops = 2.0 * T * N * K
        remote = (ws - 1) / ws * N * K * 2
        target = max(flops / gemm_peak, remote / (NVLINK_GBS * 1e9)) * 1e3
        res[f"K2_ag_gemm_{name}"] = {"ms": ms_ag, "plain_gemm_ms": ms_plain, "nccl_allgather_then_gemm_ms": ms_nccl,
                                     "tflops": flops / ms_ag / 1e9, "nvlink_bytes": remote, "target_ms": target,
                                     "fraction_of_roofline": target / ms_ag}
        del gathered
        torch.cuda.synchronize(); dist.barrier()
        shard_buf.close()
    # ---------------- K1 / K3 / pull all-gather on one block's worth of parameters ----------------
    n = 218_112_000
    n = n // (16 * ws) * (16 * ws)
    gbuf = SymmetricBuffer(n * 2, device=dev)
    g = gbuf.tensor(torch.bfloat16, n)
    g.copy_(torch.randn(n, device=dev).bfloat16())
    sig = SymmetricBuffer(L.bz_comm_signal_pad_bytes(), device=dev)
    segs = segment_table([(0, n, 0, True)], dev)
    wsb = torch.zeros(L.bz_fused_comm_workspace_bytes() // 4, device=dev)
    shard_g = torch.zeros(n // ws, device=dev, dtype=torch.bfloat16)
    sp, sr = torch.zeros(1, device=dev), torch.zeros(1, device=dev)
    flag = torch.zeros(1, dtype=torch.int32, device=dev)
    epoch = [0]
    torch.cuda.synchronize(); dist.barrier()

    def k1():
        epoch[0] += 1
        bz.check(L.bz_fused_reduce_scatter(gbuf.ptr_array(), sig.ptr_array(), segs.data_ptr(), 1, shard_g.data_ptr(), 0,
                                           1.0 / ws, wsb.data_ptr(), sp.data_ptr(), sr.data_ptr(), flag.data_ptr(),
                                           epoch[0], 0, rank, ws, 1, int(os.environ.get("B200ZERO_RS_CTAS", "128")),
                                           bz.stream()), "rs")
    ms = timeit(k1)
    remote = (ws - 1) / ws * n * 2
    target = max(remote / (NVLINK_GBS * 1e9), (n * 2 + 2 * (n // ws) * 2) / hbm) * 1e3
    out_nccl = torch.empty(n // ws, device=dev, dtype=torch.bfloat16)

    def nccl_rs():
        dist.reduce_scatter_tensor(out_nccl, g)
        shard_g.add_(out_nccl, alpha=1.0 / ws)
        (shard_g.float() ** 2).sum()
    ms_nccl = timeit(nccl_rs)
    res["K1_fused_reduce_scatter"] = {"ms": ms, "nvlink_bytes_in": remote, "achieved_gbs": remote / ms / 1e6,
                                      "target_ms": target, "fraction_of_roofline": target / ms,
                                      "nccl_rs_plus_scale_norm_ms": ms_nccl}
    # K1 with the in-switch reduction (NVLS multimem.ld_reduce), when the fabric has multicast
    nv = nvls_buffer_or_none(n * 2, dev) if os.environ.get("B200ZERO_BENCH_NVLS", "0") == "1" else None
    if nv is not None:
        nv.tensor(torch.bfloat16, n).copy_(g)
        torch.cuda.synchronize(); dist.barrier()
        for ctas in (32, 64, 128):
            def k1mc():
                epoch[0] += 1
                bz.check(L.bz_fused_reduce_scatter_mc(nv.ptr_array(), sig.ptr_array(), nv.multicast_ptr, segs.data_ptr(),
                                                      1, shard_g.data_ptr(), 0, 1.0 / ws, wsb.data_ptr(), sp.data_ptr(),
                                                      sr.data_ptr(), flag.data_ptr(), epoch[0], 0, rank, ws, 1, ctas,
                                                      bz.stream()), "rs_mc")
            ms = timeit(k1mc)
            res[f"K1_nvls_ctas{ctas}"] = {"ms": ms, "effective_gbs": remote / ms / 1e6,
                                          "vs_p2p": res["K1_fused_reduce_scatter"]["ms"] / ms}
    # K3: AdamW + push all-gather
    pbuf = SymmetricBuffer(n * 2, device=dev)
    master = torch.randn(n // ws, device=dev); m_ = torch.zeros_like(master); v_ = torch.zeros_like(master)
    coef = torch.ones(1, device=dev)
    ep3 = [0]
    torch.cuda.synchronize(); dist.barrier()

    def k3():
        ep3[0] += 1
        bz.check(L.bz_fused_adam_push(pbuf.ptr_array(), sig.ptr_array(), segs.data_ptr(), 1, master.data_ptr(),
                                      m_.data_ptr(), v_.data_ptr(), shard_g.data_ptr(), 0, None, 1e-4, 0.9, 0.999, 1e-8,
                                      0.01, 1, coef.data_ptr(), flag.data_ptr(), wsb.data_ptr(), ep3[0], 1, rank, ws, 1,
                                      1, 0, 148, bz.stream()), "adam_push")
    ms = timeit(k3)
    out_bytes = (ws - 1) * (n // ws) * 2
    hbm_bytes = (n // ws) * (24 + 2 + 2)
    target = max(out_bytes / (NVLINK_GBS * 1e9), hbm_bytes / hbm) * 1e3
    res["K3_fused_adam_push"] = {"ms": ms, "nvlink_bytes_out": out_bytes, "achieved_gbs": out_bytes / ms / 1e6,
                                 "target_ms": target, "fraction_of_roofline": target / ms}
    if nv is not None:
        nv.close()
    # pull all-gather (ZeRO-3 bulk gather used in backward)
    sbuf = SymmetricBuffer((n // ws) * 2, device=dev)
    sbuf.tensor(torch.bfloat16, n // ws).copy_(torch.randn(n // ws, device=dev).bfloat16())
    dst = torch.empty(n, device=dev, dtype=torch.bfloat16)
    torch.cuda.synchronize(); dist.barrier()
    ms = timeit(lambda: bz.check(L.bz_pull_allgather(sbuf.ptr_array(), segs.data_ptr(), 1, dst.data_ptr(), rank, ws,
                                                    128, bz.stream()), "pull"))
    ms_v2 = timeit(lambda: bz.check(L.bz_pull_allgather_v2(sbuf.ptr_array(), segs.data_ptr(), 1, dst.data_ptr(), rank, ws,
                                                           128, bz.stream()), "pull_v2")) if hasattr(L, "bz_pull_allgather_v2") else None
    ms_nccl = timeit(lambda: dist.all_gather_into_tensor(dst, sbuf.tensor(torch.bfloat16, n // ws)))
    target = remote / (NVLINK_GBS * 1e9) * 1e3
    res["pull_allgather"] = {"ms": ms, "achieved_gbs": remote / ms / 1e6, "target_ms": target,
                             "fraction_of_roofline": target / ms, "nccl_allgather_ms": ms_nccl,
                             "v2_ms": ms_v2, "v2_fraction_of_roofline": (target / ms_v2) if ms_v2 else None}
    if rank == 0:
        os.makedirs(ROOT / "gpurun_out", exist_ok=True)
        json.dump(res, open(ROOT / "gpurun_out" / f"fused_roofline_n{ws}.json", "w"), indent=1)
        for k, v in res.items():
            if isinstance(v, dict):
                print(k, {a: (round(b, 3) if isinstance(b, float) else b) for a, b in v.items()}, flush=True)
    torch.cuda.synchronize(); dist.barrier()
    for b in (gbuf, sig, pbuf, sbuf):
        b.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
