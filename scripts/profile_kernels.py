"""Runs each hot kernel a few times at Llama-3-8B shapes so that ncu can capture them
(`ncu --set full -k regex:... python scripts/profile_kernels.py`). Single GPU; the multi-rank
kernels run in their one-device emulation (peers = buffers on the same GPU, flags disabled)."""
import ctypes as C, sys
from pathlib import Path
import torch
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from distributed_llm_training_gpu_manager_b200.ops import gemm as G, lib as bz, elementwise as E
from distributed_llm_training_gpu_manager_b200.ops.flash import flash_fwd, flash_bwd
from distributed_llm_training_gpu_manager_b200.ops.ag_gemm import ag_linear_fwd, ptr_array
from distributed_llm_training_gpu_manager_b200.parallel.fused import segment_table

torch.manual_seed(0)
T, D, F_, Hq, Hk, hd, S, B = 8192, 4096, 14336, 32, 8, 128, 4096, 2
x = torch.randn(T, D, device="cuda").bfloat16()
w = (torch.randn(2 * F_, D, device="cuda") * 0.02).bfloat16()
for _ in range(3):
    y = G.linear_fwd(x, w)                                  # gate_up fwd
    dx = G.linear_dgrad(y, w)                               # dgrad (B MN-major)
    dw = G.linear_wgrad(y, x)                               # wgrad (A, B MN-major)
qkv = torch.randn(T, (Hq + 2 * Hk) * hd, device="cuda").bfloat16()
for _ in range(3):
    out, lse = flash_fwd(qkv, B, S, Hq, Hk, hd)
    dqkv = flash_bwd(qkv, out, lse, torch.randn_like(out), B, S, Hq, Hk, hd)
ws = 8
shards = [w[r * (2 * F_ // ws):(r + 1) * (2 * F_ // ws)].clone() for r in range(ws)]
wf = torch.empty_like(w)
for _ in range(3):
    ag_linear_fwd(x, ptr_array([s.data_ptr() for s in shards]), wf, 0, ws)
# K1 / K3 emulated: 8 "ranks" on one device
L = bz.load(required=True)
n = 218_112_000 // 16 * 16
n = (n // (16 * ws)) * 16 * ws
grads = [torch.randn(n, device="cuda").bfloat16() for _ in range(ws)]
segs = segment_table([(0, n, 0, True)], "cuda")
wsb = torch.zeros(L.bz_fused_comm_workspace_bytes() // 4, device="cuda")
shard = torch.zeros(n // ws, device="cuda", dtype=torch.bfloat16)
sp, sr = torch.zeros(1, device="cuda"), torch.zeros(1, device="cuda")
flag = torch.zeros(1, dtype=torch.int32, device="cuda")
ptrs = (C.c_void_p * ws)(*[g.data_ptr() for g in grads])
for _ in range(3):
    bz.check(L.bz_fused_reduce_scatter(ptrs, None, segs.data_ptr(), 1, shard.data_ptr(), 0, 1.0 / ws, wsb.data_ptr(),
                                       sp.data_ptr(), sr.data_ptr(), flag.data_ptr(), 1, 0, 0, ws, 0, 128, bz.stream()), "rs")
master = torch.randn(n // ws, device="cuda"); m = torch.zeros_like(master); v = torch.zeros_like(master)
params = [torch.zeros(n, device="cuda", dtype=torch.bfloat16) for _ in range(ws)]
pp = (C.c_void_p * ws)(*[p.data_ptr() for p in params])
coef = torch.ones(1, device="cuda")
for _ in range(3):
    bz.check(L.bz_fused_adam_push(pp, None, segs.data_ptr(), 1, master.data_ptr(), m.data_ptr(), v.data_ptr(),
                                  shard.data_ptr(), 0, None, 1e-4, 0.9, 0.999, 1e-8, 0.01, 1, coef.data_ptr(),
                                  flag.data_ptr(), wsb.data_ptr(), 1, 1, 0, ws, 1, 0, 0, 148, bz.stream()), "adam_push")
h = torch.randn(T, D, device="cuda").bfloat16(); nw = torch.ones(D, device="cuda").bfloat16()
for _ in range(3):
    yn, rstd, _ = E.rmsnorm_fwd(h, nw, 1e-5)
    E.rmsnorm_bwd(h, nw, yn, 1e-5, rstd=rstd)
    E.swiglu_fwd(y)
torch.cuda.synchronize()
print("ok")
