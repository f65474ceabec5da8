"""Tiny-shape pass over every kernel family, for `compute-sanitizer --tool memcheck|racecheck`
(one tool per gpurun call; run the script plain first). Prints OK when all results match."""
import ctypes as C, sys
from pathlib import Path
import torch
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from distributed_llm_training_gpu_manager_b200.ops import gemm as G, lib as bz, elementwise as E, optim as O, attention as A
from distributed_llm_training_gpu_manager_b200.ops.flash import flash_fwd, flash_bwd
from distributed_llm_training_gpu_manager_b200.ops.ag_gemm import ag_linear_fwd, ptr_array
from distributed_llm_training_gpu_manager_b200.parallel.fused import segment_table

torch.manual_seed(0)
ok = True
def chk(name, a, b, tol=3e-2):
    global ok
    err = (a.float() - b.float()).abs().max().item() / (b.float().abs().max().item() + 1e-6)
    print(f"{name:28s} rel_err={err:.2e}")
    ok &= err < tol

x = torch.randn(300, 136, device="cuda").bfloat16(); w = torch.randn(520, 136, device="cuda").bfloat16()
for cg in (1, 2):
    chk(f"gemm_nt_cg{cg}", G.gemm(x, w, cta_group=cg), x.float() @ w.float().t())
dy = torch.randn(300, 520, device="cuda").bfloat16()
chk("gemm_dgrad", G.linear_dgrad(dy, w), dy.float() @ w.float())
chk("gemm_wgrad", G.linear_wgrad(dy, x), dy.float().t() @ x.float())
h = torch.randn(37, 256, device="cuda").bfloat16(); nw = torch.ones(256, device="cuda").bfloat16()
y, rstd, _ = E.rmsnorm_fwd(h, nw, 1e-5); chk("rmsnorm_fwd", y, E.rmsnorm_ref(h, nw, 1e-5)[0])
dx, dw = E.rmsnorm_bwd(h, nw, y, 1e-5, rstd=rstd); chk("rmsnorm_bwd", dx, E.rmsnorm_bwd_ref(h, nw, y, 1e-5)[0])
gu = torch.randn(50, 512, device="cuda").bfloat16(); E.swiglu_bwd(gu, E.swiglu_fwd(gu))
qkv = torch.randn(256, 4 * 128, device="cuda").bfloat16(); E.rope_(qkv, 128, 3, 128)
z = torch.randn(33, 1024, device="cuda").bfloat16(); E.cross_entropy_(z, torch.randint(0, 1024, (33,), device="cuda"))
B, S, Hq, Hk, hd = 1, 256, 2, 1, 128
qkv = torch.randn(B * S, (Hq + 2 * Hk) * hd, device="cuda").bfloat16()
out, lse = flash_fwd(qkv, B, S, Hq, Hk, hd); chk("flash_fwd", out, A.attention_ref(qkv, B, S, Hq, Hk, hd))
dout = torch.randn_like(out); dqkv = flash_bwd(qkv, out, lse, dout, B, S, Hq, Hk, hd)
xr = qkv.float().requires_grad_(); (A.attention_ref(xr, B, S, Hq, Hk, hd) * dout.float()).sum().backward()
chk("flash_bwd", dqkv, xr.grad, 5e-2)
ws = 4; wq = torch.randn(1024, 256, device="cuda").bfloat16(); xs = torch.randn(300, 256, device="cuda").bfloat16()
shards = [wq[r * 256:(r + 1) * 256].clone() for r in range(ws)]; wf = torch.zeros_like(wq)
chk("ag_gemm", ag_linear_fwd(xs, ptr_array([s.data_ptr() for s in shards]), wf, 1, ws), xs.float() @ wq.float().t())
L = bz.load(required=True); n = 16 * ws * 64
grads = [torch.randn(n, device="cuda").bfloat16() for _ in range(ws)]
segs = segment_table([(0, n, 0, True)], "cuda"); wsb = torch.zeros(L.bz_fused_comm_workspace_bytes() // 4, device="cuda")
shard = torch.zeros(n // ws, device="cuda"); sp, sr = torch.zeros(1, device="cuda"), torch.zeros(1, device="cuda")
flag = torch.zeros(1, dtype=torch.int32, device="cuda")
bz.check(L.bz_fused_reduce_scatter((C.c_void_p * ws)(*[g.data_ptr() for g in grads]), None, segs.data_ptr(), 1,
                                   shard.data_ptr(), 1, 1.0 / ws, wsb.data_ptr(), sp.data_ptr(), sr.data_ptr(),
                                   flag.data_ptr(), 1, 0, 2, ws, 0, 4, bz.stream()), "rs")
chk("fused_rs", shard, torch.stack([g.float() for g in grads]).mean(0)[2 * n // ws:3 * n // ws])
m = torch.randn(4099, device="cuda"); O.adamw_step(m.clone(), torch.zeros_like(m), torch.zeros_like(m), m, lr=1e-3, step=1)
torch.cuda.synchronize()
print("OK" if ok else "MISMATCH")
sys.exit(0 if ok else 1)
