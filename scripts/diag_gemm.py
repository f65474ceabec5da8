"""GPU diagnostic: run every (cta_group, major, out_mode) GEMM variant in its own
subprocess (so a hang or fault in one variant cannot mask the others) and print a
result matrix. Usage: python scripts/diag_gemm.py [--one cg amn bmn mode M N K]"""
import json, os, subprocess, sys
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))

def one(cg, amn, bmn, mode, M, N, K):
    import torch
    from distributed_llm_training_gpu_manager_b200.ops import gemm as G
    torch.manual_seed(0)
    a = torch.randn((K, M) if amn else (M, K), device="cuda").bfloat16()
    b = torch.randn((K, N) if bmn else (N, K), device="cuda").bfloat16()
    dt = torch.bfloat16 if mode < 2 else torch.float32
    acc = mode in (1, 3)
    base = torch.randn(M, N, device="cuda").to(dt)
    out = base.clone()
    G.gemm(a, b, a_mn_major=bool(amn), b_mn_major=bool(bmn), out=out, accumulate=acc, cta_group=cg)
    torch.cuda.synchronize()
    ref = (a.t() if amn else a).float() @ (b if bmn else b.t()).float()
    if acc: ref = ref + base.float()
    err = (out.float() - ref).abs()
    print(json.dumps({"max_err": err.max().item(), "mean_err": err.mean().item(), "ref_max": ref.abs().max().item(),
                      "bad_frac": (err > 0.05 * ref.abs().max()).float().mean().item()}))

if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "--one":
        one(*map(int, sys.argv[2:9])); sys.exit(0)
    rows = []
    cases = []
    for cg in (1, 2):
        for (amn, bmn) in ((0, 0), (0, 1), (1, 1), (1, 0)):
            cases.append((cg, amn, bmn, 0, 256, 256, 64))
            cases.append((cg, amn, bmn, 0, 1024, 768, 512))
        for mode in (1, 2, 3):
            cases.append((cg, 0, 0, mode, 384, 512, 320))
        cases.append((cg, 0, 0, 0, 300, 520, 136))
        cases.append((cg, 0, 0, 0, 4096, 4096, 4096))
    for c in cases:
        try:
            r = subprocess.run([sys.executable, __file__, "--one", *map(str, c)], capture_output=True, text=True, timeout=90)
            last = (r.stdout.strip().splitlines() or [""])[-1]
            status = last if r.returncode == 0 else f"rc={r.returncode} {r.stderr.strip()[-300:]}"
        except subprocess.TimeoutExpired:
            status = "TIMEOUT(hang)"
        print(c, status, flush=True)
        rows.append({"case": c, "status": status})
    os.makedirs(ROOT / "gpurun_out", exist_ok=True)
    json.dump(rows, open(ROOT / "gpurun_out" / "diag_gemm.json", "w"), indent=1)
