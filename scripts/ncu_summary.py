"""Summarise an .ncu-rep (captured on the GPU box) into a small CSV for profiles/ — run on the CPU box:
    python scripts/ncu_summary.py gpurun_out/prof.ncu-rep profiles/ncu_summary.csv"""
import csv, subprocess, sys
METRICS = ["gpu__time_duration.sum", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
           "sm__throughput.avg.pct_of_peak_sustained_elapsed", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
           "dram__bytes_read.sum", "dram__bytes_write.sum", "sm__warps_active.avg.pct_of_peak_sustained_active",
           "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
           "launch__shared_mem_per_block_dynamic", "sm__cycles_elapsed.max",
           "lts__t_sectors_srcunit_tex_op_read.sum", "lts__t_bytes.sum"]
raw = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr, units = rows[0], rows[1]
cols = [hdr.index("Kernel Name")] + [hdr.index(m) for m in METRICS if m in hdr]
with open(sys.argv[2], "w", newline="") as f:
    w = csv.writer(f)
    w.writerow([hdr[c] + (f" [{units[c]}]" if units[c] else "") for c in cols])
    for r in rows[2:]:
        w.writerow([r[c][:90] for c in cols])
print("wrote", sys.argv[2], len(rows) - 2, "kernels")
