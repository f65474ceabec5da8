#!/usr/bin/env bash
# ncu --set full captures of the hot kernels inside a real (2-layer, 8B-shaped) training step, one kernel family per
# report so each .ncu-rep stays small; raw + source pages are exported to CSV next to it. 1 GPU only.
#   gpurun --timeout 1500 -- 'bash scripts/ncu_capture.sh'
set -u
mkdir -p gpurun_out/ncu
cap() {  # name regex count
  timeout 400 ncu --set full --clock-control none --import-source on --profile-from-start off -k "regex:$2" -c "$3" \
      -f -o "gpurun_out/ncu/$1" python bench.py --layers 2 --steps 1 --warmup 3 --skip-e2e --skip-halt > "gpurun_out/ncu/$1.log" 2>&1
  echo "ncu $1 rc=$?"
  ncu -i "gpurun_out/ncu/$1.ncu-rep" --page raw --csv > "gpurun_out/ncu/$1.raw.csv" 2>/dev/null
  ncu -i "gpurun_out/ncu/$1.ncu-rep" --page details --csv > "gpurun_out/ncu/$1.details.csv" 2>/dev/null
  ls -la "gpurun_out/ncu/$1.ncu-rep" | awk '{print $5, $9}'
}
cap flash_fwd  'flash_fwd'  2
cap flash_bwd  'flash_bwd'  2
cap gemm       'gemm_bf16'  8
cap elementwise 'rmsnorm|swiglu|rope|cross_entropy|adamw|sumsq' 12
du -sh gpurun_out/ncu
# keep the call's return payload under the 64 MiB limit: drop the largest reports first
while [ "$(du -sm gpurun_out | cut -f1)" -gt 55 ]; do
  big=$(ls -S gpurun_out/ncu/*.ncu-rep 2>/dev/null | head -1); [ -z "$big" ] && break; echo "dropping $big"; rm -f "$big"
done
