#!/usr/bin/env bash
# One multi-GPU gpurun call: tests, then the bench arms, then the NVLS experiments (last).
#   gpurun --gpus N --timeout 1500 -- 'bash scripts/multi_gpu_session.sh [quick|full]'
set -u
mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l)
MODE=${1:-full}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
run() { local name=$1; shift; local t0=$(date +%s); timeout "${TMO:-420}" "$@" > "gpurun_out/m${N}_$name.log" 2>&1; echo "rc=$? ($name, $(( $(date +%s) - t0 )) s)"; grep -h "^{" "gpurun_out/m${N}_$name.log" | cut -c1-1200 | tail -1; }
export B200ZERO_BENCH_TRACE=1
if [ "$MODE" = "full" ]; then
  B200ZERO_TEST_UNVALIDATED=1 B200ZERO_TEST_AG_DGRAD=1 TMO=600 run tests python -m pytest tests -m "gpu and multigpu" -q -rA --timeout 300
  tail -15 "gpurun_out/m${N}_tests.log" | cut -c1-300
fi
TMO=420 run bench_default $TR --master-port 29601 bench.py --gpus $N --steps 20 --warmup 5 --breakdown
TMO=420 run bench_torch   $TR --master-port 29602 bench.py --gpus $N --steps 20 --warmup 5 --impl torch
B200ZERO_AG_DGRAD=1 TMO=300 run bench_ag_dgrad $TR --master-port 29603 bench.py --gpus $N --steps 8 --warmup 4 --skip-e2e --skip-halt --breakdown
if [ "$MODE" = "full" ]; then
  B200ZERO_PULL_AG=2 TMO=300 run bench_pull_v2 $TR --master-port 29604 bench.py --gpus $N --steps 8 --warmup 4 --skip-e2e --skip-halt --breakdown
  if [ "${WITH_NCCL_ARM:-0}" = "1" ]; then
    TMO=300 run bench_nccl $TR --master-port 29605 bench.py --gpus $N --steps 8 --warmup 4 --skip-e2e --skip-halt --impl nccl --breakdown
    B200ZERO_PARAM_SLOTS=2 TMO=300 run bench_slots2 $TR --master-port 29606 bench.py --gpus $N --steps 8 --warmup 4 --skip-e2e --skip-halt --breakdown
  fi
  B200ZERO_TEST_NVLS=1 TMO=300 run nvls_test python -m pytest tests/test_fused_comm.py -m gpu -q -k nvls
  B200ZERO_NVLS=1 TMO=300 run bench_nvls $TR --master-port 29607 bench.py --gpus $N --steps 8 --warmup 4 --skip-e2e --skip-halt --breakdown
  TMO=300 run bench_after_nvls $TR --master-port 29608 bench.py --gpus $N --steps 8 --warmup 4 --skip-e2e --skip-halt
fi
