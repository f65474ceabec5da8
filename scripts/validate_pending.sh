#!/usr/bin/env bash
# First GPU call of the next session: run everything that was written after the round-1 GPU budget was spent.
# One box, 1 GPU unless noted; every step has its own time limit; results land in gpurun_out/.
#   gpurun --timeout 1200 -- 'bash scripts/validate_pending.sh tests'          (1 GPU: child-process kernel tests, ~6 min)
#   gpurun --timeout 1500 -- 'bash scripts/validate_pending.sh bench'          (1 GPU: bench.py per opt-in, ~18 min)
#   gpurun --gpus 2 --timeout 1200 -- 'bash scripts/validate_pending.sh multi' (NVLS K1, dgrad, supervisor, N=2)
set -u
mkdir -p gpurun_out
run() { local name=$1; shift; echo "=== $name"; timeout "${TMO:-600}" "$@" > "gpurun_out/pending_$name.log" 2>&1; echo "rc=$? ($name)"; tail -4 "gpurun_out/pending_$name.log"; }
if [ "${1:-tests}" = "tests" ]; then
  # the child-process tests report xfail/xpass; --runxfail turns a failure into a visible failure
  TMO=600 run rmsnorm_v2 python -m pytest tests/test_elementwise.py -m gpu -q --runxfail -k v2
  TMO=600 run gpt2_ops   python -m pytest tests/test_gpt2_ops.py -m gpu -q --runxfail
  TMO=600 run flash_v3   python -m pytest tests/test_attention.py -m gpu -q --runxfail -k v3
  TMO=900 run ag_dgrad   python -m pytest tests/test_ag_gemm_dgrad.py -m gpu -q --runxfail
  TMO=600 run pull_v2    python -m pytest tests/test_fused_comm.py -m gpu -q --runxfail -k pull_allgather_v2
  TMO=600 run nvme_gpu   python -m pytest tests/test_nvme_offload.py -m gpu -q --runxfail
  TMO=600 run adam_overlap python -m pytest tests/test_native_llama.py -m gpu -q --runxfail -k overlap
elif [ "$1" = "bench" ]; then
  TMO=900 run bench_n1   python bench.py --steps 6 --warmup 3
  B200ZERO_ADAM_OVERLAP=1 TMO=900 run bench_n1_adam_overlap python bench.py --steps 6 --warmup 3 --skip-halt
  B200ZERO_FLASH_BWD=3  TMO=900 run bench_n1_flashbwd3 python bench.py --steps 4 --warmup 3 --skip-e2e --skip-halt
  B200ZERO_RMSNORM=2 B200ZERO_SWIGLU=2 TMO=900 run bench_n1_elementwise2 python bench.py --steps 4 --warmup 3 --skip-e2e --skip-halt
  B200ZERO_FLASH_FWD=3  TMO=900 run bench_n1_flash3   python bench.py --steps 4 --warmup 3 --skip-e2e --skip-halt
  B200ZERO_FLASH_FWD=34 TMO=900 run bench_n1_flash34  python bench.py --steps 4 --warmup 3 --skip-e2e --skip-halt
  B200ZERO_FLASH_FWD=3 B200ZERO_FLASH_BWD=3 B200ZERO_RMSNORM=2 B200ZERO_SWIGLU=3 B200ZERO_ADAM_OVERLAP=1 TMO=900 run bench_n1_all_optins python bench.py --steps 6 --warmup 3 --skip-halt
  B200ZERO_GPT2_NATIVE_OPS=1 TMO=600 run gpt2xl_native_ops python bench.py --steps 4 --warmup 3 --skip-e2e --skip-halt --model gpt2-xl --zero-stage 2 --seq-len 1024 --micro-batch 8 --engine module
else
  N=$(nvidia-smi -L | wc -l)
  TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
  # bench first: the NVLS experiments come last because the one 8-GPU call that ran them before bench.py was
  # followed by two bench runs that never finished ("Tried and backed out")
  TMO=600 run bench_default   $TR --master-port 29601 bench.py --gpus $N --steps 8 --warmup 3
  TMO=600 run bench_nccl      $TR --master-port 29602 bench.py --gpus $N --steps 4 --warmup 3 --comm nccl --skip-e2e
  B200ZERO_TEST_UNVALIDATED=1 TMO=600 run peer_stats_pull_v2 python -m pytest tests/test_fused_comm.py -m gpu -q -k peer_stats
  B200ZERO_PULL_AG=2 B200ZERO_STATS_PEER=1 TMO=600 run bench_pull_v2_peer_stats $TR --master-port 29607 bench.py --gpus $N --steps 4 --warmup 3 --skip-e2e
  B200ZERO_TEST_AG_DGRAD=1 TMO=600 run ag_dgrad_engine python -m pytest tests/test_fused_comm.py -m gpu -q -k dgrad
  B200ZERO_AG_DGRAD=1 TMO=600 run bench_ag_dgrad $TR --master-port 29606 bench.py --gpus $N --steps 4 --warmup 3 --skip-e2e
  B200ZERO_TEST_NVLS=1 TMO=600 run nvls_test python -m pytest tests/test_fused_comm.py -m gpu -q -k nvls
  B200ZERO_NVLS=1 TMO=600 run bench_nvls $TR --master-port 29603 bench.py --gpus $N --steps 4 --warmup 3 --skip-e2e
  TMO=600 run bench_after_nvls $TR --master-port 29604 bench.py --gpus $N --steps 4 --warmup 3 --skip-e2e
  B200ZERO_BENCH_NVLS=1 TMO=300 run roofline $TR --master-port 29605 scripts/bench_fused_multigpu.py
fi
grep -h "_OK" gpurun_out/*_check.log 2>/dev/null | cut -c1-300     # timing lines of the child-process kernel checks
grep -h "^{" gpurun_out/pending_bench_*.log 2>/dev/null | cut -c1-400
