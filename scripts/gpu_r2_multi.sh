#!/bin/bash
# Multi-GPU programme for N = $1 ranks (default 8).  What runs depends on N; every step has its own timeout.
N=${1:-8}
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
: > gpurun_out/summary_n$N.txt
run() { local name=$1 secs=$2; shift 2
  echo "=== $name" | tee -a gpurun_out/summary_n$N.txt
  timeout -k 5 "$secs" "$@" > "gpurun_out/$name.log" 2>&1
  local rc=$?; echo "exit $rc" | tee -a gpurun_out/summary_n$N.txt
  grep -E '^\{|passed|failed|PASSED|FAILED|Error' "gpurun_out/$name.log" | tail -n ${TAILN:-4} | cut -c1-2500 | tee -a gpurun_out/summary_n$N.txt
  return $rc; }
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --nproc-per-node $N"
nvidia-smi -L | wc -l > gpurun_out/ngpu_n$N.txt
if [ "$N" = "8" ]; then
  run multi_tests_n8box 420 python -m pytest tests/test_gpu_multi.py -v --timeout 200
fi
if [ "$N" != "1" ]; then
  run c4_n$N 300 $TR --master-port 29721 scripts/bench_c4.py
  run c2_n$N 240 $TR --master-port 29722 bench.py --gpus $N --steps 40 --warmup 5
fi
C3="env OMNIA_BENCH_ROWS=10000000 OMNIA_BENCH_DIM=768 OMNIA_BENCH_K=100"
if [ "$N" = "1" ]; then
  run c3_n1 600 $C3 python bench.py --steps 20 --warmup 3 --no-cpu-baseline
else
  run c3_n$N 400 $C3 $TR --master-port 29723 bench.py --gpus $N --steps 40 --warmup 5
fi
if [ "$N" = "8" ]; then
  run c5_n8 400 env OMNIA_BENCH_ROWS=50000000 OMNIA_BENCH_DIM=384 OMNIA_BENCH_QUERIES=1 $TR --master-port 29724 bench.py --gpus 8 --steps 200 --warmup 5
fi
