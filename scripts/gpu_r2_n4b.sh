#!/bin/bash
N=${1:-4}
cd "${GRAFT_REPO_ROOT:-/root/repo}"; mkdir -p gpurun_out; : > gpurun_out/summary_n${N}b.txt
run() { local name=$1 secs=$2; shift 2
  timeout -k 5 "$secs" "$@" > "gpurun_out/$name.log" 2>&1; echo "$name exit $?" >> gpurun_out/summary_n${N}b.txt
  grep -E '^\{|passed|failed' "gpurun_out/$name.log" | tail -n 1 | cut -c1-300 >> gpurun_out/summary_n${N}b.txt; }
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --nproc-per-node $N"
C3="env OMNIA_BENCH_ROWS=10000000 OMNIA_BENCH_DIM=768 OMNIA_BENCH_K=100"
if [ "$N" = "1" ]; then
  run c3b_n1 900 $C3 python bench.py --steps 20 --warmup 3 --no-cpu-baseline
else
  run c2b_n$N 300 $TR --master-port 29741 bench.py --gpus $N --steps 20 --warmup 3
  run c3b_n$N 500 $C3 $TR --master-port 29742 bench.py --gpus $N --steps 20 --warmup 3
fi
