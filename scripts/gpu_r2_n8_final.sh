#!/bin/bash
# N = 8 on the final build: one sharded-recall parity case, then C2 and C3 the way the driver runs them.
N=8
cd "${GRAFT_REPO_ROOT:-/root/repo}"; mkdir -p gpurun_out; : > gpurun_out/summary_n8.txt
run() { local name=$1 secs=$2; shift 2
  timeout -k 5 "$secs" "$@" > "gpurun_out/$name.log" 2>&1; echo "$name exit $?" >> gpurun_out/summary_n8.txt
  grep -E '^\{|passed|failed' "gpurun_out/$name.log" | tail -n 1 | cut -c1-3000 >> gpurun_out/summary_n8.txt; }
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --nproc-per-node $N"
run multi_test_8p2p 200 python -m pytest tests/test_gpu_multi.py -q -k "8-p2p" --timeout 150
run c2f_n8 200 $TR --master-port 29732 bench.py --gpus 8 --steps 20 --warmup 5
run c3f_n8 300 env OMNIA_BENCH_ROWS=10000000 OMNIA_BENCH_DIM=768 OMNIA_BENCH_K=100 $TR --master-port 29733 bench.py --gpus 8 --steps 20 --warmup 5
cat gpurun_out/summary_n8.txt
