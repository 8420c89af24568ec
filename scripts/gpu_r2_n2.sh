#!/bin/bash
cd "${GRAFT_REPO_ROOT:-/root/repo}"; mkdir -p gpurun_out; : > gpurun_out/summary_n2.txt
run() { local name=$1 secs=$2; shift 2
  timeout -k 5 "$secs" "$@" > "gpurun_out/$name.log" 2>&1; echo "$name exit $?" >> gpurun_out/summary_n2.txt
  grep -E '^\{|passed|failed' "gpurun_out/$name.log" | tail -n 1 | cut -c1-500 >> gpurun_out/summary_n2.txt; }
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --nproc-per-node 2"
run multi2 300 python -m pytest tests/test_gpu_multi.py -q --timeout 200 -k "2-"
run c2_n2 300 $TR --master-port 29731 bench.py --gpus 2 --steps 20 --warmup 3
run c3_n2 500 env OMNIA_BENCH_ROWS=10000000 OMNIA_BENCH_DIM=768 OMNIA_BENCH_K=100 $TR --master-port 29732 bench.py --gpus 2 --steps 20 --warmup 3
