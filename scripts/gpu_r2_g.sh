#!/bin/bash
# N=2 rehearsal of the multi-GPU programme
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
: > gpurun_out/summary.txt
run() { local name=$1 secs=$2; shift 2
  echo "=== $name" | tee -a gpurun_out/summary.txt
  timeout -k 5 "$secs" "$@" > "gpurun_out/$name.log" 2>&1
  local rc=$?; echo "exit $rc" | tee -a gpurun_out/summary.txt
  tail -n ${TAILN:-6} "gpurun_out/$name.log" | cut -c1-3000 | tee -a gpurun_out/summary.txt
  return $rc; }
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
run multi 300 python -m pytest tests/test_gpu_multi.py -v --timeout 200 -k "2-"
run bench_n2 300 $TR --nproc-per-node 2 --master-port 29711 bench.py --gpus 2 --steps 20 --warmup 3
run c4_n2 400 $TR --nproc-per-node 2 --master-port 29712 scripts/bench_c4.py
