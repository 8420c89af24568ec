#!/bin/bash
# First GPU contact: scorers in isolation, reference KATs, parity, smoke, a reduced bench.
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/gpu.txt 2>&1
run() { # name, seconds, cmd...
  local name=$1 secs=$2; shift 2
  echo "=== $name" | tee -a gpurun_out/summary.txt
  timeout -k 5 "$secs" "$@" > "gpurun_out/$name.log" 2>&1
  echo "exit $?" | tee -a gpurun_out/summary.txt
  tail -n 15 "gpurun_out/$name.log" | tee -a gpurun_out/summary.txt
}
run scorers_stream 300 python -m pytest tests/test_gpu_scorers.py -q -k "1-" --timeout 120
run scorers_umma 300 python -m pytest tests/test_gpu_scorers.py -q -k "2-" --timeout 120
run kat 400 python -m pytest tests/test_gpu_kat.py -q --timeout 120
run parity 900 python -m pytest tests/test_gpu_parity.py -q --timeout 300 -k "not full_size"
run selfjoin 400 python -m pytest tests/test_gpu_selfjoin.py tests/test_gpu_merge.py -q --timeout 200
run smoke 200 python __graft_entry__.py smoke
OMNIA_BENCH_ROWS=262144 run bench_small 400 python bench.py --steps 5 --warmup 3 --no-cpu-baseline
