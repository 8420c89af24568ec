#!/bin/bash
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
run() { local name=$1 secs=$2; shift 2
  echo "=== $name" | tee -a gpurun_out/summary.txt
  timeout -k 5 "$secs" "$@" > "gpurun_out/$name.log" 2>&1
  local rc=$?; echo "exit $rc" | tee -a gpurun_out/summary.txt
  tail -n ${TAILN:-8} "gpurun_out/$name.log" | cut -c1-1700 | tee -a gpurun_out/summary.txt
  return $rc; }
run suite 1500 python -m pytest tests -q -m gpu -x --timeout 600 -k "not full_size"
TAILN=12 run post_timing 300 python scripts/post_timing.py
run bench_pipe 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline
OMNIA_BENCH_PIPELINE=0 run bench_nopipe 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline
rm -f gpurun_out/configs.jsonl
BATCHES=1 TAILN=20 run configs 900 python scripts/bench_configs.py c5 small
