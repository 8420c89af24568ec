#!/bin/bash
# Iteration run: GPU suite, then the headline bench for each scorer path given in $PATHS.
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
run() { local name=$1 secs=$2; shift 2
  echo "=== $name" | tee -a gpurun_out/summary.txt
  timeout -k 5 "$secs" "$@" > "gpurun_out/$name.log" 2>&1
  local rc=$?; echo "exit $rc" | tee -a gpurun_out/summary.txt
  tail -n 8 "gpurun_out/$name.log" | cut -c1-1800 | tee -a gpurun_out/summary.txt
  return $rc; }
run scorers 400 python -m pytest tests/test_gpu_scorers.py -q --timeout 120
run suite 1500 python -m pytest tests -q -m gpu -x --timeout 600 -k "not full_size"
for p in ${PATHS:-2 4}; do
  OMNIA_BENCH_PATH=$p run bench_path$p 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline
done
