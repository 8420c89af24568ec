#!/bin/bash
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
run() { local name=$1 secs=$2; shift 2
  echo "=== $name" | tee -a gpurun_out/summary.txt
  timeout -k 5 "$secs" "$@" > "gpurun_out/$name.log" 2>&1
  local rc=$?; echo "exit $rc" | tee -a gpurun_out/summary.txt
  tail -n ${TAILN:-6} "gpurun_out/$name.log" | cut -c1-1500 | tee -a gpurun_out/summary.txt
  return $rc; }
run suite 1800 python -m pytest tests -q -m gpu -x --timeout 900
run bench 600 python bench.py --steps 20 --warmup 3
rm -f gpurun_out/configs.jsonl
run configs 900 python scripts/bench_configs.py c3
