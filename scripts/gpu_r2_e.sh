#!/bin/bash
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
: > gpurun_out/summary.txt
run() { local name=$1 secs=$2; shift 2
  echo "=== $name" | tee -a gpurun_out/summary.txt
  timeout -k 5 "$secs" "$@" > "gpurun_out/$name.log" 2>&1
  local rc=$?; echo "exit $rc" | tee -a gpurun_out/summary.txt
  tail -n ${TAILN:-15} "gpurun_out/$name.log" | cut -c1-2500 | tee -a gpurun_out/summary.txt
  return $rc; }
run selfjoin 900 python -m pytest tests/test_gpu_selfjoin.py -q -x --timeout 600
run sj_plain 300 python scripts/selfjoin_probe.py
CORPUS=clustered ROWS=200000 DIM=384 FLOOR=0.85 LIMIT=100 run sj_clustered 300 python scripts/selfjoin_probe.py
run suite 1200 python -m pytest tests -q -m gpu -x --timeout 600 --deselect tests/test_gpu_selfjoin.py
