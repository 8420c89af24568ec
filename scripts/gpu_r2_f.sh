#!/bin/bash
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
: > gpurun_out/summary.txt
run() { local name=$1 secs=$2; shift 2
  echo "=== $name" | tee -a gpurun_out/summary.txt
  timeout -k 5 "$secs" "$@" > "gpurun_out/$name.log" 2>&1
  local rc=$?; echo "exit $rc" | tee -a gpurun_out/summary.txt
  tail -n ${TAILN:-25} "gpurun_out/$name.log" | cut -c1-1500 | tee -a gpurun_out/summary.txt
  return $rc; }
run selfjoin_new 400 python -m pytest tests/test_gpu_selfjoin.py -v -x --timeout 150 --durations=8 -k "limit or clustered or regrow or cancel or while"
run churn 300 python -m pytest tests/test_gpu_churn.py tests/test_zz_no_alloc.py tests/test_gpu_kat.py -v -x --timeout 150 --durations=5
