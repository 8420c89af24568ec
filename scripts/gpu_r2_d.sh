#!/bin/bash
# suite + bench on the committed emission-mode build; baseline ncu capture of the self-join kernel (100k rows)
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
: > gpurun_out/summary.txt
run() { local name=$1 secs=$2; shift 2
  echo "=== $name" | tee -a gpurun_out/summary.txt
  timeout -k 5 "$secs" "$@" > "gpurun_out/$name.log" 2>&1
  local rc=$?; echo "exit $rc" | tee -a gpurun_out/summary.txt
  tail -n ${TAILN:-6} "gpurun_out/$name.log" | cut -c1-3000 | tee -a gpurun_out/summary.txt
  return $rc; }
run suite 1200 python -m pytest tests -q -m gpu -x --timeout 600
run bench 600 python bench.py --steps 20 --warmup 3
run sj_plain 300 python scripts/selfjoin_probe.py
REPS=1 run sj_ncu 600 ncu --set full --clock-control none --import-source on -k regex:score_umma2 -c 1 -f -o gpurun_out/r02_selfjoin_before python scripts/selfjoin_probe.py
