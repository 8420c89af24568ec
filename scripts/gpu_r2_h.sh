#!/bin/bash
# 1-GPU: whole suite, headline bench (+ reference arm), C4 full size with its CPU baseline, ncu captures
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
: > gpurun_out/summary.txt
run() { local name=$1 secs=$2; shift 2
  echo "=== $name" | tee -a gpurun_out/summary.txt
  timeout -k 5 "$secs" "$@" > "gpurun_out/$name.log" 2>&1
  local rc=$?; echo "exit $rc" | tee -a gpurun_out/summary.txt
  tail -n ${TAILN:-5} "gpurun_out/$name.log" | cut -c1-3500 | tee -a gpurun_out/summary.txt
  return $rc; }
run suite 900 python -m pytest tests -q -m gpu -x --timeout 300 --durations=6
run bench 400 python bench.py --steps 20 --warmup 3
run c4_n1 400 python scripts/bench_c4.py --cpu-baseline
CORPUS=planted ROWS=100000 REPS=1 run sj_ncu 400 ncu --set full --clock-control none --import-source on -k regex:score_umma2 -c 1 -f -o gpurun_out/r02_selfjoin_after python scripts/selfjoin_probe.py
run ncu_full 600 ncu --set full --clock-control none --import-source on -k regex:"score_umma2|emit_post" -s 8 -c 4 -f -o gpurun_out/r02_prof_recall2 python bench.py --steps 2 --warmup 3 --no-cpu-baseline
run ncu_launches 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/r02_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline
