#!/bin/bash
# Full-size measurement: whole GPU suite, headline bench (+ reference arm), ncu launch list and one full capture.
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
run() { # name, seconds, cmd...
  local name=$1 secs=$2; shift 2
  echo "=== $name" | tee -a gpurun_out/summary.txt
  timeout -k 5 "$secs" "$@" > "gpurun_out/$name.log" 2>&1
  local rc=$?
  echo "exit $rc" | tee -a gpurun_out/summary.txt
  tail -n 12 "gpurun_out/$name.log" | cut -c1-3000 | tee -a gpurun_out/summary.txt
  return $rc
}
nproc > gpurun_out/nproc.txt; free -g >> gpurun_out/nproc.txt
run pytest_gpu 1500 python -m pytest tests -q -m gpu -x --timeout 900
run bench 900 python bench.py
run bench_ref 600 python bench.py --impl reference --steps 5 --warmup 1
if run bench_short 600 python bench.py --steps 2 --warmup 3 --no-cpu-baseline; then
  run ncu_launches 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
      --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline
  run ncu_full 900 ncu --set full --clock-control none --import-source on -k regex:score_umma -s 3 -c 2 \
      -f -o gpurun_out/prof_umma python bench.py --steps 2 --warmup 3 --no-cpu-baseline
fi
