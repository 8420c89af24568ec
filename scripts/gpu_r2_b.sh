#!/bin/bash
# Round-2 iteration: pipeline test, headline bench with phases (default path and path 2), ncu captures.
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
: > gpurun_out/summary.txt
run() { local name=$1 secs=$2; shift 2
  echo "=== $name" | tee -a gpurun_out/summary.txt
  timeout -k 5 "$secs" "$@" > "gpurun_out/$name.log" 2>&1
  local rc=$?; echo "exit $rc" | tee -a gpurun_out/summary.txt
  tail -n ${TAILN:-6} "gpurun_out/$name.log" | cut -c1-3500 | tee -a gpurun_out/summary.txt
  return $rc; }
run pipeline 600 python -m pytest tests/test_gpu_pipeline.py tests/test_gpu_parity.py -q -x --timeout 300 -k "not full_size"
run bench_default 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline
OMNIA_BENCH_PATH=2 run bench_path2 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline
if [ "${NCU:-1}" = "1" ]; then
  run ncu_launches 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
      --log-file gpurun_out/r02_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline
  run ncu_full 900 ncu --set full --clock-control none --import-source on -k regex:"score_umma|emit_" -s 12 -c 6 \
      -f -o gpurun_out/r02_prof_recall python bench.py --steps 2 --warmup 3 --no-cpu-baseline
fi
