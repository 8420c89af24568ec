#!/bin/bash
# Round-2 iteration run: focused parity tests of the emission-mode recall, whole GPU suite, headline bench.
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
: > gpurun_out/summary.txt
run() { local name=$1 secs=$2; shift 2
  echo "=== $name" | tee -a gpurun_out/summary.txt
  timeout -k 5 "$secs" "$@" > "gpurun_out/$name.log" 2>&1
  local rc=$?; echo "exit $rc" | tee -a gpurun_out/summary.txt
  tail -n ${TAILN:-12} "gpurun_out/$name.log" | cut -c1-2500 | tee -a gpurun_out/summary.txt
  return $rc; }
run parity 600 python -m pytest tests/test_gpu_parity.py -q -x --timeout 300 -k "not full_size"
run adversarial 600 python -m pytest tests/test_gpu_adversarial.py -q -x --timeout 400 -s
run suite 900 python -m pytest tests -q -m gpu -x --timeout 600 --deselect tests/test_gpu_adversarial.py
for p in ${PATHS:-0 4}; do
  OMNIA_BENCH_PATH=$p run bench_path$p 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline
done
