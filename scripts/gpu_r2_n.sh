#!/bin/bash
cd "${GRAFT_REPO_ROOT:-/root/repo}"; mkdir -p gpurun_out; : > gpurun_out/summary.txt
for cfg in "125000 1536 10 1024" "1000000 1536 10 1024" "1250000 768 100 1024"; do
  set -- $cfg
  OMNIA_BENCH_ROWS=$1 OMNIA_BENCH_DIM=$2 OMNIA_BENCH_K=$3 OMNIA_BENCH_QUERIES=$4 timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/n_$1_$2.log 2>&1
  python - <<PY >> gpurun_out/summary.txt
import json
d=json.loads([l for l in open("gpurun_out/n_$1_$2.log") if l.startswith("{")][-1]); c=d["config"]
print("$1x$2 k$3 value %.0f ms %.4f kernel %.4f" % (d["value"], d["ms_per_step"], d["roofline"]["kernel_ms"]), d["roofline"]["kernel_span"], "sust", c["sustained"]["scorer_kernel_span"])
PY
done
