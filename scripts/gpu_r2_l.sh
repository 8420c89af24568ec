#!/bin/bash
cd "${GRAFT_REPO_ROOT:-/root/repo}"; mkdir -p gpurun_out; : > gpurun_out/summary.txt
for q in 256 512 1024; do
  OMNIA_BENCH_SUSTAINED_S=0 OMNIA_BENCH_ROWS=125000 OMNIA_BENCH_QUERIES=$q timeout 200 python bench.py --steps 100 --warmup 10 --no-cpu-baseline > gpurun_out/h_$q.log 2>&1
  python - <<PY >> gpurun_out/summary.txt
import json
d=json.loads([l for l in open("gpurun_out/h_$q.log") if l.startswith("{")][-1]); c=d["config"]
print("125k q=$q ms %.4f kernel %.4f unpip %.4f lat %.4f e2e_ms %.4f launches/step %.1f" % (d["ms_per_step"], d["roofline"]["kernel_ms"], d["roofline"]["kernel_ms_unpipelined"], c["single_step_latency_ms"], d["e2e"]["ms_per_step"], d["gpu_launches"]/d["steps"]))
PY
done
