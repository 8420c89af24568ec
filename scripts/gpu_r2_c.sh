#!/bin/bash
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
: > gpurun_out/summary.txt
for rep in 1 2; do
for f in 0 1 2 3; do
  OMR_RECALL2=$f timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/bench_r2flag${f}_$rep.log 2>&1
  python - <<PY | tee -a gpurun_out/summary.txt
import json
d=json.loads(open("gpurun_out/bench_r2flag${f}_$rep.log").read().strip().splitlines()[-1])
print("flag $f rep $rep value %.0f ms %.4f e2e %.0f sustained %.0f (%.4f ms, kernel %.4f) kernel_ms %.4f lat %.4f" % (d["value"], d["ms_per_step"], d["e2e"]["value"], d["config"]["sustained"]["value"], d["config"]["sustained"]["ms_per_step"], d["config"]["sustained"]["scorer_kernel_ms"], d["roofline"]["kernel_ms_unpipelined"], d["config"]["single_step_latency_ms"]))
PY
done
done
