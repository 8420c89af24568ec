#!/bin/bash
cd "${GRAFT_REPO_ROOT:-/root/repo}"; mkdir -p gpurun_out; : > gpurun_out/summary.txt
for t in 128 256 512 1024; do
  OMR_EMIT_TRIGGER=$t OMNIA_BENCH_SUSTAINED_S=0 OMNIA_BENCH_ROWS=1250000 OMNIA_BENCH_DIM=768 OMNIA_BENCH_K=100 timeout 200 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/c3_t$t.log 2>&1
  python - <<PY >> gpurun_out/summary.txt
import json
d=json.loads([l for l in open("gpurun_out/c3_t$t.log") if l.startswith("{")][-1]); c=d["config"]
print("trigger $t value %.0f ms %.4f kernel %.4f unpip %.4f listed/q %.0f fb %d" % (d["value"], d["ms_per_step"], d["roofline"]["kernel_ms"], d["roofline"]["kernel_ms_unpipelined"], c["listed_entries"]/d["steps"]/c["queries_per_step"], c["fallback_queries"]))
PY
done
for t in 32 64 128; do
  OMR_EMIT_TRIGGER=$t OMNIA_BENCH_SUSTAINED_S=0 OMNIA_BENCH_ROWS=125000 timeout 200 python bench.py --steps 40 --warmup 5 --no-cpu-baseline > gpurun_out/c2s_t$t.log 2>&1
  python - <<PY >> gpurun_out/summary.txt
import json
d=json.loads([l for l in open("gpurun_out/c2s_t$t.log") if l.startswith("{")][-1]); c=d["config"]
print("125k trigger $t value %.0f ms %.4f kernel %.4f unpip %.4f listed/q %.0f fb %d" % (d["value"], d["ms_per_step"], d["roofline"]["kernel_ms"], d["roofline"]["kernel_ms_unpipelined"], c["listed_entries"]/d["steps"]/c["queries_per_step"], c["fallback_queries"]))
PY
done
