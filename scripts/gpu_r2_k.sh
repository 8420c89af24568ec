#!/bin/bash
cd "${GRAFT_REPO_ROOT:-/root/repo}"; mkdir -p gpurun_out; : > gpurun_out/summary.txt
run() { local name=$1 secs=$2; shift 2
  timeout -k 5 "$secs" "$@" > "gpurun_out/$name.log" 2>&1; echo "$name exit $?" >> gpurun_out/summary.txt
  grep -E 'passed|failed' "gpurun_out/$name.log" | tail -n 1 >> gpurun_out/summary.txt; }
run tests 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_adversarial.py tests/test_gpu_golden.py tests/test_gpu_kat.py -q -x --timeout 300
OMNIA_BENCH_ROWS=125000 run bench_125k 300 python bench.py --steps 40 --warmup 5 --no-cpu-baseline
run bench_1m 300 python bench.py --steps 20 --warmup 3 --no-cpu-baseline
OMNIA_BENCH_ROWS=1250000 OMNIA_BENCH_DIM=768 OMNIA_BENCH_K=100 run bench_c3shard 300 python bench.py --steps 20 --warmup 3 --no-cpu-baseline
OMNIA_BENCH_ROWS=6250000 OMNIA_BENCH_DIM=384 OMNIA_BENCH_QUERIES=1 run bench_c5shard 300 python bench.py --steps 200 --warmup 5 --no-cpu-baseline
python - <<'PY' >> gpurun_out/summary.txt
import json
for n in ["bench_125k","bench_1m","bench_c3shard","bench_c5shard"]:
    try:
        d=json.loads([l for l in open(f"gpurun_out/{n}.log") if l.startswith("{")][-1]); c=d["config"]
        print(n, "value %.0f ms %.4f kernel %.4f unpip %.4f lat %.4f sust %.0f e2e %.0f listed/q %.0f fb %d roof %s %.3f" % (d["value"], d["ms_per_step"], d["roofline"]["kernel_ms"], d["roofline"]["kernel_ms_unpipelined"], c["single_step_latency_ms"], c["sustained"]["value"], d["e2e"]["value"], c["listed_entries"]/d["steps"]/c["queries_per_step"], c["fallback_queries"], d["roofline"]["bound"], d["roofline"]["frac"]))
    except Exception as e: print(n, "ERR", e)
PY
