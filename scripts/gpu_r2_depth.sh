#!/bin/bash
# Batches in flight (3 vs 4) at the N=8 shard size and at C2, seed chain by programmatic dependent launch.
cd "${GRAFT_REPO_ROOT:-/root/repo}"; mkdir -p gpurun_out; : > gpurun_out/summary.txt
run() { local name=$1 secs=$2; shift 2
  timeout -k 5 "$secs" "$@" > "gpurun_out/$name.log" 2>&1; echo "$name exit $?" >> gpurun_out/summary.txt
  grep -E '^\{|passed|failed' "gpurun_out/$name.log" | tail -n 1 | python -c "
import sys, json
for t in sys.stdin:
    try:
        d = json.loads(t); c = d['config']; r = d['roofline']; s = c.get('sustained') or {}
        print('  value %.0f ms %.4f kernel %.4f lat %.3f sust %.0f e2e %.0f fb %s span %s' % (d['value'], d['ms_per_step'], r.get('kernel_ms') or 0,
              c.get('single_step_latency_ms') or 0, s.get('value') or 0, d['e2e']['value'], c.get('fallback_queries'), json.dumps(r.get('kernel_span'))))
    except Exception:
        print('  ' + t.strip()[:300])
" >> gpurun_out/summary.txt; }
B="python bench.py --steps 40 --warmup 5 --no-cpu-baseline"
for d in 3 4 2; do
  OMNIA_BENCH_DEPTH=$d OMNIA_BENCH_ROWS=125000 run s125k_depth$d 300 $B
done
OMNIA_BENCH_DEPTH=4 OMNIA_BENCH_ROWS=250000 run s250k_depth4 300 $B
OMNIA_BENCH_DEPTH=3 OMNIA_BENCH_ROWS=250000 run s250k_depth3 300 $B
OMNIA_BENCH_DEPTH=4 run c2_depth4 300 python bench.py --steps 20 --warmup 3 --no-cpu-baseline
cat gpurun_out/summary.txt
