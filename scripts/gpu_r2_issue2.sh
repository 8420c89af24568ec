#!/bin/bash
# After the lighter profiling: subset of tests, C2 three times, 125k shard, C3 shard.
cd "${GRAFT_REPO_ROOT:-/root/repo}"; mkdir -p gpurun_out; : > gpurun_out/summary.txt
run() { local name=$1 secs=$2; shift 2
  timeout -k 5 "$secs" "$@" > "gpurun_out/$name.log" 2>&1; echo "$name exit $?" >> gpurun_out/summary.txt
  grep -E '^\{|passed|failed' "gpurun_out/$name.log" | tail -n 1 | python -c "
import sys, json
for t in sys.stdin:
    try:
        d = json.loads(t)
        c = d['config']; r = d['roofline']; s = c.get('sustained') or {}; ks = r.get('kernel_span') or {}
        print('  value %.0f ms %.4f kernel %.4f frac %.3f lat %.3f sust %.0f (kernel %.4f, frac %.3f, %s MHz) e2e %.0f fb %s listed/q %.0f span %.4f skew %.3f/%.3f idle %.3f post %.3f/%.3f %s' % (d['value'], d['ms_per_step'], r.get('kernel_ms') or 0, r.get('frac') or 0,
              c.get('single_step_latency_ms') or 0, s.get('value') or 0, s.get('scorer_kernel_ms') or 0, (r.get('sustained') or {}).get('frac') or 0, (s.get('clocks') or {}).get('sm_mhz'), d['e2e']['value'], c.get('fallback_queries'),
              c.get('listed_entries', 0) / max(1, d['steps'] * c['queries_per_step']), ks.get('span_ms', 0), ks.get('start_skew_ms', 0), ks.get('end_skew_ms', 0), ks.get('idle_before_next_scorer_ms', 0),
              ks.get('post_kernel_span_ms', 0), ks.get('post_end_after_scorer_end_ms', 0), json.dumps(c.get('remeasured'))))
    except Exception as e:
        print('  ' + t.strip()[:300])
" >> gpurun_out/summary.txt; }
run tests 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_golden.py tests/test_gpu_concurrency.py -q -m gpu -x --timeout 300
B="python bench.py --steps 20 --warmup 5 --no-cpu-baseline"
C3="OMNIA_BENCH_ROWS=1250000 OMNIA_BENCH_DIM=768 OMNIA_BENCH_K=100"
run c2_a 300 $B
run c2_b 300 $B
run s125k 300 env OMNIA_BENCH_ROWS=125000 $B
run c3shard 300 env $C3 $B
run c2_c 300 $B
cat gpurun_out/summary.txt
