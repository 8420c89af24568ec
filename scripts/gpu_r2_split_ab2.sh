#!/bin/bash
# Second A/B on one box after the refresh warp learnt to stop mid-sweep: k = 10 at two shard sizes, k = 100 on the C3 shard.
cd "${GRAFT_REPO_ROOT:-/root/repo}"; mkdir -p gpurun_out; : > gpurun_out/summary.txt
run() { local name=$1 secs=$2; shift 2
  timeout -k 5 "$secs" "$@" > "gpurun_out/$name.log" 2>&1; echo "$name exit $?" >> gpurun_out/summary.txt
  grep -E '^\{|passed|failed' "gpurun_out/$name.log" | tail -n 1 | python -c "
import sys, json
for t in sys.stdin:
    try:
        d = json.loads(t); c = d['config']; r = d['roofline']; s = c.get('sustained') or {}; ks = r.get('kernel_span') or {}
        print('  value %.0f ms %.4f kernel %.4f lat %.3f sust %.0f e2e %.0f fb %s listed/q %.0f span %.4f skew %.3f/%.3f idle %.3f' % (d['value'], d['ms_per_step'], r.get('kernel_ms') or 0,
              c.get('single_step_latency_ms') or 0, s.get('value') or 0, d['e2e']['value'], c.get('fallback_queries'),
              c.get('listed_entries', 0) / max(1, d['steps'] * c['queries_per_step']), ks.get('span_ms', 0), ks.get('start_skew_ms', 0), ks.get('end_skew_ms', 0), ks.get('idle_before_next_scorer_ms', 0)))
    except Exception:
        print('  ' + t.strip()[:300])
" >> gpurun_out/summary.txt; }
B="python bench.py --steps 40 --warmup 5 --no-cpu-baseline"
C3="OMNIA_BENCH_ROWS=1250000 OMNIA_BENCH_DIM=768 OMNIA_BENCH_K=100"
for rep in 1 2; do
  for sp in 0 1; do
    run s125k_split${sp}_r$rep 300 env OMNIA_BENCH_ROWS=125000 OMR_EMIT_SPLIT=$sp $B
  done
done
for sp in 0 1; do
  run c2_split$sp 300 env OMR_EMIT_SPLIT=$sp python bench.py --steps 20 --warmup 3 --no-cpu-baseline
done
run c3shard_default 300 env $C3 $B
run c3shard_split1_t256 300 env $C3 OMR_EMIT_TRIGGER=256 $B
run tests 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_adversarial.py tests/test_gpu_golden.py -q -m gpu -x --timeout 300
cat gpurun_out/summary.txt
