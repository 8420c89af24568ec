cd "${GRAFT_REPO_ROOT:-/root/repo}"; mkdir -p gpurun_out; : > gpurun_out/summary.txt
run() { local name=$1 secs=$2; shift 2
  timeout -k 5 "$secs" "$@" > "gpurun_out/$name.log" 2>&1; echo "$name exit $?" >> gpurun_out/summary.txt
  grep -E '^\{' "gpurun_out/$name.log" | tail -n 1 | python -c "
import sys, json
for t in sys.stdin:
    d = json.loads(t); c = d['config']; r = d['roofline']; s = c.get('sustained') or {}; ks = r.get('kernel_span') or {}
    print('  value %.0f ms %.4f kernel %.4f sust %.0f e2e %.0f fb %s listed/q %.0f span %.4f skew %.3f/%.3f idle %.3f' % (d['value'], d['ms_per_step'], r.get('kernel_ms') or 0,
          s.get('value') or 0, d['e2e']['value'], c.get('fallback_queries'), c.get('listed_entries', 0) / max(1, d['steps'] * c['queries_per_step']), ks.get('span_ms', 0), ks.get('start_skew_ms', 0), ks.get('end_skew_ms', 0), ks.get('idle_before_next_scorer_ms', 0)))
" >> gpurun_out/summary.txt; }
B="python bench.py --steps 20 --warmup 5 --no-cpu-baseline"
for n in 18 12 8 18 12; do
  run s125k_tiles$n 200 env OMNIA_BENCH_ROWS=125000 OMR_SEED_TILES=$n $B
done
run c2_tiles12 200 env OMR_SEED_TILES=12 $B
cat gpurun_out/summary.txt
