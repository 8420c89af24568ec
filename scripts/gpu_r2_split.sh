#!/bin/bash
# Drain (warp 3) and refresh (warp 2) split: correctness subset, then C3 shard, C2 and the N=8 shard size.
cd "${GRAFT_REPO_ROOT:-/root/repo}"; mkdir -p gpurun_out; : > gpurun_out/summary.txt
run() { local name=$1 secs=$2; shift 2
  timeout -k 5 "$secs" "$@" > "gpurun_out/$name.log" 2>&1; echo "$name exit $?" >> gpurun_out/summary.txt
  grep -E '^\{|passed|failed' "gpurun_out/$name.log" | tail -n 1 | python -c "
import sys, json
for t in sys.stdin:
    try:
        d = json.loads(t); c = d['config']; r = d['roofline']; s = c.get('sustained') or {}
        print('  value %.0f ms %.4f kernel %.4f lat %.3f sust %.0f e2e %.0f fb %s listed/q %.0f span %s' % (d['value'], d['ms_per_step'], r.get('kernel_ms') or 0,
              c.get('single_step_latency_ms') or 0, s.get('value') or 0, d['e2e']['value'], c.get('fallback_queries'),
              c.get('listed_entries', 0) / max(1, d['steps'] * c['queries_per_step']), json.dumps(r.get('kernel_span'))))
    except Exception:
        print('  ' + t.strip()[:300])
" >> gpurun_out/summary.txt; }
run tests 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_adversarial.py tests/test_gpu_golden.py tests/test_gpu_kat.py -q -m gpu -x --timeout 300
B="python bench.py --steps 20 --warmup 3 --no-cpu-baseline"
C3="env OMNIA_BENCH_ROWS=1250000 OMNIA_BENCH_DIM=768 OMNIA_BENCH_K=100"
run c3shard 300 $C3 $B
run c3shard_t128 300 $C3 OMR_EMIT_TRIGGER=128 $B
run c3shard_t512 300 $C3 OMR_EMIT_TRIGGER=512 $B
run c2 300 $B
run c2_t64 300 env OMR_EMIT_TRIGGER=64 $B
run s125k 300 env OMNIA_BENCH_ROWS=125000 $B
run s125k_t64 300 env OMNIA_BENCH_ROWS=125000 OMR_EMIT_TRIGGER=64 $B
cat gpurun_out/summary.txt
