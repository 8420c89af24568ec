#!/bin/bash
# Programmatic dependent launch of the seed chain: correctness subset, then C2 and the N=8 shard size with the
# chain off / on and with a smaller seed.
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
run tests_pdl 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_adversarial.py tests/test_gpu_golden.py tests/test_gpu_kat.py tests/test_gpu_concurrency.py -q -m gpu -x --timeout 300
B="python bench.py --steps 20 --warmup 3 --no-cpu-baseline"
OMR_PDL=0 run c2_off 300 $B
OMR_PDL=1 run c2_on 300 $B
OMR_PDL=1 OMR_SEED_TILES=4 run c2_on_seed4 300 $B
OMR_PDL=0 OMNIA_BENCH_ROWS=125000 run s125k_off 300 $B
OMR_PDL=1 OMNIA_BENCH_ROWS=125000 run s125k_on 300 $B
OMR_PDL=1 OMR_SEED_TILES=4 OMNIA_BENCH_ROWS=125000 run s125k_on_seed4 300 $B
OMR_PDL=0 OMR_SEED_TILES=4 OMNIA_BENCH_ROWS=125000 run s125k_off_seed4 300 $B
cat gpurun_out/summary.txt
