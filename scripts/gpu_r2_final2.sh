#!/bin/bash
# Final 1-GPU evidence of the round on one box: full suite, smoke, headline bench + reference arm, C1, C3 at N=1,
# the allocation-free probe, then (after the plain runs) the ncu launch list and one full capture of the scorer.
cd "${GRAFT_REPO_ROOT:-/root/repo}"; mkdir -p gpurun_out; : > gpurun_out/summary.txt
run() { local name=$1 secs=$2; shift 2
  timeout -k 5 "$secs" "$@" > "gpurun_out/$name.log" 2>&1; echo "$name exit $?" >> gpurun_out/summary.txt
  grep -E '^\{|passed|failed' "gpurun_out/$name.log" | tail -n 1 | cut -c1-400 >> gpurun_out/summary.txt; }
run suite 900 python -m pytest tests -q -m gpu -x --timeout 300
run smoke 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')"
run bench 400 python bench.py --steps 20 --warmup 5
run bench_ref 400 python bench.py --impl reference --steps 3 --warmup 1
OMNIA_BENCH_ROWS=10000 OMNIA_BENCH_QUERIES=64 run c1 300 python bench.py --steps 200 --warmup 10 --no-cpu-baseline
OMNIA_BENCH_ROWS=10000000 OMNIA_BENCH_DIM=768 OMNIA_BENCH_K=100 run c3_n1 900 python bench.py --steps 20 --warmup 5 --no-cpu-baseline
run noalloc 300 python scripts/noalloc_probe.py
run c4 300 python scripts/bench_c4.py
run ncu_launches 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/r02_launches_final.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline
run ncu_full 600 ncu --set full --clock-control none --import-source on -k regex:"score_umma2|emit_post" -s 8 -c 4 -f -o gpurun_out/r02_prof_recall_final python bench.py --steps 2 --warmup 3 --no-cpu-baseline
cat gpurun_out/summary.txt
