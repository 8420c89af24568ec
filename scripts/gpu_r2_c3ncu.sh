#!/bin/bash
# ncu source-level capture of the emission scorer on the C3 shard (1.25M x 768, top-100), after a plain run.
cd "${GRAFT_REPO_ROOT:-/root/repo}"; mkdir -p gpurun_out; : > gpurun_out/summary.txt
export OMNIA_BENCH_ROWS=1250000 OMNIA_BENCH_DIM=768 OMNIA_BENCH_K=100
timeout -k 5 300 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/c3shard_plain.log 2>&1; echo "plain exit $?" >> gpurun_out/summary.txt
grep -E '^\{' gpurun_out/c3shard_plain.log | tail -n 1 | cut -c1-1500 >> gpurun_out/summary.txt
timeout -k 5 600 ncu --set full --clock-control none --import-source on -k regex:score_umma2 -s 6 -c 1 -f -o gpurun_out/r02_c3shard_emit python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/c3shard_ncu.log 2>&1; echo "ncu exit $?" >> gpurun_out/summary.txt
tail -n 3 gpurun_out/c3shard_ncu.log >> gpurun_out/summary.txt
cat gpurun_out/summary.txt
