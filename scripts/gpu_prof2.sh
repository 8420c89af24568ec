#!/bin/bash
# Round-1 evidence: launch list of the default bench command, full captures of the post kernel and of the
# single-query bulk-staged scorer (C5 shard), smoke().
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/smoke.log
python bench.py --steps 4 --warmup 3 --no-cpu-baseline > gpurun_out/plain.log 2>&1 || { echo plain failed; tail gpurun_out/plain.log; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_r01b.csv \
    python bench.py --steps 4 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_l.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"post_kernel" -s 4 -c 1 -f -o gpurun_out/prof_post \
    python bench.py --steps 4 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_p.log 2>&1
BATCHES=1 python scripts/bench_configs.py c5 > gpurun_out/plain_c5.log 2>&1 || { echo plain c5 failed; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:"score_stream1" -s 4 -c 1 -f -o gpurun_out/prof_stream1 \
    python scripts/bench_configs.py c5 > gpurun_out/ncu_s.log 2>&1
ls -la gpurun_out/*.ncu-rep
