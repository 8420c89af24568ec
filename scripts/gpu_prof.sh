#!/bin/bash
# Launch list + full captures of the scorers (1-CTA and 2-CTA) and the fused post kernel.
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
export OMNIA_BENCH_PATH=${BPATH:-2}
python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/plain.log 2>&1 || { echo plain failed; tail gpurun_out/plain.log; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/launches_p${OMNIA_BENCH_PATH}.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_l.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"score_umma|post_kernel" -s 6 -c 4 -f -o gpurun_out/prof_p${OMNIA_BENCH_PATH} \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_f.log 2>&1
tail -3 gpurun_out/ncu_f.log | cut -c1-300
