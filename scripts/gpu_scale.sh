#!/bin/bash
# Scaling run on one box: N = 1, 2, 4, 8 back to back, then the NCCL parity test.
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
nvidia-smi -L | wc -l > gpurun_out/ngpu.txt
python bench.py --gpus 1 --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/scale_n1.log 2>&1; tail -1 gpurun_out/scale_n1.log | cut -c1-400
for n in 2 4 8; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500+n)) \
      bench.py --gpus $n --steps 20 --warmup 3 > gpurun_out/scale_n$n.log 2>&1
  echo "N=$n rc=$?"; tail -1 gpurun_out/scale_n$n.log | cut -c1-400
done
timeout 900 python -m pytest tests/test_gpu_multi.py -q -m gpu > gpurun_out/multi.log 2>&1; tail -3 gpurun_out/multi.log
