#!/usr/bin/env python
"""BASELINE config C4: near-duplicate self-join (1M x 1536, cosine >= 0.92), on 1 GPU or block-partitioned over
the ranks of a torchrun launch (matrix replicated, pair space split, device gather of counts then payload).

    [ROWS=1000000 DIM=1536 FLOOR=0.92 LIMIT=0 REPS=2] python scripts/bench_c4.py [--cpu-baseline]
    torchrun --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P scripts/bench_c4.py

Prints one JSON line per repetition (rank 0): seconds = max over ranks of the whole sharded_selfjoin call
(host post-processing and the gather included), kernel_ms = this rank's tensor-core pass, the pair set checked
against the planted pairs.  --cpu-baseline adds the CPU arm BASELINE.md section 4 item 5 promises: a blocked fp32
M @ M.T with the same threshold on all host threads, at a reduced, stated row count."""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.distributed as dist
from omnia_b200 import GPUMemoryStore, synth
from omnia_b200.dist import sharded_selfjoin
from bench import ClockSampler, measured_peaks, host_threads

n, dim = int(os.environ.get("ROWS", 1_000_000)), int(os.environ.get("DIM", 1536))
floor, limit, reps = float(os.environ.get("FLOOR", 0.92)), int(os.environ.get("LIMIT", 0)), int(os.environ.get("REPS", 2))
rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)


def cpu_baseline(corpus, rows):
    """Blocked fp32 sgemm self-join of the first `rows` rows (unit vectors), pairs above the floor counted."""
    m = np.ascontiguousarray(corpus[:rows])
    t0 = time.perf_counter()
    found, blk = 0, 4096
    for i in range(0, rows, blk):
        for j in range(i, rows, blk):
            s = m[i:i + blk] @ m[j:j + blk].T
            if i == j:
                s = np.triu(s, 1)
            found += int((s >= np.float32(floor)).sum())
    dt = time.perf_counter() - t0
    return {"rows": rows, "seconds": dt, "pairs": found, "cores": host_threads(),
            "gflops": rows * rows * dim / dt / 1e9, "seconds_at_full_size_extrapolated": dt * (n / rows) ** 2,
            "kind": "numpy sgemm blocks of 4096 rows + threshold count (BASELINE.md section 4 item 5)"}


t0 = time.perf_counter()
if rank == 0:
    corpus, planted = synth.planted_dups(n, dim, n_pairs=2000)
else:
    corpus, planted = np.empty((n, dim), np.float32), None
if world > 1:   # one generation, NVLink broadcast in 1 GB pieces
    step = max(1, (1 << 30) // (dim * 4))
    for lo in range(0, n, step):
        t = torch.from_numpy(corpus[lo:lo + step]).to(dev)
        dist.broadcast(t, 0)
        if rank != 0:
            corpus[lo:lo + step] = t.cpu().numpy()
setup_gen = time.perf_counter() - t0
peaks, _ = measured_peaks()
with GPUMemoryStore(dim, n, device=local, force_path=int(os.environ.get("C4_PATH", 0))) as s:
    s.bulk_load(corpus)
    torch.cuda.synchronize()
    for rep in range(reps):
        s.set_profiling(True)
        st0 = s.stats()
        cs = ClockSampler(local, 0.02)
        if rank == 0:
            cs.start()
        if world > 1:
            dist.barrier(device_ids=[local])
        t0 = time.perf_counter()
        pairs = sharded_selfjoin(s, "default", floor, limit, require_same_kind=False)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        st1 = s.stats()
        kernel_ms = st1.scorer_ms_sum - st0.scorer_ms_sum
        t = torch.tensor([dt, kernel_ms / 1e3], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        clocks = cs.stop() if rank == 0 else None
        s.set_profiling(False)
        if rank == 0:
            sec, ksec = float(t[0]), float(t[1])
            flops = 2.0 * dim * n * (n + 256) / 2
            want = {(min(int(i), int(j)), max(int(i), int(j))) for i, j, sim in planted if sim >= float(np.float32(floor))}
            got = {(min(p.canonical_row, p.dupe_row), max(p.canonical_row, p.dupe_row)) for p in pairs}
            line = {"config": f"C4: self-join {n} x {dim}, cosine >= {floor}" + (f", LIMIT {limit}" if limit else ""),
                    "n_gpus": world, "seconds": sec, "kernel_seconds_max_over_ranks": ksec, "pairs": len(pairs),
                    "join_candidates_rank0": int(st1.join_candidates - st0.join_candidates),
                    "planted_above_floor": len(want), "planted_found": len(want & got), "unplanted_pairs": len(got - want) if not limit else None,
                    "clocks": clocks, "setup_generate_broadcast_s": round(setup_gen, 1),
                    "roofline": {"bound": "tensor", "achieved": flops / sec / 1e12, "unit": "TFLOP/s (all ranks, whole call)",
                                 "kernel_achieved_per_gpu": flops / world / ksec / 1e12 if ksec > 0 else None,
                                 "peak_per_gpu_sustained": peaks.get("bf16_tflops_sustained"), "peak_per_gpu_burst": peaks["bf16_tflops"],
                                 "frac_of_sustained": flops / world / ksec / 1e12 / peaks.get("bf16_tflops_sustained", peaks["bf16_tflops"]) if ksec > 0 else None}}
            print(json.dumps(line), flush=True)
if rank == 0 and "--cpu-baseline" in sys.argv:
    print(json.dumps({"config": "C4 CPU baseline", **cpu_baseline(corpus, int(os.environ.get("CPU_ROWS", 40_000)))}), flush=True)
if world > 1:
    dist.barrier(device_ids=[local])
    dist.destroy_process_group()
