"""Shape of the power transient: per-block step times of the pipelined recall loop (1M x 1536, 1024 queries)
after different idle periods.  python scripts/power_transient.py"""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from omnia_b200 import GPUMemoryStore, synth

N, D, NQ, K = 1_000_000, 1536, 1024, 10
corpus = synth.gauss_unit(N, D)
dq = [torch.from_numpy(synth.queries_from(corpus, NQ, seed=2 + b)).cuda() for b in range(4)]
r = [torch.empty((NQ, K), dtype=torch.int64, device="cuda") for _ in range(4)]
d = [torch.empty((NQ, K), dtype=torch.float64, device="cuda") for _ in range(4)]
c = [torch.empty(NQ, dtype=torch.int32, device="cuda") for _ in range(4)]


def run(s, steps, block=5, depth=3):
    stream = torch.cuda.current_stream()
    st = stream.cuda_stream
    evs = [torch.cuda.Event(enable_timing=True) for _ in range(steps // block + 1)]
    pend, done = [], 0
    evs[0].record(stream)
    for i in range(steps):
        j = i & 3
        pend.append(s.device_search_begin(dq[j], K, r[j], d[j], c[j], stream=st))
        if len(pend) >= depth:
            s.device_search_finish(pend.pop(0)); done += 1
            if done % block == 0: evs[done // block].record(stream)
    for t in pend:
        s.device_search_finish(t); done += 1
        if done % block == 0: evs[done // block].record(stream)
    torch.cuda.synchronize()
    return [round(evs[i].elapsed_time(evs[i + 1]) / block, 3) for i in range(steps // block)]


if os.environ.get("COMPARE_PATHS"):
    # sustained comparison of scorer variants: 400 steps each, second half reported, alternating
    stores = {p: GPUMemoryStore(D, N, force_path=p) for p in (2, 4)}
    for st_ in stores.values():
        st_.bulk_load(corpus)
    for rep in range(3):
        for p, st_ in stores.items():
            time.sleep(1.0)
            t = run(st_, 400, block=50)
            print(json.dumps({"force_path": p, "rep": rep, "ms_per_step_blocks_of_50": t}), flush=True)
    sys.exit(0)

with GPUMemoryStore(D, N) as s:
    s.bulk_load(corpus)
    run(s, 10)
    for idle in (0.0, 0.25, 0.5, 1.0, 2.0, 0.0):
        time.sleep(idle)
        t = run(s, 60)
        print(json.dumps({"idle_s": idle, "ms_per_step_blocks_of_5": t}), flush=True)
    time.sleep(2.0)
    t = run(s, 600, block=50)
    print(json.dumps({"idle_s": 2.0, "ms_per_step_blocks_of_50": t}), flush=True)
