"""Where the host-buffer pipeline loses time against the device-resident one (1M x 1536, 1024 queries, top-10).
    python scripts/e2e_probe.py"""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from omnia_b200 import GPUMemoryStore, FilterSpec, synth

N, D, NQ, K = int(os.environ.get("PROBE_ROWS", 1_000_000)), 1536, 1024, 10
corpus = synth.gauss_unit(N, D)
batches = [torch.from_numpy(synth.queries_from(corpus, NQ, seed=2 + b)).pin_memory() for b in range(4)]
pinned = [t.numpy() for t in batches]
pageable = [np.array(p) for p in pinned]
flt = [FilterSpec("default", now_us=0)]
outs = [(np.empty((NQ, K), np.int64), np.empty((NQ, K), np.float64), np.empty(NQ, np.int32)) for _ in range(4)]


def host_loop(s, src, depth, steps):
    pend = []
    for i in range(steps):
        pend.append((s.search_rows_begin(src[i % 4], flt, K), i & 3))
        if len(pend) >= depth:
            t, o = pend.pop(0); s.search_rows_finish(t, outs[o])
    for t, o in pend:
        s.search_rows_finish(t, outs[o])


def device_loop(s, dq, depth, steps, bitmaps=None):
    r = [torch.empty((NQ, K), dtype=torch.int64, device="cuda") for _ in range(4)]
    d = [torch.empty((NQ, K), dtype=torch.float64, device="cuda") for _ in range(4)]
    c = [torch.empty(NQ, dtype=torch.int32, device="cuda") for _ in range(4)]
    st = torch.cuda.current_stream().cuda_stream
    pend = []
    for i in range(steps):
        j = i & 3
        pend.append(s.device_search_begin(dq[j], K, r[j], d[j], c[j], stream=st))
        if len(pend) >= depth:
            s.device_search_finish(pend.pop(0))
    for t in pend:
        s.device_search_finish(t)


with GPUMemoryStore(D, N) as s:
    s.bulk_load(corpus)
    dq = [t.cuda() for t in batches]
    res = {}
    for name, fn in [("device depth3", lambda n: device_loop(s, dq, 3, n)),
                     ("host pinned depth3", lambda n: host_loop(s, pinned, 3, n)),
                     ("host pinned depth4", lambda n: host_loop(s, pinned, 4, n)),
                     ("host pinned depth2", lambda n: host_loop(s, pinned, 2, n)),
                     ("host pageable depth3", lambda n: host_loop(s, pageable, 3, n)),
                     ("device depth3 again", lambda n: device_loop(s, dq, 3, n))]:
        fn(4); torch.cuda.synchronize()
        time.sleep(0.5)
        t0 = time.perf_counter(); fn(20); torch.cuda.synchronize()
        res[name] = round(1e3 * (time.perf_counter() - t0) / 20, 4)
    print(json.dumps(res))
