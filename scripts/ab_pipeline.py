"""A/B of the begin/finish pipeline against back-to-back synchronous calls, interleaved in one process so
that clock drift hits both arms alike.  Prints per-arm medians of ms/step."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from omnia_b200 import synth
from omnia_b200.dist import ShardedRecall
n, dim, nq, k = int(os.environ.get("ROWS", 1_000_000)), int(os.environ.get("DIM", 1536)), 1024, int(os.environ.get("K", 10))
corpus = synth.gauss_unit(n, dim)
sr = ShardedRecall(dim, n, 0)
sr.store.bulk_load(corpus)
dq = [torch.from_numpy(synth.queries_from(corpus, nq, seed=s)).cuda() for s in range(4)]
def sync_steps(c):
    for i in range(c): sr.search_device(dq[i % 4], k)
def pipe_steps(c):
    pend = None
    for i in range(c):
        h = sr.begin(dq[i % 4], k, i & 1)
        if pend is not None: sr.finish(pend)
        pend = h
    sr.finish(pend)
def timed(fn, c=20):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record(); fn(c); e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / c
sync_steps(3); pipe_steps(3)
res = {"sync": [], "pipe": []}
for rep in range(6):
    res["sync"].append(timed(sync_steps)); res["pipe"].append(timed(pipe_steps))
for kx, v in res.items():
    print(kx, "ms/step median %.4f  all %s  -> %.0f queries/s" % (np.median(v), " ".join("%.3f" % x for x in v), nq / np.median(v) * 1e3))
