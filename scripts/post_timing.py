"""Per-phase cycle breakdown of the fused post kernel (diagnostics; OMR_POST_TIMING=1)."""
import os, sys
os.environ["OMR_POST_TIMING"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, ctypes
from omnia_b200 import GPUMemoryStore, synth, _lib as L
n, dim, nq, k = int(os.environ.get("ROWS", 1_000_000)), 1536, 1024, 10
corpus = synth.gauss_unit(n, dim)
q = torch.from_numpy(synth.queries_from(corpus, nq)).cuda()
s = GPUMemoryStore(dim, n, sub_bucket=int(os.environ.get("S", 0)))
s.bulk_load(corpus)
r = torch.empty((nq, k), dtype=torch.int64, device="cuda"); d = torch.empty((nq, k), dtype=torch.float64, device="cuda")
c = torch.empty(nq, dtype=torch.int32, device="cuda")
for _ in range(3):
    s.device_search(q, k, r, d, c)
torch.cuda.synchronize()
t = np.zeros((nq, 8), np.int64)
L.check(L.load().omr_debug_post_timing(s.handle, t.ctypes.data, nq))
ph = np.diff(t, axis=1)
names = ["pass1 (max)", "sort 512 maxima", "pass2 (collect)", "rescore", "sort+write+topup", "certificate pass", "rescue"]
print("S =", s.stats().last_sub_bucket, "block start spread (cycles):", int(t[:, 0].max() - t[:, 0].min()))
for i, nm in enumerate(names):
    print(f"{nm:20s} median {int(np.median(ph[:, i])):8d}  p90 {int(np.percentile(ph[:, i], 90)):8d}  max {int(ph[:, i].max()):8d} cycles")
tot = t[:, 7] - t[:, 0]
print(f"{'block total':20s} median {int(np.median(tot)):8d}  p90 {int(np.percentile(tot, 90)):8d}  max {int(tot.max()):8d} cycles")
print("kernel span (first start -> last end):", int(t[:, 7].max() - t[:, 0].min()), "cycles")
