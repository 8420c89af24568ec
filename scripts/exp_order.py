"""Work-order experiment for the CTA-pair recall scorer (diagnostics): device time of the emission-mode scorer
alone for each OMR_RECALL2 setting (bit 0 = row-tile-major order, bit 1 = evict_last on the row tiles), arms
interleaved so the power state hits them alike.  Run under `ncu --metrics dram__bytes_read.sum` with ITERS=1 to
get the DRAM traffic of each arm."""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from omnia_b200 import GPUMemoryStore, synth, _lib as L
n, dim, nq = int(os.environ.get("ROWS", 1_000_000)), int(os.environ.get("DIM", 1536)), int(os.environ.get("NQ", 1024))
iters = int(os.environ.get("ITERS", 20))
s = GPUMemoryStore(dim, n)
s.bulk_load(synth.gauss_unit(n, dim))
lib = L.load()
arms = [("pairs a-fastest", 4, "0"), ("pairs row-major", 4, "1"), ("pairs row-major evict_last", 4, "3"),
        ("pairs a-fastest evict_last", 4, "2"), ("1-CTA", 2, "0")]
res = {a[0]: [] for a in arms}
for rep in range(int(os.environ.get("REPS", 5))):
    for name, path, flag in arms:
        os.environ["OMR_RECALL2"] = flag
        ms = ctypes.c_float(0)
        L.check(lib.omr_debug_time_scorer(s.handle, nq, path, 0, iters, ctypes.byref(ms)))
        res[name].append(ms.value)
for name, v in res.items():
    print(f"{name:28s} median {np.median(v):.3f} ms ({2.0*nq*n*dim/np.median(v)/1e9:.0f} TFLOP/s)  all " + " ".join("%.3f" % x for x in v), flush=True)
