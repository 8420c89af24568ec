"""Scorer experiments (diagnostics): device time of the scorer alone, variants interleaved so that the
power-limited clock state hits every arm alike."""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from omnia_b200 import GPUMemoryStore, synth, _lib as L
n, dim, nq = 1_000_000, int(os.environ.get("DIM", 1536)), int(os.environ.get("NQ", 1024))
s = GPUMemoryStore(dim, n)
s.bulk_load(synth.gauss_unit(n, dim))
lib = L.load()
arms = [("1-CTA 4 stages", 2, {}), ("2-CTA 6 stages", 4, {}), ("1-CTA epilogue off", 2, {"OMR_UMMA_EXP": "1"}),
        ("1-CTA TMA off", 2, {"OMR_UMMA_EXP": "2"})]
res = {a[0]: [] for a in arms}
for rep in range(int(os.environ.get("REPS", 6))):
    for name, path, env in arms:
        for k in ("OMR_UMMA_EXP", "OMR_UMMA_STAGES"): os.environ.pop(k, None)
        os.environ.update(env)
        ms = ctypes.c_float(0)
        L.check(lib.omr_debug_time_scorer(s.handle, nq, path, 128, 20, ctypes.byref(ms)))
        res[name].append(ms.value)
for name, v in res.items():
    print(f"{name:24s} median {np.median(v):.3f} ms ({2.0*nq*n*dim/np.median(v)/1e9:.0f} TFLOP/s)  all " + " ".join("%.3f" % x for x in v), flush=True)
