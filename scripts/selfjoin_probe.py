"""Self-join probe (diagnostics): one near-duplicate join over N planted-duplicate rows, printing wall time,
kernel time and pair count.  Meant to be run under ncu for a capture of the join kernel on a short join.
    ROWS=100000 DIM=1536 [CORPUS=planted|clustered] [FLOOR=0.92] [LIMIT=0] python scripts/selfjoin_probe.py"""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from omnia_b200 import GPUMemoryStore, synth
n, dim = int(os.environ.get("ROWS", 100_000)), int(os.environ.get("DIM", 1536))
floor, limit = float(os.environ.get("FLOOR", 0.92)), int(os.environ.get("LIMIT", 0))
if os.environ.get("CORPUS", "planted") == "clustered":
    corpus = synth.clustered(n, dim)
else:
    corpus, _ = synth.planted_dups(n, dim, n_pairs=500)
with GPUMemoryStore(dim, n, force_path=int(os.environ.get("FPATH", 0))) as s:
    s.bulk_load(corpus)
    s.set_profiling(True)
    for rep in range(int(os.environ.get("REPS", 2))):
        st0 = s.stats()
        t0 = time.perf_counter()
        pairs = s.run_entity_duplicate_candidates("default", floor, limit, require_same_kind=False)
        dt = time.perf_counter() - t0
        st1 = s.stats()
        kms = st1.scorer_ms_sum - st0.scorer_ms_sum
        print(json.dumps({"rows": n, "dim": dim, "floor": floor, "limit": limit, "pairs": len(pairs), "wall_s": dt,
                          "kernel_ms": kms, "tflops": (n * n * dim / 1e12) / (kms / 1e3) if kms > 0 else None,
                          "top": [(p.canonical_row, p.dupe_row, p.similarity) for p in pairs[:3]]}), flush=True)
