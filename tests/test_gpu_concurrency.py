"""Concurrent callers (one goroutine per HTTP request in the reference, cmd/memory-api/main.go:1137-1157):
the store must be safe to call from many threads, and the calls that arrive while a search is running must
be coalesced into one pass over the matrix (dedup-on-write micro-batching)."""
import threading

import numpy as np
import pytest

from omnia_b200 import GPUMemoryStore, FilterSpec, RecallError, synth
from oracle import oracle as O
from helpers import assert_topk_matches

pytestmark = pytest.mark.gpu


def test_concurrent_find_similar_is_coalesced_and_exact():
    n, dim, n_threads, per_thread = 50_000, 256, 16, 12
    corpus = synth.gauss_unit(n, dim)
    ws = (np.arange(n) % 3).astype(np.uint32)
    queries = synth.queries_from(corpus, n_threads * per_thread)
    with GPUMemoryStore(dim, n) as s:
        s.bulk_load(corpus, workspaces=["w0", "w1", "w2"], workspace_idx=ws)
        out = [None] * len(queries)
        errs = []

        def worker(t):
            try:
                for j in range(per_thread):
                    i = t * per_thread + j
                    f = FilterSpec(f"w{i % 3}", max_distance=0.9)     # k = 5, a cutoff: the dedup-on-write shape
                    out[i] = s.search_rows(queries[i], [f], 5)
            except Exception as e:  # noqa: BLE001
                errs.append(e)

        threads = [threading.Thread(target=worker, args=(t,)) for t in range(n_threads)]
        for t in threads: t.start()
        for t in threads: t.join()
        assert not errs, errs
        st = s.stats()
        assert st.host_calls == len(queries)
        assert st.host_batches < st.host_calls, "concurrent calls were not coalesced"
        # a bad call among good ones fails alone
        bad = []
        def bad_worker():
            try:
                s.search_rows(queries[0], [FilterSpec("")], 5)
            except RecallError as e:
                bad.append(str(e))
        ths = [threading.Thread(target=bad_worker)] + [threading.Thread(target=worker, args=(t,)) for t in range(4)]
        for t in ths: t.start()
        for t in ths: t.join()
        assert bad and "workspace_id is required" in bad[0] and not errs
    for g in range(3):
        idx = [i for i in range(len(queries)) if i % 3 == g]
        r = O.search(corpus, queries[idx], 5, (ws == g).astype(np.uint8), max_distance=0.9, variant=O.VARIANT_F64)
        g_rows = np.stack([out[i][0][0] for i in idx]); g_dist = np.stack([out[i][1][0] for i in idx])
        g_cnt = np.array([out[i][2][0] for i in idx])
        assert_topk_matches(g_rows, g_dist, g_cnt, r[0], r[1], r[2], what=f"workspace {g}")
