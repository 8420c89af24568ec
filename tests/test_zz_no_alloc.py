"""omr_exact_topk_noalloc -- the entry point for hosts that cannot allocate device memory while
serving (HBM owned by another framework's pool, latency-critical threads, fixed memory budgets).

Collected last on purpose (file name sorts after every other test file).  It calls the new entry
point through the C ABI on caller-owned device buffers and checks it against the existing entry
point, omr_device_search, on the same rows -- twice in a row, because every call shares one
library-resident scratch array that still holds the previous call's contents."""
import numpy as np
import pytest
import torch

from omnia_b200 import GPUMemoryStore, RecallError, synth
from omnia_b200 import _lib as L
from omnia_b200.store import exact_topk_noalloc

pytestmark = pytest.mark.gpu
ATOL = 1e-9   # the tolerance the other exact-path tests use for distances


def existing_entry_point(corpus, query, k):
    n, dim = corpus.shape
    with GPUMemoryStore(dim, n) as s:
        s.bulk_load(corpus)
        dq = torch.from_numpy(query[None, :]).cuda()
        rows = torch.empty((1, k), dtype=torch.int64, device="cuda")
        dist = torch.empty((1, k), dtype=torch.float64, device="cuda")
        cnt = torch.empty(1, dtype=torch.int32, device="cuda")
        s.device_search(dq, k, rows, dist, cnt)
        torch.cuda.synchronize()
        return rows.cpu().numpy()[0], dist.cpu().numpy()[0], int(cnt.item())


def new_entry_point(d_rows, query, k, check_driver_free=False):
    dq = torch.from_numpy(query).cuda()
    rows = torch.empty(k, dtype=torch.int64, device="cuda")
    dist = torch.empty(k, dtype=torch.float64, device="cuda")
    cnt = torch.empty(1, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    before, free_before = torch.cuda.memory_allocated(), torch.cuda.mem_get_info()[0]
    exact_topk_noalloc(d_rows, dq, k, rows, dist, cnt, torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated() == before
    if check_driver_free:   # once the kernels are loaded, the driver's free-memory figure must not move
        assert torch.cuda.mem_get_info()[0] == free_before
    return rows.cpu().numpy(), dist.cpu().numpy(), int(cnt.item())


def test_two_calls_in_a_row_match_the_existing_entry_point():
    # second call: fewer rows, another dimension, a row pitch wider than dim, a zero row, another k --
    # anything stale in the shared scratch from the first call would show up here
    cases = [(20_000, 384, 10, 0), (3_000, 100, 50, 28)]
    for i, (n, dim, k, pad) in enumerate(cases):
        corpus = synth.gauss_unit(n, dim)
        corpus[n // 2] = 0.0                                   # no cosine: must never rank
        query = synth.queries_from(corpus, 1)[0]
        padded = np.zeros((n, dim + pad), np.float32)
        padded[:, :dim] = corpus
        d_rows = torch.from_numpy(padded).cuda()[:, :dim]      # stride(0) = dim + pad
        g_rows, g_dist, g_cnt = new_entry_point(d_rows, query, k, check_driver_free=i > 0)
        r_rows, r_dist, r_cnt = existing_entry_point(corpus, query, k)
        assert g_cnt == r_cnt == k
        assert np.allclose(g_dist, r_dist, rtol=0, atol=ATOL)
        same = g_rows == r_rows
        assert same.all() or np.abs(g_dist[~same] - r_dist[~same]).max() <= 1e-6   # ties only
        assert n // 2 not in g_rows.tolist()


def test_fewer_rows_than_k_and_limits_are_refused_clearly():
    corpus = synth.gauss_unit(7, 64)
    d_rows = torch.from_numpy(corpus).cuda()
    rows, dist, cnt = new_entry_point(d_rows, corpus[3], 16)
    assert cnt == 7 and rows[0] == 3 and dist[0] < 1e-12 and (rows[7:] == -1).all() and np.isinf(dist[7:]).all()
    q = torch.from_numpy(corpus[0]).cuda()
    o_r = torch.empty(200, dtype=torch.int64, device="cuda")
    o_d = torch.empty(200, dtype=torch.float64, device="cuda")
    o_c = torch.empty(1, dtype=torch.int32, device="cuda")
    with pytest.raises(RecallError, match="k must be in 1..128"):
        exact_topk_noalloc(d_rows, q, L.NOALLOC_MAX_K + 1, o_r, o_d, o_c)
    lib = L.load()   # a row count beyond the fixed scratch is refused before anything is launched
    rc = lib.omr_exact_topk_noalloc(d_rows.data_ptr(), L.NOALLOC_MAX_ROWS + 1, 64, 64, q.data_ptr(), 5,
                                    o_r.data_ptr(), o_d.data_ptr(), o_c.data_ptr(), None)
    assert rc == L.E_CAPACITY and b"at most 524288 rows" in lib.omr_last_error()


def test_against_the_oracle_with_duplicates_straddling_the_kth_place():
    """Held to the CPU oracle directly, and on a tie: 40 copies of one row are the nearest neighbours of the
    query, k = 16 cuts through them.  The copies are bit-identical, so their distances are too, and the 16 that
    come back must be the 16 smallest row numbers (the oracle's ORDER BY distance, row) -- on every run."""
    from oracle import oracle as O
    n, dim, k = 9_000, 96, 16
    corpus = synth.gauss_unit(n, dim, seed=17)
    rng = np.random.default_rng(3)
    dup_rows = np.sort(rng.choice(n, 40, replace=False))
    corpus[dup_rows] = corpus[dup_rows[0]]
    query = (corpus[dup_rows[0]] + 0.01 * synth.gauss_unit(1, dim, seed=5)[0]).astype(np.float32)
    d_rows = torch.from_numpy(corpus).cuda()
    r_rows, r_dist, r_cnt = O.search(corpus, query[None, :], k, variant=O.VARIANT_F64)
    assert r_rows[0].tolist() == dup_rows[:k].tolist()
    for _ in range(3):
        g_rows, g_dist, g_cnt = new_entry_point(d_rows, query, k)
        assert g_cnt == k and g_rows.tolist() == r_rows[0].tolist()
        assert np.allclose(g_dist, r_dist[0], rtol=0, atol=ATOL)
