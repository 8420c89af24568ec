"""Shared checks for the parity tests.

Acceptance rule (BASELINE.json): returned ids and order bit-exact except ties within 1e-6;
scores within 1e-5 relative of the fp32 reference.  The GPU rescores in fp64, so against the
fp64 oracle the distances agree to ~1e-12; against the pgvector-faithful fp32 oracle they agree
to its own accumulation error (<= 1e-5 relative on the synthetic corpora)."""
import numpy as np

from oracle import oracle as O

TIE = 1e-6
SCORE_RTOL = 1e-5


def assert_topk_matches(g_rows, g_dist, g_cnt, r_rows, r_dist, r_cnt, corpus=None, queries=None,
                        tie=TIE, rtol=SCORE_RTOL, what=""):
    g_rows, g_dist, r_rows, r_dist = map(np.asarray, (g_rows, g_dist, r_rows, r_dist))
    assert np.array_equal(np.asarray(g_cnt), np.asarray(r_cnt)), f"{what}: counts differ {g_cnt} vs {r_cnt}"
    nq = g_rows.shape[0]
    for i in range(nq):
        c = int(r_cnt[i])
        gs, rs = 1.0 - g_dist[i, :c], 1.0 - r_dist[i, :c]
        bad = np.abs(gs - rs) > rtol * np.maximum(np.abs(rs), 1e-3) + 1e-9
        assert not bad.any(), f"{what}: query {i}: similarity off by {np.abs(gs - rs).max():.3e}"
        diff = np.flatnonzero(g_rows[i, :c] != r_rows[i, :c])
        for j in diff:
            # a different id is only legal inside a tie: same distance within `tie`
            assert abs(g_dist[i, j] - r_dist[i, j]) <= tie, (
                f"{what}: query {i} rank {j}: row {g_rows[i, j]} (d={g_dist[i, j]:.9f}) vs "
                f"oracle row {r_rows[i, j]} (d={r_dist[i, j]:.9f})")
            if corpus is not None:
                true_d = O.cosine_distance(corpus[g_rows[i, j]], queries[i], O.VARIANT_F64)
                assert abs(true_d - g_dist[i, j]) <= 1e-9, f"{what}: reported distance is not the row's distance"
        assert len(set(g_rows[i, :c].tolist())) == c, f"{what}: query {i} returned a row twice"
        assert (g_rows[i, c:] == -1).all()
        assert np.all(np.diff(g_dist[i, :c]) >= 0), f"{what}: query {i} not sorted"


def ids_from_rows(n):
    ids = np.zeros((n, 16), np.uint8)
    v = np.arange(n, dtype=np.uint64)
    for j in range(8):
        ids[:, 15 - j] = (v >> np.uint64(8 * j)).astype(np.uint8)
    return ids
