"""Pins the CPU oracle against the known-answer cases the reference's own tests hold at the
`<=>` boundary.  Each case cites the Go test it transcribes.  The
reference tests use one-hot or constant vectors, so similarities are exactly 0, ~0.7071 or 1."""
import math

import numpy as np
import pytest

from oracle import oracle as O

DIM = 1536


def repeat_float(v, n=DIM):  # store_test.go:912-918
    return np.full(n, v, np.float32)


def one_hot(pos, n=DIM):  # store_test.go:926-933
    out = np.zeros(n, np.float32)
    if 0 <= pos < n:
        out[pos] = 1.0
    return out


@pytest.mark.parametrize("variant", [O.VARIANT_F32_SEQ, O.VARIANT_F64, O.VARIANT_F32_VEC])
def test_distance_known_answers(variant):
    assert O.cosine_distance(one_hot(0), one_hot(0), variant) == 0.0
    assert O.cosine_distance(one_hot(0), one_hot(100), variant) == 1.0
    t = np.zeros(DIM, np.float32); t[7] = t[8] = 0.7071  # hnsw_ef_search_test.go:176-178
    assert abs((1.0 - O.cosine_distance(one_hot(7), t, variant)) - math.sqrt(0.5)) < 1e-6
    assert O.cosine_distance(one_hot(3), -one_hot(3), variant) == 2.0
    assert math.isnan(O.cosine_distance(np.zeros(DIM, np.float32), one_hot(1), variant))
    # Parallel constant vectors (store_test.go:429-434).  Sequential fp32 accumulation of 1536
    # equal terms drifts: the pgvector-faithful variant is ~3.5e-5 away from the true distance 0,
    # which is why the reference only asserts `Similarity > 0.99`.  The clamp keeps d in [0, 2].
    d = O.cosine_distance(repeat_float(0.1), repeat_float(0.9), variant)
    assert 0.0 <= d < (1e-12 if variant == O.VARIANT_F64 else 1e-4)


def test_dimension_mismatch_is_an_error():
    # hnsw_ef_search_test.go:126-144 -- wrongDim := []float32{1, 2, 3}
    with pytest.raises(ValueError, match="different vector dimensions"):
        O.cosine_distance(one_hot(1), np.array([1, 2, 3], np.float32))
    with pytest.raises(ValueError):
        O.search(one_hot(1)[None, :], np.array([[1, 2, 3]], np.float32), 5)


def test_find_similar_ranks_by_cosine():
    # store_test.go:422-442.  Both rows are parallel to the query (cosine 1 for both), so the
    # reference's asserted order rests on fp32 rounding: a latent tie.  Pin what IS determined:
    # both rows come back, both with similarity > 0.99, and within 1e-6 of each other.
    rows = np.stack([repeat_float(0.1), repeat_float(0.9)])
    meta = O.RowMeta.uniform(2)
    got = O.find_similar_observations(rows, meta, O.Filter(workspace=1), repeat_float(0.1), 5, 0.5)
    assert sorted(r for r, _ in got) == [0, 1]
    assert all(s > 0.99 for _, s in got)
    assert abs(got[0][1] - got[1][1]) < 1e-4   # fp32 drift, see test_distance_known_answers
    assert got[0][1] >= got[1][1]


def test_find_similar_honours_threshold():
    # store_test.go:891-905: orthogonal one-hot rejected at minSimilarity 0.5
    rows = one_hot(0)[None, :]
    got = O.find_similar_observations(rows, O.RowMeta.uniform(1), O.Filter(workspace=1), one_hot(100), 5, 0.5)
    assert got == []
    # ...and the boundary is inclusive: d <= maxDistance (store_read.go:278)
    got = O.find_similar_observations(rows, O.RowMeta.uniform(1), O.Filter(workspace=1), one_hot(100), 5, 0.0)
    assert [r for r, _ in got] == [0]


def test_default_k_is_five():
    # store_read.go:254-256 / types.go:137-141 DefaultDuplicateCandidateLimit
    rows = np.stack([one_hot(0)] * 9)
    got = O.find_similar_observations(rows, O.RowMeta.uniform(9), O.Filter(workspace=1), one_hot(0), 0, 0.5)
    assert len(got) == 5


def test_workspace_isolation_with_foreign_rows_at_distance_zero():
    # hnsw_ef_search_test.go:167-187: 60 workspace-2 rows equal to the query, one workspace-1
    # target at cosine 0.7071; only the target may surface.
    q = one_hot(7)
    target = np.zeros(DIM, np.float32); target[7] = target[8] = 0.7071
    rows = np.stack([q] * 60 + [target])
    meta = O.RowMeta.uniform(61, workspace=2)
    meta.workspace[60] = 1
    entity = np.arange(61)
    ranked = O.cosine_ranked(rows, meta, entity, O.Filter(workspace=1), q, fanout=50)
    assert [e for e, _, _ in ranked] == [60]
    assert abs(ranked[0][1] - (1 - math.sqrt(0.5))) < 1e-6
    assert ranked[0][2] == 1


def test_scope_predicate_single_and_multi_tier():
    # store.go:250-258 (nil user/agent = no constraint) vs retrieve_multi_tier.go:313-328
    # (empty = IS NULL, set = IS NULL OR = x).
    meta = O.RowMeta.uniform(6)
    meta.user[:] = [0, 5, 6, 5, 0, 5]
    meta.agent[:] = [0, 0, 0, 9, 9, 8]
    f = lambda **kw: list(np.flatnonzero(O.eligible(meta, O.Filter(workspace=1, **kw))))
    assert f() == [0, 1, 2, 3, 4, 5]
    assert f(user=5, user_mode=O.USER_EQ) == [1, 3, 5]
    assert f(user=5, user_mode=O.USER_EQ, agent=9, agent_mode=O.USER_EQ) == [3]
    assert f(user_mode=O.USER_NULL, agent_mode=O.USER_NULL) == [0]
    assert f(user=5, user_mode=O.USER_NULL_OR_EQ, agent_mode=O.USER_NULL) == [0, 1]
    assert f(user=5, user_mode=O.USER_NULL_OR_EQ, agent=9, agent_mode=O.USER_NULL_OR_EQ) == [0, 1, 3, 4]


def test_row_state_predicates():
    # NOT e.forgotten, superseded_by IS NULL, valid_until IS NULL OR > now(), embedding IS NOT
    # NULL, confidence >= min (store.go:84-86, 250-258; store_test.go:770-791 confidence filter)
    meta = O.RowMeta.uniform(7)
    meta.flags[1] |= O.FLAG_FORGOTTEN
    meta.flags[2] |= O.FLAG_SUPERSEDED
    meta.flags[3] &= ~np.uint8(O.FLAG_HAS_EMBEDDING)
    meta.valid_until_us[4] = 1000      # expired at now=1000 (strict >)
    meta.valid_until_us[5] = 1001      # still valid
    meta.confidence[6] = 0.3
    got = list(np.flatnonzero(O.eligible(meta, O.Filter(workspace=1, now_us=1000, min_confidence=0.5))))
    assert got == [0, 5]
    meta.kind[:] = [0, 0, 0, 0, 0, 3, 0]
    assert list(np.flatnonzero(O.eligible(meta, O.Filter(workspace=1, now_us=1000, kinds=(3,))))) == [5]


def test_cosine_arm_keeps_nearest_observation_per_entity():
    # store.go:261-272: DISTINCT ON (entity_id) ... ORDER BY entity_id, cos_dist
    q = one_hot(0)
    def mix(c):  # cosine c to the query
        v = np.zeros(DIM, np.float32); v[0] = c; v[1] = math.sqrt(1 - c * c); return v
    rows = np.stack([mix(0.9), mix(0.5), mix(0.8), mix(0.95), mix(0.1)])
    entity = np.array([10, 10, 11, 12, 12])
    ranked = O.cosine_ranked(rows, O.RowMeta.uniform(5), entity, O.Filter(workspace=1), q, fanout=50)
    assert [(e, rn) for e, _, rn in ranked] == [(12, 1), (10, 2), (11, 3)]
    assert np.allclose([d for _, d, _ in ranked], [0.05, 0.1, 0.2], atol=1e-6)


def test_rrf_constant():
    # store.go:272-282, rrfK = 60
    s = O.rrf_fuse([1, 2], [2, 3])
    assert s[1] == pytest.approx(1 / 61) and s[2] == pytest.approx(1 / 62 + 1 / 61) and s[3] == pytest.approx(1 / 62)


def test_selfjoin_predicates_and_float32_floor():
    # consolidation/prefilter.go:113-131 (SQL shape pinned by prefilter_test.go:62-77; no
    # reference test executes it, so pair values are "parity unpinned").
    def mix(c):
        v = np.zeros(8, np.float32); v[0] = c; v[1] = math.sqrt(max(0.0, 1 - c * c)); return v
    base = mix(1.0)
    rows = np.stack([base, mix(0.95), mix(0.90), base, base])
    ids = np.zeros((5, 16), np.uint8); ids[:, 15] = [5, 4, 3, 2, 1]   # uuid order reversed vs row order
    kind = np.array([1, 1, 1, 2, 1], np.uint16)
    flags = np.full(5, O.FLAG_HAS_EMBEDDING | O.FLAG_MUTABLE, np.uint8)
    flags[4] &= ~np.uint8(O.FLAG_MUTABLE)
    a, b, s = O.selfjoin(rows, None, kind, flags, ids, 0.92)
    # eligible: rows 0,1,2 (kind 1, mutable). cos(0,1)=.95 ok; cos(0,2)=.90 no; cos(1,2)=~0.991 ok
    pairs = {(int(x), int(y)) for x, y in zip(a, b)}
    assert pairs == {(1, 0), (2, 1)}          # canonical = smaller uuid = larger row here
    assert s[0] >= s[1]                        # ORDER BY similarity DESC
    # float32 widening: (double)(float)0.92 = 0.92000001668..., so an exact-0.92 pair is excluded
    assert float(np.float32(0.92)) > 0.92
    a2, _, _ = O.selfjoin(rows, None, kind, flags, ids, 0.92, require_same_kind=False, require_mutable=False)
    assert len(a2) > len(a)
    a3, _, _ = O.selfjoin(rows, None, kind, flags, ids, 0.92, limit=1)
    assert len(a3) == 1


def test_variants_agree_on_random_data():
    rng = np.random.default_rng(7)
    rows = rng.standard_normal((500, 96)).astype(np.float32)
    q = rng.standard_normal((4, 96)).astype(np.float32)
    r0, d0, _ = O.search(rows, q, 10, variant=O.VARIANT_F32_SEQ)
    r1, d1, _ = O.search(rows, q, 10, variant=O.VARIANT_F64)
    r2, d2, _ = O.search(rows, q, 10, variant=O.VARIANT_F32_VEC)
    assert np.array_equal(r0, r1) and np.array_equal(r0, r2)
    assert np.allclose(d0, d1, rtol=0, atol=2e-6) and np.allclose(d2, d1, rtol=0, atol=2e-6)
    rn, dn = O.numpy_topk(rows, q, 10)
    assert np.array_equal(rn, r1) and np.allclose(dn, d1, atol=2e-6)


def test_pca_reduce_restatement_known_answers_and_reference_shapes():
    # ee/pkg/memory/projection/pca_test.go:17-99 (shapes, clamping, degenerate sets) and a hand-checkable case
    t = np.linspace(-1.0, 1.0, 21)
    u = np.array([3.0, 4.0]) / 5.0
    m = np.outer(t, u) + np.array([10.0, -2.0])               # points on a line, far from the origin
    out = O.pca_reduce(m, 1)
    assert out.shape == (21, 1)
    want = m @ u                                               # the UNCENTRED rows times the first axis (pca.go:60)
    assert min(np.abs(out[:, 0] - want).max(), np.abs(out[:, 0] + want).max()) < 1e-12
    data = (np.arange(6 * 8) * 7 % 13 / 13.0).reshape(6, 8)
    assert O.pca_reduce(data, 3).shape == (6, 3)
    assert O.pca_reduce(np.array([[1, 0], [0, 1], [1, 1], [0, 0]], float), 50).shape == (4, 2)
    wide = (np.arange(6 * 1536) * 7 % 97 / 97.0).reshape(6, 1536)
    assert O.pca_reduce(wide, 50).shape == (6, 6)
    one = (np.arange(64) / 64.0)[None, :]
    np.testing.assert_array_equal(O.pca_reduce(one, 50), one[:, :1])
    with pytest.raises(ValueError, match="empty input matrix"):
        O.pca_reduce(np.zeros((0, 0)), 50)
