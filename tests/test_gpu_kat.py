"""The reference's own known-answer cases at the `<=>` boundary, run through the CUDA engine.
Each test transcribes a Go test (cited) against GPUMemoryStore, whose methods mirror memory.Store."""
import math
import uuid

import numpy as np
import pytest

from omnia_b200 import (GPUMemoryStore, FilterSpec, RecallError, NotFound, SCOPE_WORKSPACE_ID,
                        SCOPE_VIRTUAL_USER_ID, SCOPE_AGENT_ID, MATCH_NULL_ONLY, MATCH_NULL_OR_EQ,
                        ROW_HAS_EMBEDDING, ROW_MUTABLE)

pytestmark = pytest.mark.gpu
DIM = 1536
WS1 = "00000000-0000-0000-0000-000000000001"   # testWorkspace1
WS2 = "00000000-0000-0000-0000-000000000002"


def repeat_float(v, n=DIM):
    return np.full(n, v, np.float32)


def one_hot(pos, n=DIM):
    out = np.zeros(n, np.float32)
    out[pos] = 1.0
    return out


def new_store(**kw):
    return GPUMemoryStore(DIM, 4096, **kw)


def save(store, scope, vec, **kw):
    oid, eid = uuid.uuid4(), uuid.uuid4()
    store.upsert(oid, eid, scope, vec, **kw)
    return str(oid), str(eid)


@pytest.mark.parametrize("path", [0, 1, 2, 3])
def test_find_similar_ranks_by_cosine(path):
    # store_test.go:422-442
    with new_store(force_path=path) as s:
        scope = {SCOPE_WORKSPACE_ID: WS1}
        near, _ = save(s, scope, repeat_float(0.1), kind="preference", confidence=0.9)
        far, _ = save(s, scope, repeat_float(0.9), kind="fact", confidence=0.9)
        m = s.find_similar_observations(scope, repeat_float(0.1), 5, 0.5)
        assert m, "require.NotEmpty"
        assert m[0].similarity > 0.99
        # both rows are parallel to the query: an exact tie in exact arithmetic
        assert {x.observation_id for x in m} == {near, far}
        assert abs(m[0].similarity - 1.0) < 1e-12 and abs(m[1].similarity - 1.0) < 1e-12


@pytest.mark.parametrize("path", [0, 2, 3])
def test_find_similar_honours_threshold(path):
    # store_test.go:891-905
    with new_store(force_path=path) as s:
        scope = {SCOPE_WORKSPACE_ID: WS1}
        oid, _ = save(s, scope, one_hot(0), kind="fact", confidence=0.9)
        assert s.find_similar_observations(scope, one_hot(100), 5, 0.5) == []
        got = s.find_similar_observations(scope, one_hot(100), 5, 0.0)     # d <= maxDistance is inclusive
        assert [x.observation_id for x in got] == [oid] and got[0].similarity == 0.0


def test_dimension_mismatch_is_an_error_not_a_panic():
    # hnsw_ef_search_test.go:126-144
    with new_store() as s:
        scope = {SCOPE_WORKSPACE_ID: WS1}
        save(s, scope, one_hot(1))
        wrong = np.array([1, 2, 3], np.float32)
        with pytest.raises(RecallError, match="different vector dimensions"):
            s.find_similar_observations(scope, wrong, 5, 0.5)
        with pytest.raises(RecallError, match="different vector dimensions"):
            s.cosine_ranked(FilterSpec(WS1), wrong, 50)
        with pytest.raises(RecallError):
            s.upsert(uuid.uuid4(), uuid.uuid4(), scope, wrong)


def test_workspace_required_and_default_k():
    with new_store() as s:
        with pytest.raises(RecallError, match="workspace_id is required"):   # store.go:60
            s.find_similar_observations({}, one_hot(0), 5, 0.5)
        scope = {SCOPE_WORKSPACE_ID: WS1}
        for _ in range(9):
            save(s, scope, one_hot(0))
        assert len(s.find_similar_observations(scope, one_hot(0), 0, 0.5)) == 5   # types.go:137-141


@pytest.mark.parametrize("path", [0, 1, 2, 3])
def test_workspace_isolation_with_foreign_rows_at_distance_zero(path):
    # hnsw_ef_search_test.go:167-187
    with new_store(force_path=path) as s:
        q = one_hot(7)
        for _ in range(60):
            save(s, {SCOPE_WORKSPACE_ID: WS2}, q, kind="fact", confidence=0.9)
        target = np.zeros(DIM, np.float32)
        target[7] = target[8] = 0.7071
        _, eid = save(s, {SCOPE_WORKSPACE_ID: WS1}, target, kind="fact", confidence=0.9)
        got = s.cosine_ranked(FilterSpec(WS1), q, 50)
        assert [e for e, _, _ in got] == [eid]
        assert abs(got[0][1] - (1 - math.sqrt(0.5))) < 1e-6 and got[0][2] == 1
        # and the foreign rows are visible from their own workspace only
        assert len(s.cosine_ranked(FilterSpec(WS2), q, 100)) == 60


def test_scope_modes_user_and_agent():
    # store.go:250-258 (nil = no constraint) and retrieve_multi_tier.go:313-328 (IS NULL / IS NULL OR =)
    with new_store() as s:
        v = one_hot(3)
        rows = {}
        for name, user, agent in [("none", None, None), ("u5", "u5", None), ("u6", "u6", None),
                                  ("u5a9", "u5", "a9"), ("a9", None, "a9")]:
            scope = {SCOPE_WORKSPACE_ID: WS1}
            if user: scope[SCOPE_VIRTUAL_USER_ID] = user
            if agent: scope[SCOPE_AGENT_ID] = agent
            rows[save(s, scope, v)[0]] = name
        def names(f):
            _, _, cnt, hits = s.search(v[None, :], [f], 10)
            from omnia_b200.store import id_str
            return sorted(rows[id_str(hits[j].observation_id)] for j in range(int(cnt[0])))
        assert names(FilterSpec(WS1)) == sorted(rows.values())
        assert names(FilterSpec(WS1, "u5")) == ["u5", "u5a9"]
        assert names(FilterSpec(WS1, "u5", None, "a9")) == ["u5a9"]
        assert names(FilterSpec(WS1, None, MATCH_NULL_ONLY, None, MATCH_NULL_ONLY)) == ["none"]
        assert names(FilterSpec(WS1, "u5", MATCH_NULL_OR_EQ, None, MATCH_NULL_ONLY)) == ["none", "u5"]
        assert names(FilterSpec(WS1, "u5", MATCH_NULL_OR_EQ, "a9", MATCH_NULL_OR_EQ)) == ["a9", "none", "u5", "u5a9"]
        assert names(FilterSpec(WS1, "nobody")) == []


def test_row_state_edits_change_eligibility():
    # forgotten / superseded / valid_until / confidence / kind (store.go:84-86, 250-258; appendix A)
    with new_store() as s:
        scope = {SCOPE_WORKSPACE_ID: WS1}
        v = one_hot(5)
        ids = [save(s, scope, v, kind=("fact" if i != 5 else "preference")) for i in range(7)]
        def count(**kw):
            return len(s.find_similar_observations(scope, v, 50, 0.5, **kw))
        assert count() == 7
        s.forget(ids[1][1]);                      assert count() == 6     # store_delete.go:42
        s.supersede(ids[2][0]);                   assert count() == 5     # consolidation_writes.go:65
        s.set_valid_until(ids[3][0], 1000)
        assert count(now_us=999) == 5 and count(now_us=1000) == 4         # valid_until > now(), strict
        s.set_confidence(ids[4][0], 0.3)
        f = FilterSpec(WS1, min_confidence=0.5, max_distance=0.5, now_us=0)
        assert int(s.search(v[None, :], [f], 50)[2][0]) == 4
        f = FilterSpec(WS1, kinds=["preference"], now_us=0)
        assert int(s.search(v[None, :], [f], 50)[2][0]) == 1
        s.delete_entity(ids[5][1]);               assert int(s.search(v[None, :], [f], 50)[2][0]) == 0
        with pytest.raises(NotFound):
            s.forget(uuid.uuid4())                                         # store.go:69 ErrNotFound


def test_update_embedding_targets_latest_observation_and_zero_vectors_never_match():
    # store_write.go:557-578; pgvector returns NaN for a zero vector, which fails every `<=` cutoff
    with new_store() as s:
        scope = {SCOPE_WORKSPACE_ID: WS1}
        eid = uuid.uuid4()
        o1, o2 = uuid.uuid4(), uuid.uuid4()
        s.upsert(o1, eid, scope, one_hot(1))
        s.upsert(o2, eid, scope, one_hot(2))
        s.update_embedding(eid, one_hot(9))
        got = s.find_similar_observations(scope, one_hot(9), 5, 0.9)
        assert [g.observation_id for g in got] == [str(o2)]
        assert [g.observation_id for g in s.find_similar_observations(scope, one_hot(1), 5, 0.9)] == [str(o1)]
        s.upsert(uuid.uuid4(), uuid.uuid4(), scope, np.zeros(DIM, np.float32))
        assert len(s.find_similar_observations(scope, one_hot(9), 50, -1.0)) == 2
        assert s.find_similar_observations(scope, np.zeros(DIM, np.float32), 50, -1.0) == []


def test_cosine_arm_keeps_nearest_observation_per_entity():
    # store.go:261-272
    with new_store() as s:
        scope = {SCOPE_WORKSPACE_ID: WS1}
        def mix(c):
            v = np.zeros(DIM, np.float32); v[0] = c; v[1] = math.sqrt(1 - c * c); return v
        e10, e11, e12 = uuid.uuid4(), uuid.uuid4(), uuid.uuid4()
        for e, c in [(e10, 0.9), (e10, 0.5), (e11, 0.8), (e12, 0.95), (e12, 0.1)]:
            s.upsert(uuid.uuid4(), e, scope, mix(c))
        got = s.cosine_ranked(FilterSpec(WS1), one_hot(0), 50)
        assert [(e, rn) for e, _, rn in got] == [(str(e12), 1), (str(e10), 2), (str(e11), 3)]
        assert np.allclose([d for _, d, _ in got], [0.05, 0.1, 0.2], atol=1e-6)
        assert len(s.cosine_ranked(FilterSpec(WS1), one_hot(0), 2)) == 2          # LIMIT fanout


def test_pre_write_dedup_thresholds():
    # service_write.go:63-104 with the defaults of types.go:137-141 (0.95 / 0.85, 5 candidates);
    # mirrors TestMemoryService_AutoSupersedesByEmbeddingSimilarity (internal/memory/api/service_test.go:898-942)
    with new_store() as s:
        scope = {SCOPE_WORKSPACE_ID: WS1, SCOPE_VIRTUAL_USER_ID: "u1"}
        def mix(c):
            v = np.zeros(DIM, np.float32); v[0] = c; v[1] = math.sqrt(1 - c * c); return v
        ids = {}
        for name, c in [("same", 0.99), ("close", 0.90), ("far", 0.50)]:
            ids[name] = save(s, scope, mix(c))[0]
        action, top, matches = s.pre_write_dedup(scope, one_hot(0))
        assert action == "auto_supersede" and top.observation_id == ids["same"]
        assert [m.observation_id for m in matches] == [ids["same"], ids["close"]]      # >= 0.85 only, nearest first
        s.supersede(ids["same"])
        action, top, matches = s.pre_write_dedup(scope, one_hot(0))
        assert action == "insert" and top is None and [m.observation_id for m in matches] == [ids["close"]]
        assert s.pre_write_dedup({SCOPE_WORKSPACE_ID: WS1, SCOPE_VIRTUAL_USER_ID: "someone-else"}, one_hot(0))[2] == []


def test_non_finite_vectors_are_refused_like_pgvector():
    # pgvector rejects NaN / infinity when a vector is parsed, so the reference's statement fails as a whole
    with new_store() as s:
        scope = {SCOPE_WORKSPACE_ID: WS1}
        save(s, scope, one_hot(1))
        bad = one_hot(2); bad[5] = np.nan
        inf = one_hot(2); inf[5] = np.inf
        with pytest.raises(RecallError, match="NaN not allowed in vector"):
            s.find_similar_observations(scope, bad, 5, 0.5)
        with pytest.raises(RecallError, match="infinite value not allowed in vector"):
            s.cosine_ranked(FilterSpec(WS1), inf, 10)
        with pytest.raises(RecallError, match="NaN not allowed"):
            s.upsert(uuid.uuid4(), uuid.uuid4(), scope, bad)
        assert len(s.find_similar_observations(scope, one_hot(1), 5, 0.5)) == 1      # the store is unharmed


def test_consolidation_writes_and_scope_deletes_move_the_mask():
    # Rescope / MergeEntities (postgres/consolidation_writes.go:75-127), tombstone GC
    # (tombstone.go:124), DeleteAll / BatchDelete (store_delete.go:60-125), and the re-embed flush
    with new_store() as s:
        v = one_hot(7)
        ws = {SCOPE_WORKSPACE_ID: WS1}
        alice = {SCOPE_WORKSPACE_ID: WS1, SCOPE_VIRTUAL_USER_ID: "alice"}
        bob = {SCOPE_WORKSPACE_ID: WS1, SCOPE_VIRTUAL_USER_ID: "bob"}
        other = {SCOPE_WORKSPACE_ID: WS2, SCOPE_VIRTUAL_USER_ID: "alice"}
        a = [save(s, alice, v) for _ in range(3)]
        b = [save(s, bob, v) for _ in range(2)]
        o = [save(s, other, v) for _ in range(2)]
        def count(scope):
            return len(s.find_similar_observations(scope, v, 50, 0.5))
        assert (count(alice), count(bob), count(ws), count(other)) == (3, 2, 5, 2)

        # Rescope: alice's first entity becomes bob's; a second observation of that entity follows it
        extra = uuid.uuid4()
        s.upsert(extra, a[0][1], alice, v)
        assert count(alice) == 4
        s.rescope([a[0][0]], user_id="bob")
        assert (count(alice), count(bob)) == (2, 4)
        s.rescope([uuid.uuid4()], user_id="carol")                        # unknown id: UPDATE matches nothing
        assert count({SCOPE_WORKSPACE_ID: WS1, SCOPE_VIRTUAL_USER_ID: "carol"}) == 0
        s.rescope([b[0][0]], user_id="")                                  # NULLIF($1, '') -> institutional row
        assert count(bob) == 3
        f = FilterSpec(WS1, None, MATCH_NULL_ONLY, None, None, now_us=0)
        assert int(s.search(v[None, :], [f], 50)[2][0]) == 1

        # MergeEntities: alice's remaining two entities collapse onto the first; per-entity dedup sees one
        ranked = s.cosine_ranked(FilterSpec(WS1, "alice", now_us=0), v, 50)
        assert len(ranked) == 2
        s.merge_entities(a[1][1], [a[2][1]])
        ranked = s.cosine_ranked(FilterSpec(WS1, "alice", now_us=0), v, 50)
        assert [r[0] for r in ranked] == [a[1][1]] and count(alice) == 2
        s.forget(a[1][1])                                                  # the canonical now owns both rows
        assert count(alice) == 0
        # rows merged into a bob entity take bob's entity-level columns
        s.merge_entities(b[1][1], [a[1][1]])
        assert count(bob) == 5 and count(alice) == 0
        s.update_embedding(b[1][1], one_hot(8))                            # targets the newest row of the merged entity
        assert len(s.find_similar_observations(bob, one_hot(8), 5, 0.9)) == 1 and count(bob) == 4

        # tombstone GC of one observation
        s.delete_observation(extra)
        assert count(bob) == 3
        with pytest.raises(NotFound):
            s.delete_observation(uuid.uuid4())

        # BatchDelete in pages, then DeleteAll of the workspace; the other workspace is untouched
        with pytest.raises(RecallError, match="workspace_id is required"):
            s.delete_all({})
        assert s.batch_delete(bob, 1) == 1
        total = 1
        while True:
            nd = s.batch_delete(bob, 1)
            if nd == 0:
                break
            total += nd
        assert total == 2 and count(bob) == 0                              # a[0] and b[1] (which absorbed a[1], a[2])
        assert s.delete_all(ws) == 1 and count(ws) == 0
        assert s.delete_all({SCOPE_WORKSPACE_ID: "no-such-workspace"}) == 0
        assert count(other) == 2

        # flush and reload (re-embed with a new model of the same dimension)
        s.clear()
        assert len(s) == 0 and count(other) == 0
        save(s, other, v)
        assert count(other) == 1
