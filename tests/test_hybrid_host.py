"""Host-side arithmetic of the hybrid recall paths (omnia_b200/hybrid.py) against the oracle's restatement
and the reference's own unit tests.  No GPU: the cosine arm is replaced by a stub store."""
import math

import pytest

from omnia_b200 import hybrid as H
from oracle import oracle as O


class StubStore:
    """Answers cosine_ranked from a fixed list, recording the fanout it was asked for."""

    def __init__(self, ranked):
        self.ranked, self.fanouts, self.filters = ranked, [], []

    def cosine_ranked(self, flt, query_embedding, fanout):
        self.fanouts.append(fanout)
        self.filters.append(flt)
        return [(e, 0.0, i + 1) for i, e in enumerate(self.ranked[:fanout])]


def test_rrf_matches_the_oracle_and_the_sql_formula():
    fts = ["e3", "e1", "e9"]
    cos = ["e1", "e2", "e3", "e4"]
    got = H.rrf_fuse(fts, cos)
    want = O.rrf_fuse(fts, cos)
    assert got.keys() == want.keys()
    for e in got:
        assert got[e] == want[e]
    assert got["e1"] == 1.0 / 62.0 + 1.0 / 61.0          # store.go:279-280
    assert got["e9"] == 1.0 / 63.0 and got["e4"] == 1.0 / 64.0


def test_source_type_weights_and_fanout():
    # store.go:170-178, store_read.go:365-372
    assert [H.source_type_weight(s) for s in ("user_requested", "operator_curated", "reflection",
                                              "conversation_extraction", "system_generated", "other", None)] \
        == [1.0, 1.0, 0.85, 0.7, 0.5, 0.7, 0.7]
    assert [H.single_tier_fanout(n) for n in (0, 1, 10, 11, 50)] == [250, 50, 50, 55, 250]


def test_single_tier_final_score_and_fallback():
    rows = {f"e{i}": H.EntityRow(f"e{i}", confidence=0.8, source_type="reflection", age_seconds=2592000.0)
            for i in range(1, 5)}
    rows["e2"].confidence = 0.4
    s = StubStore(["e1", "e2", "e3", "ghost"])
    scope = {"workspace_id": "w"}
    got = H.retrieve_hybrid(s, scope, "refund", [1.0], ["e3"], rows, limit=3)
    assert s.fanouts == [50]
    want = {"e1": (1 / 61.0) * 0.85 * 0.8 * math.exp(-1.0),
            "e3": (1 / 61.0 + 1 / 63.0) * 0.85 * 0.8 * math.exp(-1.0),
            "e2": (1 / 62.0) * 0.85 * 0.4 * math.exp(-1.0)}
    assert [m.entity_id for m in got] == ["e3", "e1", "e2"]      # "ghost" has no projection row: dropped by the JOIN
    for m in got:
        assert m.score == pytest.approx(want[m.entity_id], rel=1e-15)
    assert H.retrieve_hybrid(s, scope, "", [1.0], [], rows) is None       # store_read.go:361-363
    assert H.retrieve_hybrid(s, scope, "q", [], [], rows) is None
    with pytest.raises(H.RecallError, match="workspace_id is required"):
        H.retrieve_hybrid(s, {}, "q", [1.0], [], rows)
    assert len(H.retrieve_hybrid(s, scope, "q", [1.0], [], rows, limit=2)) == 2


def test_tier_classification_and_rankers():
    # retrieve_multi_tier.go:437-450; tier_ranking_test.go:15; ee/pkg/memory/tier_ranking.go:28-38
    assert H.classify_tier(None, None) == H.TIER_INSTITUTIONAL
    assert H.classify_tier("u", None) == H.TIER_USER
    assert H.classify_tier(None, "a") == H.TIER_AGENT
    assert H.classify_tier("u", "a") == H.TIER_USER_FOR_AGENT
    for t in (H.TIER_INSTITUTIONAL, H.TIER_AGENT, H.TIER_USER, H.TIER_USER_FOR_AGENT):
        assert H.IdentityTierRanker().adjust(0.37, t) == 0.37
    m = H.MultiplicativeTierRanker({H.TIER_USER: 2.0, H.TIER_INSTITUTIONAL: 0.5})
    assert m.adjust(1.0, H.TIER_USER_FOR_AGENT) == 2.0 and m.adjust(1.0, H.TIER_AGENT) == 1.0
    assert m.adjust(1.0, H.TIER_INSTITUTIONAL) == 0.5


def test_rank_hybrid_results_is_a_stable_descending_sort():
    mk = lambda e, s, t: H.ScoredMemory(e, s, H.EntityRow(e), t)
    r = [mk("a", 0.5, H.TIER_USER), mk("b", 0.7, H.TIER_AGENT), mk("c", 0.5, H.TIER_USER), mk("d", 0.7, H.TIER_AGENT)]
    H.rank_hybrid_results(r)
    assert [m.entity_id for m in r] == ["b", "d", "a", "c"]
    H.rank_hybrid_results(r, H.MultiplicativeTierRanker({H.TIER_USER: 2.0}))
    assert [m.entity_id for m in r] == ["a", "c", "b", "d"] and r[0].score == 1.0


def test_multi_tier_score_half_life_tiers_and_limit():
    day = 86400.0
    rows = {
        "i": H.EntityRow("i", confidence=0.9, age_seconds=30 * day),
        "u": H.EntityRow("u", confidence=0.9, age_seconds=30 * day, virtual_user_id="u1"),
        "a": H.EntityRow("a", confidence=None, age_seconds=0.0, agent_id="a1"),
    }
    s = StubStore(["i", "u", "a"])
    req = H.MultiTierRequest(workspace_id="w", user_id="u1", agent_id="a1", query="topic", limit=10)
    got = H.retrieve_multi_tier_hybrid(s, req, [1.0], [], rows)
    assert s.fanouts == [100]
    by = {m.entity_id: m for m in got}
    assert by["i"].score == pytest.approx((1 / 61.0) * 0.7 * 0.9 * 0.5, rel=1e-12)     # one half-life old
    assert by["a"].score == pytest.approx((1 / 63.0) * 0.7 * 0.7 * 1.0, rel=1e-12)     # coalesce(confidence, 0.7)
    assert (by["i"].tier, by["u"].tier, by["a"].tier) == (H.TIER_INSTITUTIONAL, H.TIER_USER, H.TIER_AGENT)
    # retrieve_multi_tier_hybrid_test.go:181-213: the half-life, not a fixed bias, decides the order
    req.half_life = H.TierHalfLife(user=1.0, institutional=365 * day)
    assert H.retrieve_multi_tier_hybrid(s, req, [1.0], [], rows)[0].entity_id in ("a", "i")
    got = H.retrieve_multi_tier_hybrid(s, req, [1.0], [], rows)
    assert [m.entity_id for m in got if m.entity_id != "a"] == ["i", "u"]
    assert next(m for m in got if m.entity_id == "u").score == (1 / 62.0) * 0.7 * 0.9 * math.exp(-700.0)
    req.half_life = H.TierHalfLife(user=365 * day, institutional=1.0)
    got = H.retrieve_multi_tier_hybrid(s, req, [1.0], [], rows)
    assert [m.entity_id for m in got if m.entity_id != "a"] == ["u", "i"]
    # filterTiers and truncation
    req.tiers = [H.TIER_USER, H.TIER_AGENT]
    assert {m.entity_id for m in H.retrieve_multi_tier_hybrid(s, req, [1.0], [], rows)} == {"u", "a"}
    req.tiers, req.limit = (), 2
    assert len(H.retrieve_multi_tier_hybrid(s, req, [1.0], [], rows)) == 2
    req.limit = 0
    many = {f"e{i:03d}": H.EntityRow(f"e{i:03d}", confidence=0.9) for i in range(40)}
    assert len(H.retrieve_multi_tier_hybrid(StubStore(sorted(many)), req, [1.0], [], many)) == 15
    # guards: retrieve_multi_tier_hybrid_test.go:217-224 and :109-124
    with pytest.raises(H.RecallError, match="workspace_id"):
        H.retrieve_multi_tier_hybrid(s, H.MultiTierRequest(query="q"), [1.0], [], rows)
    assert H.retrieve_multi_tier_hybrid(s, H.MultiTierRequest(workspace_id="w", query="zlexmatch"), None, [], rows) is None


def test_semantic_over_fetch_and_deny_filter():
    # api/service_semantic.go:47-86; api/service_test.go:898-942
    assert [H.over_fetch_limit(n) for n in (0, -1, 1, 10, 50, 51, 1000)] == [200, 200, 4, 40, 200, 200, 200]
    mems = [H.ScoredMemory(f"e{i}", 1.0 - i * 0.01, H.EntityRow(f"e{i}", metadata={"restricted": i % 2 == 0}))
            for i in range(12)]
    asked = []
    def search(n):
        asked.append(n)
        return mems[:n]
    got = H.retrieve_semantic(search, lambda md: not md["restricted"], 3)
    assert asked == [12] and [m.entity_id for m in got] == ["e1", "e3", "e5"]
    assert len(H.retrieve_semantic(search, lambda md: True, 0)) == 12
