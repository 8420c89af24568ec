"""The reference's hybrid-recall tests with the cosine arm answered by the CUDA engine
(internal/memory/retrieve_multi_tier_hybrid_test.go, internal/memory/store_test.go).  FTS and the
projection stay in Postgres in the drop-in, so each test hands in the FTS rank list the reference's
statement would have computed (here: empty or explicit) and the projection rows of what it saved."""
import uuid

import numpy as np
import pytest

from omnia_b200 import GPUMemoryStore, SCOPE_WORKSPACE_ID, SCOPE_VIRTUAL_USER_ID, SCOPE_AGENT_ID
from omnia_b200 import hybrid as H

pytestmark = pytest.mark.gpu
DIM = 1536
WS1 = "00000000-0000-0000-0000-000000000001"
USER = "hybrid-user"
AGENT = "00000000-0000-0000-0000-0000000000aa"
DAY = 86400.0


def one_hot(pos, n=DIM):
    out = np.zeros(n, np.float32)
    out[pos] = 1.0
    return out


class Fixture:
    """insertHybridMemory (retrieve_multi_tier_hybrid_test.go:48-84): one entity, one active observation."""

    def __init__(self):
        self.store = GPUMemoryStore(DIM, 1024)
        self.rows = {}

    def insert(self, user, agent, kind, confidence, emb, age_days=0.0, source_type="conversation_extraction"):
        oid, eid = uuid.uuid4(), uuid.uuid4()
        scope = {SCOPE_WORKSPACE_ID: WS1}
        if user: scope[SCOPE_VIRTUAL_USER_ID] = user
        if agent: scope[SCOPE_AGENT_ID] = agent
        self.store.upsert(oid, eid, scope, emb, kind=kind, confidence=confidence)
        self.rows[str(eid)] = H.EntityRow(str(eid), kind=kind, source_type=source_type, confidence=confidence,
                                          age_seconds=age_days * DAY, virtual_user_id=user or None,
                                          agent_id=agent or None)
        return str(eid)

    def close(self):
        self.store.close()


@pytest.fixture
def fx():
    f = Fixture()
    yield f
    f.close()


def test_semantic_only_match_across_tiers(fx):
    # retrieve_multi_tier_hybrid_test.go:87-104: the FTS arm matches nothing, cosine alone surfaces the row
    emb = one_hot(7)
    fx.insert("", "", "fact", 0.9, emb)
    res = H.retrieve_multi_tier_hybrid(fx.store, H.MultiTierRequest(workspace_id=WS1, agent_id=AGENT,
                                                                    query="refund window", limit=10), emb, [], fx.rows)
    assert res and res[0].tier == H.TIER_INSTITUTIONAL


def test_tier_ranker_reorders(fx):
    # :127-151
    emb = one_hot(3)
    fx.insert("", "", "fact", 0.9, emb)
    fx.insert(USER, AGENT, "fact", 0.9, emb)
    ranker = H.MultiplicativeTierRanker({H.TIER_INSTITUTIONAL: 0.1, H.TIER_USER: 1.0})
    res = H.retrieve_multi_tier_hybrid(fx.store, H.MultiTierRequest(workspace_id=WS1, user_id=USER, agent_id=AGENT,
                                                                    query="refund speed", limit=10, ranker=ranker),
                                       emb, [], fx.rows)
    assert len(res) == 2 and res[0].tier == H.TIER_USER_FOR_AGENT


def test_type_filter_reaches_the_cosine_arm(fx):
    # :155-174
    emb = one_hot(5)
    fx.insert("", "", "fact", 0.8, emb)
    fx.insert("", "", "preference", 0.8, emb)
    res = H.retrieve_multi_tier_hybrid(fx.store, H.MultiTierRequest(workspace_id=WS1, query="row", types=["fact"],
                                                                    limit=10), emb, [], fx.rows)
    assert res and all(m.row.kind == "fact" for m in res)


def test_per_tier_half_life_changes_ranking(fx):
    # :181-213
    emb = one_hot(9)
    fx.insert("", "", "fact", 0.9, emb, age_days=30)
    fx.insert(USER, "", "fact", 0.9, emb, age_days=30)
    base = dict(workspace_id=WS1, user_id=USER, query="topic", limit=10)
    inst = H.retrieve_multi_tier_hybrid(fx.store, H.MultiTierRequest(
        half_life=H.TierHalfLife(user=1.0, institutional=365 * DAY), **base), emb, [], fx.rows)
    assert len(inst) == 2 and inst[0].tier == H.TIER_INSTITUTIONAL
    usr = H.retrieve_multi_tier_hybrid(fx.store, H.MultiTierRequest(
        half_life=H.TierHalfLife(user=365 * DAY, institutional=1.0), **base), emb, [], fx.rows)
    assert len(usr) == 2 and usr[0].tier == H.TIER_USER


def test_requires_workspace_and_truncates_to_limit(fx):
    # :217-247
    with pytest.raises(H.RecallError, match="workspace_id"):
        H.retrieve_multi_tier_hybrid(fx.store, H.MultiTierRequest(query="q"), one_hot(0), [], fx.rows)
    emb = one_hot(2)
    for _ in range(4):
        fx.insert("", "", "fact", 0.8, emb)
    res = H.retrieve_multi_tier_hybrid(fx.store, H.MultiTierRequest(workspace_id=WS1, query="topic", limit=2),
                                       emb, [], fx.rows)
    assert len(res) == 2


def test_multi_tier_null_anchoring_excludes_other_users(fx):
    # retrieve_multi_tier.go:313-328: a request for USER sees institutional and USER rows, never another user's
    emb = one_hot(11)
    mine = fx.insert(USER, "", "fact", 0.9, emb)
    inst = fx.insert("", "", "fact", 0.9, emb)
    fx.insert("someone-else", "", "fact", 0.9, emb)
    fx.insert("", AGENT, "fact", 0.9, emb)        # agent-scoped: needs AgentID on the request
    res = H.retrieve_multi_tier_hybrid(fx.store, H.MultiTierRequest(workspace_id=WS1, user_id=USER, query="q"),
                                       emb, [], fx.rows)
    assert {m.entity_id for m in res} == {mine, inst}
    res = H.retrieve_multi_tier_hybrid(fx.store, H.MultiTierRequest(workspace_id=WS1, query="q"), emb, [], fx.rows)
    assert {m.entity_id for m in res} == {inst}


def test_single_tier_hybrid_fuses_fts_and_cosine(fx):
    # store.go:220-317 through RetrieveHybrid (store_read.go:351-406): an entity in both lists outranks
    # one that is only nearest by cosine; min_confidence applies inside the cosine arm
    near = fx.insert(USER, "", "fact", 0.9, one_hot(1), source_type="user_requested")
    both = fx.insert(USER, "", "fact", 0.9, (one_hot(1) + 0.2 * one_hot(2)).astype(np.float32),
                     source_type="user_requested")
    low = fx.insert(USER, "", "fact", 0.2, one_hot(1), source_type="user_requested")
    fx.insert("other-user", "", "fact", 0.9, one_hot(1), source_type="user_requested")
    scope = {SCOPE_WORKSPACE_ID: WS1, SCOPE_VIRTUAL_USER_ID: USER}
    res = H.retrieve_hybrid(fx.store, scope, "refund", one_hot(1), [both], fx.rows, limit=10, min_confidence=0.5)
    assert [m.entity_id for m in res] == [both, near]
    assert res[0].score == pytest.approx((1 / 61.0 + 1 / 62.0) * 0.9, rel=1e-12)
    assert res[1].score == pytest.approx((1 / 61.0) * 0.9, rel=1e-12)
    res = H.retrieve_hybrid(fx.store, scope, "refund", one_hot(1), [], fx.rows, limit=10)
    assert {m.entity_id for m in res} == {near, both, low}
    assert H.retrieve_hybrid(fx.store, scope, "", one_hot(1), [], fx.rows) is None
