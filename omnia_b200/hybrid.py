"""Host side of the hybrid recall paths: what happens to the cosine arm's rank list after the GPU.

SURVEY.md section 8 rows a5, a7 and a13.  In the reference one SQL statement computes an FTS rank list,
a cosine rank list (cosine_ann -> cosine -> cosine_ranked, which `GPUMemoryStore.cosine_ranked`
answers from HBM), fuses them by reciprocal rank, multiplies by quality weights and projects the
surviving entities; the multi-tier variant then applies a tier bias in Go.  The drop-in keeps FTS and
the projection in Postgres and splices the GPU's candidate list in, so this
module restates only the arithmetic that sits between the two -- it is what the `gpustore` Go package
does after `omr_cosine_ranked` returns, written in Python because the image has no Go toolchain.

Inputs Postgres still supplies: `fts_ranked` (entity ids in fts_rn order, at most `fanout` of them)
and `rows`, the projection's columns for each entity that survives the statement's own WHERE clause
(an entity missing from `rows` is dropped, exactly as the final JOIN drops it).
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field
from typing import Callable, Dict, Iterable, List, Mapping, Optional, Sequence

from . import _lib as L
from ._lib import RecallError
from .store import FilterSpec, GPUMemoryStore, SCOPE_WORKSPACE_ID

RRF_K = 60.0                     # retrieve_multi_tier_hybrid.go:41; store.go:279-280
HYBRID_FANOUT = 100              # retrieve_multi_tier_hybrid.go:37
DEFAULT_MEMORY_LIMIT = 50        # store.go:330
DEFAULT_MULTI_TIER_LIMIT = 15    # retrieve_multi_tier.go:122
DEFAULT_HALF_LIFE_S = 30 * 24 * 3600.0   # tier_ranking.go:18; store.go:306 (2592000.0)
SEMANTIC_OVER_FETCH_FACTOR = 4   # api/service_semantic.go:33
SEMANTIC_MAX_FETCH = 200         # api/service_semantic.go:36

TIER_INSTITUTIONAL = "institutional"     # retrieve_multi_tier.go:40-45
TIER_AGENT = "agent"
TIER_USER = "user"
TIER_USER_FOR_AGENT = "user_for_agent"

_SOURCE_TYPE_WEIGHT = {                  # store.go:170-178
    "user_requested": 1.0,
    "operator_curated": 1.0,
    "reflection": 0.85,
    "conversation_extraction": 0.7,
    "system_generated": 0.5,
}


def source_type_weight(source_type: Optional[str]) -> float:
    return _SOURCE_TYPE_WEIGHT.get(source_type or "", 0.7)


@dataclass
class EntityRow:
    """Projection columns of one entity: its newest active observation joined to memory_entities."""
    entity_id: str
    kind: str = ""
    source_type: str = ""
    confidence: Optional[float] = None       # o.confidence
    age_seconds: float = 0.0                 # EXTRACT(EPOCH FROM now() - o.observed_at)
    virtual_user_id: Optional[str] = None
    agent_id: Optional[str] = None
    metadata: Dict[str, object] = field(default_factory=dict)


@dataclass
class ScoredMemory:
    entity_id: str
    score: float
    row: EntityRow
    tier: str = ""


def single_tier_fanout(limit: int) -> int:
    """store_read.go:365-372: limit defaults to 50; fanout = max(5 * limit, 50)."""
    if limit <= 0:
        limit = DEFAULT_MEMORY_LIMIT
    return max(limit * 5, 50)


def rrf_fuse(fts_ranked: Sequence[str], cos_ranked: Sequence[str], k: float = RRF_K) -> Dict[str, float]:
    """`fused` CTE (store.go:272-282): coalesce(1/(k + fts_rn), 0) + coalesce(1/(k + cos_rn), 0) over the
    FULL OUTER JOIN of the two rank lists; ranks are 1-based row numbers."""
    fused: Dict[str, float] = {}
    for rn, e in enumerate(fts_ranked, 1):
        fused[e] = 1.0 / (k + rn)
    for rn, e in enumerate(cos_ranked, 1):
        fused[e] = fused.get(e, 0.0) + 1.0 / (k + rn)
    return fused


def retrieve_hybrid(store: GPUMemoryStore, scope: Mapping[str, str], query: str, query_embedding,
                    fts_ranked: Sequence[str], rows: Mapping[str, EntityRow], limit: int = 0,
                    min_confidence: float = 0.0, now_us: Optional[int] = None) -> Optional[List[ScoredMemory]]:
    """RetrieveHybrid (store_read.go:351-406) with the cosine arm on the GPU.

    Returns None when the reference falls through to the FTS-only Retrieve (empty query text or empty
    embedding, store_read.go:361-363): that statement has no cosine term and stays in Postgres."""
    if not scope.get(SCOPE_WORKSPACE_ID):
        raise RecallError(L.E_WORKSPACE, "memory: workspace_id is required")
    if not query or query_embedding is None or len(query_embedding) == 0:
        return None
    if limit <= 0:
        limit = DEFAULT_MEMORY_LIMIT
    fanout = single_tier_fanout(limit)
    flt = FilterSpec.from_scope(dict(scope), min_confidence=max(min_confidence, 0.0), now_us=now_us)
    cos = [e for e, _, _ in store.cosine_ranked(flt, query_embedding, fanout)]
    fused = rrf_fuse(list(fts_ranked)[:fanout], cos)
    out = []
    for e, rrf in fused.items():
        r = rows.get(e)
        if r is None:
            continue
        # store.go:298-306: rrf * source weight * o.confidence * exp(-age / 2592000)
        conf = r.confidence if r.confidence is not None else float("nan")
        score = rrf * source_type_weight(r.source_type) * conf * math.exp(-r.age_seconds / DEFAULT_HALF_LIFE_S)
        out.append(ScoredMemory(e, score, r))
    # ORDER BY final_score DESC LIMIT $7; SQL leaves ties open, the entity id settles them here
    out.sort(key=lambda m: (-m.score, m.entity_id))
    return out[:limit]


# ---------------------------------------------------------------- multi-tier

def classify_tier(virtual_user_id: Optional[str], agent_id: Optional[str]) -> str:
    """classifyTier (retrieve_multi_tier.go:437-450): NULL columns, not empty strings, decide."""
    if virtual_user_id is not None and agent_id is not None:
        return TIER_USER_FOR_AGENT
    if virtual_user_id is not None:
        return TIER_USER
    if agent_id is not None:
        return TIER_AGENT
    return TIER_INSTITUTIONAL


class IdentityTierRanker:
    """tier_ranking.go:66-70."""

    def adjust(self, base: float, tier: str) -> float:
        return base


class MultiplicativeTierRanker:
    """ee/pkg/memory/tier_ranking.go:19-38: base * weight[tier]; user_for_agent inherits the user weight;
    a tier with no weight keeps its base."""

    def __init__(self, weights: Mapping[str, float]):
        self.weights = dict(weights)

    def adjust(self, base: float, tier: str) -> float:
        w = self.weights.get(tier)
        if w is None and tier == TIER_USER_FOR_AGENT:
            w = self.weights.get(TIER_USER)
        return base if w is None else base * w


@dataclass
class TierHalfLife:
    """tier_ranking.go:24-56; seconds, zero or negative = unset (30 days)."""
    user: float = 0.0
    agent: float = 0.0
    institutional: float = 0.0

    def or_defaults(self) -> "TierHalfLife":
        d = DEFAULT_HALF_LIFE_S
        return TierHalfLife(self.user if self.user > 0 else d, self.agent if self.agent > 0 else d,
                            self.institutional if self.institutional > 0 else d)


@dataclass
class MultiTierRequest:
    """retrieve_multi_tier.go:57-118, the fields the hybrid path reads."""
    workspace_id: str = ""
    user_id: str = ""
    agent_id: str = ""
    query: str = ""
    types: Sequence[str] = ()
    purposes: Sequence[str] = ()
    min_confidence: float = 0.0
    limit: int = 0
    tiers: Sequence[str] = ()
    ranker: Optional[object] = None
    half_life: TierHalfLife = field(default_factory=TierHalfLife)


def multi_tier_filter(req: MultiTierRequest, now_us: Optional[int] = None) -> FilterSpec:
    """The shared `candidates` predicate (retrieve_multi_tier_hybrid.go:132-138; retrieve_multi_tier.go:313-328):
    an unset user or agent means IS NULL, a set one means IS NULL OR = x."""
    return FilterSpec(req.workspace_id,
                      req.user_id or None, L.MATCH_NULL_OR_EQ if req.user_id else L.MATCH_NULL_ONLY,
                      req.agent_id or None, L.MATCH_NULL_OR_EQ if req.agent_id else L.MATCH_NULL_ONLY,
                      kinds=list(req.types), purposes=list(req.purposes),
                      min_confidence=req.min_confidence if req.min_confidence > 0 else 0.0, now_us=now_us)


def rank_hybrid_results(results: List[ScoredMemory], ranker=None) -> None:
    """rankHybridResults (retrieve_multi_tier_hybrid.go:278-288): tier bias on the SQL score, then a
    stable descending sort."""
    ranker = ranker or IdentityTierRanker()
    for r in results:
        r.score = ranker.adjust(r.score, r.tier)
    results.sort(key=lambda m: -m.score)   # list.sort is stable, as sort.SliceStable is


def retrieve_multi_tier_hybrid(store: GPUMemoryStore, req: MultiTierRequest, query_embedding,
                               fts_ranked: Sequence[str], rows: Mapping[str, EntityRow],
                               now_us: Optional[int] = None) -> Optional[List[ScoredMemory]]:
    """RetrieveMultiTierHybrid (retrieve_multi_tier_hybrid.go:180-227) with the cosine arm on the GPU.
    Returns None where the reference falls through to the FTS-only RetrieveMultiTier (:181-183)."""
    if not req.query or query_embedding is None or len(query_embedding) == 0:
        if not req.workspace_id:
            raise RecallError(L.E_WORKSPACE, "memory: workspace_id is required")
        return None
    if not req.workspace_id:
        raise RecallError(L.E_WORKSPACE, "memory: workspace_id is required")
    hl = req.half_life.or_defaults()
    cos = [e for e, _, _ in store.cosine_ranked(multi_tier_filter(req, now_us), query_embedding, HYBRID_FANOUT)]
    fused = rrf_fuse(list(fts_ranked)[:HYBRID_FANOUT], cos)
    out = []
    # the statement returns one row per entity ordered by entity id (DISTINCT ON ... ORDER BY e.entity_id)
    for e in sorted(fused):
        r = rows.get(e)
        if r is None:
            continue
        tier = classify_tier(r.virtual_user_id, r.agent_id)
        # recencyExpr (:158-164): a user-bearing row decays with the user half-life, agent-only with the
        # agent one, the rest institutional; the exponent is floored at -700
        half = hl.user if r.virtual_user_id is not None else (hl.agent if r.agent_id is not None else hl.institutional)
        recency = math.exp(max(-700.0, -math.log(2.0) * r.age_seconds / half))
        conf = r.confidence if r.confidence is not None else 0.7        # coalesce(e.confidence, 0.7) (:116)
        out.append(ScoredMemory(e, fused[e] * source_type_weight(r.source_type) * conf * recency, r, tier))
    if req.tiers:                                                      # filterTiers, retrieve_multi_tier.go:454-469
        allow = set(req.tiers)
        out = [m for m in out if m.tier in allow]
    rank_hybrid_results(out, req.ranker)
    limit = req.limit if req.limit > 0 else DEFAULT_MULTI_TIER_LIMIT
    return out[:limit]


# ---------------------------------------------------------------- semantic retrieve

def over_fetch_limit(limit: int) -> int:
    """overFetchLimit (api/service_semantic.go:77-86)."""
    if limit <= 0:
        return SEMANTIC_MAX_FETCH
    return min(limit * SEMANTIC_OVER_FETCH_FACTOR, SEMANTIC_MAX_FETCH)


def retrieve_semantic(search: Callable[[int], Iterable[ScoredMemory]], allowed: Callable[[Dict[str, object]], bool],
                      limit: int) -> List[ScoredMemory]:
    """RetrieveSemantic (api/service_semantic.go:47-72): over-fetch, drop what the deny filter refuses,
    stop at `limit`.  `search(n)` is the hybrid search asked for n rows; `allowed` is the CEL deny filter."""
    out: List[ScoredMemory] = []
    for m in search(over_fetch_limit(limit)):
        if m is None or not allowed(m.row.metadata):
            continue
        out.append(m)
        if limit > 0 and len(out) >= limit:
            break
    return out
