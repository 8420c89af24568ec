"""omnia_b200 -- B200-native semantic recall for the Omnia agent memory store.

One hot path only: the pgvector `<=>` scan behind internal/memory (recall, dedup-on-write)
and the consolidation near-duplicate self-join, rebuilt as hand-written sm_100a kernels behind
a C ABI (include/omnia_recall.h).  Importing the package needs no GPU; opening a store does.
"""
from ._lib import (MATCH_ANY, MATCH_EQ, MATCH_NULL_ONLY, MATCH_NULL_OR_EQ, NO_EXPIRY, ROW_FORGOTTEN,
                   ROW_HAS_EMBEDDING, ROW_MUTABLE, ROW_SUPERSEDED, NotFound, RecallError)
from .store import (SCOPE_AGENT_ID, SCOPE_VIRTUAL_USER_ID, SCOPE_WORKSPACE_ID, DuplicatePair, FilterSpec,
                    GPUMemoryStore, SimilarObservation)

__all__ = ["GPUMemoryStore", "FilterSpec", "SimilarObservation", "DuplicatePair", "RecallError", "NotFound",
           "SCOPE_WORKSPACE_ID", "SCOPE_VIRTUAL_USER_ID", "SCOPE_AGENT_ID", "MATCH_ANY", "MATCH_EQ",
           "MATCH_NULL_ONLY", "MATCH_NULL_OR_EQ", "NO_EXPIRY", "ROW_FORGOTTEN", "ROW_SUPERSEDED",
           "ROW_HAS_EMBEDDING", "ROW_MUTABLE"]
