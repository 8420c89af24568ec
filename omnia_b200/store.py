"""Host-side mirror of the reference's store interface for the recall path.

`GPUMemoryStore` carries the methods of Go's ``memory.Store`` that reach pgvector
(internal/memory/types.go:231-332) with the same names, argument meaning and error
behaviour, so the parity tests read like the reference's own tests:

    FindSimilarObservations   -> find_similar_observations   (store_read.go:243-306)
    UpdateEmbedding           -> update_embedding            (store_write.go:557-578)
    UpdateObservationEmbedding-> upsert                      (store_write.go:580-600)
    cosine arm of RetrieveHybrid -> cosine_ranked            (store.go:240-272)
    RunEntityDuplicateCandidates -> run_entity_duplicate_candidates
                                  (internal/memory/postgres/consolidation_prefilter.go:176-203)

Everything here is a thin layer over the C ABI (include/omnia_recall.h); the arithmetic runs
in libomnia_recall.so on the GPU and nowhere else.
"""
from __future__ import annotations

import ctypes as C
import uuid as _uuid
from dataclasses import dataclass
from typing import Dict, List, Optional, Sequence, Union

import numpy as np

from . import _lib as L

SCOPE_WORKSPACE_ID = "workspace_id"        # internal/memory/store.go:35
SCOPE_VIRTUAL_USER_ID = "virtual_user_id"  # internal/memory/store.go:41
SCOPE_AGENT_ID = "agent_id"                # internal/memory/store.go:47

DEFAULT_DUPLICATE_CANDIDATE_LIMIT = 5      # internal/memory/types.go:137-141
DEFAULT_AUTO_SUPERSEDE_SIMILARITY = 0.95
DEFAULT_SURFACE_DUPLICATE_SIMILARITY = 0.85

IdLike = Union[str, bytes, _uuid.UUID, int]


def id_bytes(v: IdLike) -> bytes:
    if isinstance(v, _uuid.UUID):
        return v.bytes
    if isinstance(v, bytes):
        if len(v) != 16:
            raise ValueError("uuid bytes must have length 16")
        return v
    if isinstance(v, int):
        return int(v).to_bytes(16, "big")
    return _uuid.UUID(str(v)).bytes


def id_str(b) -> str:
    return str(_uuid.UUID(bytes=bytes(b)))


@dataclass
class SimilarObservation:  # internal/memory/types.go:126-131 (Content stays in Postgres)
    observation_id: str
    entity_id: str
    similarity: float
    row: int


@dataclass
class DuplicatePair:
    """One bucket of RunEntityDuplicateCandidates: Key "dupe:<canonical>:<dupe>", Stats.similarity."""
    canonical_id: str
    dupe_id: str
    similarity: float
    kind: str
    canonical_row: int
    dupe_row: int

    @property
    def key(self) -> str:
        return f"dupe:{self.canonical_id}:{self.dupe_id}"


def _cstr(s: Optional[str]) -> Optional[bytes]:
    return None if s is None else s.encode()


def _str_array(items: Sequence[str]):
    arr = (C.c_char_p * max(1, len(items)))()
    for i, s in enumerate(items):
        arr[i] = s.encode()
    return arr


def _now_us(now_us) -> int:
    """The query's now() for `valid_until > now()`: the wall clock unless the caller pins it (tests do)."""
    import time
    return int(time.time() * 1e6) if now_us is None else int(now_us)


class FilterSpec:
    """Reference row a8 -- the WHERE clause of a recall query.  Keeps its C strings alive."""

    def __init__(self, workspace_id: str, virtual_user_id: Optional[str] = None, user_mode: Optional[int] = None,
                 agent_id: Optional[str] = None, agent_mode: Optional[int] = None, kinds: Sequence[str] = (),
                 purposes: Sequence[str] = (), min_confidence: float = 0.0, max_distance: float = float("nan"),
                 now_us: Optional[int] = None):
        # single-tier default: a nil scope value means "no constraint" (store.go:252-253)
        if user_mode is None:
            user_mode = L.MATCH_EQ if virtual_user_id is not None else L.MATCH_ANY
        if agent_mode is None:
            agent_mode = L.MATCH_EQ if agent_id is not None else L.MATCH_ANY
        self._kinds = _str_array(kinds)
        self._purposes = _str_array(purposes)
        self.c = L.Filter(_cstr(workspace_id), _cstr(virtual_user_id), user_mode, _cstr(agent_id), agent_mode,
                          C.cast(self._kinds, L.c_str_p), len(kinds), C.cast(self._purposes, L.c_str_p),
                          len(purposes), float(min_confidence), float(max_distance), _now_us(now_us))

    @staticmethod
    def from_scope(scope: Dict[str, str], **kw) -> "FilterSpec":
        return FilterSpec(scope.get(SCOPE_WORKSPACE_ID, ""), scope.get(SCOPE_VIRTUAL_USER_ID) or None, None,
                          scope.get(SCOPE_AGENT_ID) or None, None, **kw)


def _filters_array(filters: Sequence[FilterSpec]):
    arr = (L.Filter * len(filters))()
    for i, f in enumerate(filters):
        arr[i] = f.c
    return arr


class GPUMemoryStore:
    def __init__(self, dim: int, capacity: int, device: int = 0, row_base: int = 0, row_stride: int = 1,
                 candidate_margin: float = 0.0, sub_bucket: int = 0, rescore_width: int = 0, force_path: int = 0):
        self._lib = L.load()
        self.dim = int(dim)
        self.capacity = int(capacity)
        self._h = C.c_void_p()
        cfg = L.Config(self.dim, int(device), self.capacity, int(row_base), int(row_stride), float(candidate_margin),
                       int(sub_bucket), int(rescore_width), int(force_path))
        L.check(self._lib.omr_open(C.byref(cfg), C.byref(self._h)))

    # -- lifecycle -------------------------------------------------------------
    def close(self) -> None:
        if getattr(self, "_h", None) is not None and self._h:
            self._lib.omr_close(self._h)
            self._h = C.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def handle(self) -> C.c_void_p:
        return self._h

    def __len__(self) -> int:
        return int(self._lib.omr_count(self._h))

    def stats(self) -> L.Stats:
        st = L.Stats()
        L.check(self._lib.omr_get_stats(self._h, C.byref(st)))
        return st

    def set_profiling(self, on: bool) -> None:
        L.check(self._lib.omr_set_profiling(self._h, int(on)))

    def bitmap_words(self) -> int:
        return int(self._lib.omr_bitmap_words(self._h))

    # -- ingest ----------------------------------------------------------------
    def bulk_load(self, vectors: np.ndarray, *, observation_ids: Optional[np.ndarray] = None,
                  entity_ids: Optional[np.ndarray] = None, workspaces: Sequence[str] = ("default",),
                  workspace_idx: Optional[np.ndarray] = None, users: Sequence[str] = (),
                  user_idx: Optional[np.ndarray] = None, agents: Sequence[str] = (),
                  agent_idx: Optional[np.ndarray] = None, kinds: Sequence[str] = (),
                  kind_idx: Optional[np.ndarray] = None, purposes: Sequence[str] = (),
                  purpose_idx: Optional[np.ndarray] = None, confidence: Optional[np.ndarray] = None,
                  flags: Optional[np.ndarray] = None, valid_until_us: Optional[np.ndarray] = None) -> None:
        v = np.ascontiguousarray(vectors, np.float32)
        if v.ndim != 2 or v.shape[1] != self.dim:
            raise L.RecallError(L.E_DIM, f"memory: embedding dimension mismatch: got {v.shape[-1]}, store has {self.dim}")
        n = v.shape[0]
        keep = [v]

        def arr(a, dt, width=None):
            if a is None:
                return None
            a = np.ascontiguousarray(a, dt)
            if a.shape[0] != n or (width and a.reshape(n, -1).shape[1] != width):
                raise ValueError("bulk_load: array length mismatch")
            keep.append(a)
            return a.ctypes.data

        tabs = [_str_array(t) for t in (workspaces, users, agents, kinds, purposes)]
        b = L.Bulk(n, v.ctypes.data, arr(observation_ids, np.uint8, 16), arr(entity_ids, np.uint8, 16),
                   C.cast(tabs[0], L.c_str_p), len(workspaces), arr(workspace_idx, np.uint32),
                   C.cast(tabs[1], L.c_str_p), len(users), arr(user_idx, np.int32),
                   C.cast(tabs[2], L.c_str_p), len(agents), arr(agent_idx, np.int32),
                   C.cast(tabs[3], L.c_str_p), len(kinds), arr(kind_idx, np.uint16),
                   C.cast(tabs[4], L.c_str_p), len(purposes), arr(purpose_idx, np.uint16),
                   arr(confidence, np.float32), arr(flags, np.uint8), arr(valid_until_us, np.int64))
        L.check(self._lib.omr_bulk_load(self._h, C.byref(b)))

    def upsert(self, observation_id: IdLike, entity_id: IdLike, scope: Dict[str, str], embedding, *,
               kind: str = "", purpose: str = "", confidence: float = 1.0,
               flags: int = L.ROW_HAS_EMBEDDING | L.ROW_MUTABLE, valid_until_us: int = L.NO_EXPIRY) -> None:
        """UpdateObservationEmbedding (store_write.go:580-600) plus the metadata the filter reads."""
        v = np.ascontiguousarray(embedding, np.float32).reshape(-1)
        r = L.Row()
        C.memmove(r.observation_id, id_bytes(observation_id), 16)
        C.memmove(r.entity_id, id_bytes(entity_id), 16)
        r.workspace_id = _cstr(scope.get(SCOPE_WORKSPACE_ID, ""))
        r.virtual_user_id = _cstr(scope.get(SCOPE_VIRTUAL_USER_ID) or None)
        r.agent_id = _cstr(scope.get(SCOPE_AGENT_ID) or None)
        r.kind, r.purpose = kind.encode(), purpose.encode()
        r.confidence, r.flags, r.valid_until_us = float(confidence), int(flags), int(valid_until_us)
        L.check(self._lib.omr_upsert(self._h, C.byref(r), v.ctypes.data, v.shape[0]))

    def upsert_batch(self, observation_ids, entity_ids, scope: Dict[str, str], embeddings, *, kind: str = "",
                     purpose: str = "", confidence: float = 1.0, flags: int = L.ROW_HAS_EMBEDDING | L.ROW_MUTABLE,
                     valid_until_us: int = L.NO_EXPIRY) -> None:
        """n upserts under one scope in one call (omr_upsert_batch): one staged copy, one ingest launch."""
        v = np.ascontiguousarray(embeddings, np.float32)
        n = v.shape[0]
        rows = (L.Row * max(1, n))()
        ws = _cstr(scope.get(SCOPE_WORKSPACE_ID, ""))
        user, agent = _cstr(scope.get(SCOPE_VIRTUAL_USER_ID) or None), _cstr(scope.get(SCOPE_AGENT_ID) or None)
        for i in range(n):
            r = rows[i]
            C.memmove(r.observation_id, id_bytes(observation_ids[i]), 16)
            C.memmove(r.entity_id, id_bytes(entity_ids[i]), 16)
            r.workspace_id, r.virtual_user_id, r.agent_id = ws, user, agent
            r.kind, r.purpose = kind.encode(), purpose.encode()
            r.confidence, r.flags, r.valid_until_us = float(confidence), int(flags), int(valid_until_us)
        L.check(self._lib.omr_upsert_batch(self._h, rows, v.ctypes.data, n, v.shape[1] if n else self.dim))

    def compact(self) -> int:
        """Squeezes hard-deleted rows out of the shard (omr_compact); returns how many rows were reclaimed."""
        got = C.c_int64(0)
        L.check(self._lib.omr_compact(self._h, C.byref(got)))
        return int(got.value)

    def free_slots(self) -> int:
        return int(self._lib.omr_free_slots(self._h))

    def update_embedding(self, entity_id: IdLike, embedding) -> None:
        """UpdateEmbedding (store_write.go:557-578)."""
        v = np.ascontiguousarray(embedding, np.float32).reshape(-1)
        L.check(self._lib.omr_update_embedding(self._h, id_bytes(entity_id), v.ctypes.data, v.shape[0]))

    def set_row_flags(self, observation_id: IdLike, set_mask: int = 0, clear_mask: int = 0) -> None:
        L.check(self._lib.omr_set_row_flags(self._h, id_bytes(observation_id), set_mask, clear_mask))

    def forget(self, entity_id: IdLike) -> None:
        """Soft delete: forgotten = true (store_delete.go:30-57)."""
        L.check(self._lib.omr_set_entity_flags(self._h, id_bytes(entity_id), L.ROW_FORGOTTEN, 0))

    def supersede(self, observation_id: IdLike) -> None:
        L.check(self._lib.omr_set_row_flags(self._h, id_bytes(observation_id), L.ROW_SUPERSEDED, 0))

    def set_valid_until(self, observation_id: IdLike, valid_until_us: int) -> None:
        L.check(self._lib.omr_set_valid_until(self._h, id_bytes(observation_id), int(valid_until_us)))

    def set_confidence(self, observation_id: IdLike, confidence: float) -> None:
        L.check(self._lib.omr_set_confidence(self._h, id_bytes(observation_id), float(confidence)))

    def delete_entity(self, entity_id: IdLike) -> None:
        L.check(self._lib.omr_delete_entity(self._h, id_bytes(entity_id)))

    def delete_observation(self, observation_id: IdLike) -> None:
        """Hard delete of one observation (tombstone.go:124, supersession_store.go:43)."""
        L.check(self._lib.omr_delete_observation(self._h, id_bytes(observation_id)))

    def delete_all(self, scope: Dict[str, str]) -> int:
        """DeleteAll (store_delete.go:60-87): every entity of the workspace, or of one user in it."""
        return self.batch_delete(scope, 0)

    def batch_delete(self, scope: Dict[str, str], limit: int) -> int:
        """BatchDelete (store_delete.go:89-125): at most `limit` entities; returns the count."""
        n = C.c_int64(0)
        L.check(self._lib.omr_delete_scope(self._h, _cstr(scope.get(SCOPE_WORKSPACE_ID, "")),
                                           _cstr(scope.get(SCOPE_VIRTUAL_USER_ID) or None), int(limit), C.byref(n)))
        return int(n.value)

    def rescope(self, observation_ids: Sequence[IdLike], user_id: str = "", agent_id: str = "") -> None:
        """ConsolidationWriter.Rescope (postgres/consolidation_writes.go:75-92)."""
        buf = b"".join(id_bytes(o) for o in observation_ids)
        L.check(self._lib.omr_rescope(self._h, buf, len(observation_ids), _cstr(user_id or None),
                                      _cstr(agent_id or None)))

    def merge_entities(self, canonical_id: IdLike, merge_ids: Sequence[IdLike]) -> None:
        """ConsolidationWriter.MergeEntities (postgres/consolidation_writes.go:104-127)."""
        buf = b"".join(id_bytes(e) for e in merge_ids)
        L.check(self._lib.omr_merge_entities(self._h, id_bytes(canonical_id), buf, len(merge_ids)))

    def clear(self) -> None:
        """Drops every row (re-embed flush; a dimension change is close + reopen)."""
        L.check(self._lib.omr_clear(self._h))

    # -- recall ----------------------------------------------------------------
    def search(self, queries, filters: Sequence[FilterSpec], k: int, query_filter: Optional[np.ndarray] = None):
        """Batched `ORDER BY embedding <=> q LIMIT k`.  Returns (rows [nq,k] int64 (-1 pad),
        distances [nq,k] float64 (inf pad), counts [nq], hits ndarray of L.Hit)."""
        q = np.ascontiguousarray(queries, np.float32)
        if q.ndim == 1:
            q = q[None, :]
        nq = q.shape[0]
        kk = k if k > 0 else DEFAULT_DUPLICATE_CANDIDATE_LIMIT
        hits = (L.Hit * (nq * kk))()
        counts = np.zeros(nq, np.int32)
        qf = None if query_filter is None else np.ascontiguousarray(query_filter, np.int32)
        fa = _filters_array(filters)
        L.check(self._lib.omr_search(self._h, q.ctypes.data, nq, q.shape[1], fa, len(filters),
                                     None if qf is None else qf.ctypes.data, k, hits, counts.ctypes.data))
        rows = np.full((nq, kk), -1, np.int64)
        dist = np.full((nq, kk), np.inf, np.float64)
        for i in range(nq):
            for j in range(int(counts[i])):
                h = hits[i * kk + j]
                rows[i, j], dist[i, j] = h.row, h.distance
        return rows, dist, counts, hits

    def search_rows(self, queries, filters: Sequence[FilterSpec], k: int, query_filter: Optional[np.ndarray] = None,
                    out=None):
        """omr_search_rows: the same recall with array results only (rows, distances, counts)."""
        q = np.ascontiguousarray(queries, np.float32)
        if q.ndim == 1:
            q = q[None, :]
        nq = q.shape[0]
        kk = k if k > 0 else DEFAULT_DUPLICATE_CANDIDATE_LIMIT
        rows, dist, counts = out if out is not None else (np.empty((nq, kk), np.int64), np.empty((nq, kk), np.float64),
                                                          np.empty(nq, np.int32))
        qf = None if query_filter is None else np.ascontiguousarray(query_filter, np.int32)
        fa = _filters_array(filters)
        L.check(self._lib.omr_search_rows(self._h, q.ctypes.data, nq, q.shape[1], fa, len(filters),
                                          None if qf is None else qf.ctypes.data, k, rows.ctypes.data,
                                          dist.ctypes.data, counts.ctypes.data))
        return rows, dist, counts

    def search_rows_begin(self, queries, filters: Sequence[FilterSpec], k: int,
                          query_filter: Optional[np.ndarray] = None) -> int:
        """omr_search_rows_begin: stage a host batch and return its ticket (at most four in flight)."""
        q = np.ascontiguousarray(queries, np.float32)
        if q.ndim == 1:
            q = q[None, :]
        qf = None if query_filter is None else np.ascontiguousarray(query_filter, np.int32)
        fa = _filters_array(filters)
        ticket = C.c_int32(0)
        L.check(self._lib.omr_search_rows_begin(self._h, q.ctypes.data, q.shape[0], q.shape[1], fa, len(filters),
                                                None if qf is None else qf.ctypes.data, k, C.byref(ticket)))
        return ticket.value

    def search_rows_finish(self, ticket: int, out):
        rows, dist, counts = out
        L.check(self._lib.omr_search_rows_finish(self._h, ticket, rows.ctypes.data, dist.ctypes.data,
                                                 counts.ctypes.data))
        return rows, dist, counts

    def find_similar_observations(self, scope: Dict[str, str], query_embedding, k: int,
                                  min_similarity: float, now_us: Optional[int] = None) -> List[SimilarObservation]:
        """FindSimilarObservations (store_read.go:243-306): most similar first, similarity >= min."""
        if not scope.get(SCOPE_WORKSPACE_ID):
            raise L.RecallError(L.E_WORKSPACE, "memory: workspace_id is required")
        kk = k if k > 0 else DEFAULT_DUPLICATE_CANDIDATE_LIMIT
        f = FilterSpec.from_scope(scope, max_distance=1.0 - float(min_similarity), now_us=now_us)
        _, _, counts, hits = self.search(np.asarray(query_embedding, np.float32)[None, :], [f], kk)
        return [SimilarObservation(id_str(hits[j].observation_id), id_str(hits[j].entity_id),
                                   1.0 - hits[j].distance, hits[j].row) for j in range(int(counts[0]))]

    def pre_write_dedup(self, scope: Dict[str, str], embedding, *,
                        auto_supersede_above: float = DEFAULT_AUTO_SUPERSEDE_SIMILARITY,
                        surface_duplicates_above: float = DEFAULT_SURFACE_DUPLICATE_SIMILARITY,
                        candidate_limit: int = DEFAULT_DUPLICATE_CANDIDATE_LIMIT, now_us: Optional[int] = None):
        """The embedding-similarity branch of SaveMemoryWithResult (internal/memory/api/service_write.go:63-104,
        213-228): FindSimilarObservations(scope, vec, candidateLimit, surfaceDuplicatesAbove); a top match at or
        above autoSupersedeAbove routes to AppendObservationToEntity (auto-supersede), anything else lands
        normally with the matches surfaced as PotentialDuplicates.
        Returns ("auto_supersede", top_match, matches) or ("insert", None, matches)."""
        matches = self.find_similar_observations(scope, embedding, candidate_limit, surface_duplicates_above, now_us)
        if matches and matches[0].similarity >= auto_supersede_above:
            return "auto_supersede", matches[0], matches
        return "insert", None, matches

    def cosine_ranked(self, flt: FilterSpec, query_embedding, fanout: int):
        """cosine_ann -> cosine -> cosine_ranked (store.go:240-272): [(entity_id, cos_dist, cos_rn)]."""
        q = np.ascontiguousarray(query_embedding, np.float32).reshape(-1)
        hits = (L.Hit * max(1, fanout))()
        cnt = C.c_int32(0)
        L.check(self._lib.omr_cosine_ranked(self._h, q.ctypes.data, q.shape[0], C.byref(flt.c), fanout, hits,
                                            C.byref(cnt)))
        return [(id_str(hits[i].entity_id), hits[i].distance, i + 1) for i in range(cnt.value)]

    # -- consolidation -----------------------------------------------------------
    def _join_opts(self, workspace_id, similarity_floor, limit, require_same_kind, require_mutable, part_index,
                   part_count, cancel):
        return L.SelfJoinOpts(_cstr(workspace_id), float(similarity_floor), int(limit), int(require_same_kind),
                              int(require_mutable), int(part_index), int(part_count),
                              C.pointer(cancel) if cancel is not None else None)

    def run_entity_duplicate_candidates(self, workspace_id: str, similarity_floor: float = 0.85,
                                        max_buckets_per_pass: int = 100, require_same_kind: bool = True,
                                        require_mutable: bool = True, part_index: int = 0,
                                        part_count: int = 1, cancel=None) -> List[DuplicatePair]:
        """RunEntityDuplicateCandidates (consolidation_prefilter.go:176-203; defaults worker.go:432-451).
        part_index / part_count: scan only that block partition of the pair space (multi-GPU, see dist.py).
        cancel: a ctypes.c_int32 another thread may set to 1 (the Go binding's ctx.Done watcher); the join then
        stops at its next chunk boundary and RecallError(E_CANCELLED) is raised."""
        opts = self._join_opts(workspace_id, similarity_floor, max_buckets_per_pass, require_same_kind, require_mutable,
                               part_index, part_count, cancel)
        cap = max_buckets_per_pass if max_buckets_per_pass > 0 else 1 << 16
        while True:
            out = (L.Pair * max(1, cap))()
            found = C.c_int64(0)
            L.check(self._lib.omr_selfjoin(self._h, C.byref(opts), out, cap, C.byref(found)))
            if found.value <= cap:
                break
            cap = int(found.value)
        return [DuplicatePair(id_str(p.canonical_id), id_str(p.dupe_id), p.similarity,
                              (self._lib.omr_kind_name(self._h, p.kind) or b"").decode(), p.canonical_row, p.dupe_row)
                for p in out[:found.value]]

    def selfjoin_rows(self, workspace_id: str, similarity_floor: float = 0.85, limit: int = 100,
                      require_same_kind: bool = True, require_mutable: bool = True, part_index: int = 0,
                      part_count: int = 1, cancel=None):
        """The same join as struct-of-arrays (canonical_row, dupe_row, similarity), best first, no uuid lookup
        (omr_selfjoin_rows): what the ranks of a partitioned join exchange."""
        opts = self._join_opts(workspace_id, similarity_floor, limit, require_same_kind, require_mutable, part_index,
                               part_count, cancel)
        cap = limit if limit > 0 else 1 << 16
        while True:
            a, b, s = np.empty(max(1, cap), np.int64), np.empty(max(1, cap), np.int64), np.empty(max(1, cap), np.float64)
            found = C.c_int64(0)
            L.check(self._lib.omr_selfjoin_rows(self._h, C.byref(opts), a.ctypes.data, b.ctypes.data, s.ctypes.data, cap,
                                                C.byref(found)))
            if found.value <= cap:
                break
            cap = int(found.value)
        n = int(found.value)
        return a[:n], b[:n], s[:n]

    def pairs_from_rows(self, canonical_rows, dupe_rows, similarities) -> List[DuplicatePair]:
        """Rows of a (gathered) join result -> the Bucket records of RunEntityDuplicateCandidates, through this
        store's identity maps (omr_row_ids)."""
        out = []
        for c, d, v in zip(canonical_rows, dupe_rows, similarities):
            ent_c = (C.c_uint8 * 16)(); ent_d = (C.c_uint8 * 16)(); kind = C.c_uint32(0)
            L.check(self._lib.omr_row_entity(self._h, int(c), ent_c, C.byref(kind)))
            L.check(self._lib.omr_row_entity(self._h, int(d), ent_d, None))
            out.append(DuplicatePair(id_str(ent_c), id_str(ent_d), float(v),
                                     (self._lib.omr_kind_name(self._h, kind.value) or b"").decode(), int(c), int(d)))
        return out

    # -- Memory Galaxy projection input ------------------------
    def projection_rows(self, flt: FilterSpec, cap: int = 0) -> np.ndarray:
        """One row per live entity in scope, newest eligible observation (LoadProjectionInputs,
        internal/memory/projection_store.go:33-62); `cap` > 0 truncates the list."""
        found = C.c_int64(0)
        L.check(self._lib.omr_projection_rows(self._h, C.byref(flt.c), None, 0, C.byref(found)))
        n = found.value if cap <= 0 else min(cap, found.value)
        rows = np.empty(max(n, 1), np.int64)
        L.check(self._lib.omr_projection_rows(self._h, C.byref(flt.c), rows.ctypes.data, n, C.byref(found)))
        return rows[:n]

    def pca_moments(self, rows):
        """Column means and sample covariance (float64) of the listed rows, computed on the GPU."""
        r = np.ascontiguousarray(rows, np.int64)
        mean = np.empty(self.dim, np.float64)
        cov = np.empty((self.dim, self.dim), np.float64)
        L.check(self._lib.omr_pca_moments(self._h, r.ctypes.data, r.shape[0], mean.ctypes.data, cov.ctypes.data))
        return mean, cov

    def pca_project(self, rows, basis) -> np.ndarray:
        """X @ basis in float64 over the listed (uncentred) rows; basis is [dim, k]."""
        r = np.ascontiguousarray(rows, np.int64)
        b = np.ascontiguousarray(basis, np.float64)
        out = np.empty((r.shape[0], b.shape[1]), np.float64)
        L.check(self._lib.omr_pca_project(self._h, r.ctypes.data, r.shape[0], b.ctypes.data, b.shape[1],
                                          out.ctypes.data))
        return out

    def pca_reduce(self, rows, k: int) -> np.ndarray:
        """PCAReduce (ee/pkg/memory/projection/pca.go:28-67) over resident rows.

        k is clamped to min(rows, dim) and to at least 1; a single row is truncated to its first k
        features; an empty selection is an error.  The moments and the projection run on the GPU, the
        dim x dim symmetric eigenproblem in between on the host (LAPACK here, gonum EigenSym in Go).
        Components are ordered by decreasing variance; their signs are the eigensolver's, as the
        reference's are its SVD's."""
        r = np.ascontiguousarray(rows, np.int64)
        n = r.shape[0]
        if n == 0 or self.dim == 0:
            raise L.RecallError(L.E_INVALID, f"projection: empty input matrix ({n}x{self.dim})")
        k = max(1, min(k, self.dim, n))
        if n < 2:
            basis = np.zeros((self.dim, k), np.float64)      # firstKCols (pca.go:70-83)
            basis[np.arange(k), np.arange(k)] = 1.0
            return self.pca_project(r, basis)
        _, cov = self.pca_moments(r)
        w, v = np.linalg.eigh(cov)
        order = np.argsort(-w, kind="stable")[:k]
        return self.pca_project(r, v[:, order])

    def debug_candidate_keys(self, queries, path: int, S: int) -> np.ndarray:
        q = np.ascontiguousarray(queries, np.float32)
        nq = q.shape[0]
        n_tiles = (len(self) + 255) // 256
        out = np.empty((nq, n_tiles * (256 // S) * 2), np.int32)
        L.check(self._lib.omr_debug_candidate_keys(self._h, q.ctypes.data, nq, path, S, out.ctypes.data, out.size))
        return out

    # -- device-pointer path (torch tensors are only carriers of device memory) -------
    def device_build_bitmaps(self, filters: Sequence[FilterSpec], out_tensor, stream: int = 0) -> None:
        fa = _filters_array(filters)
        L.check(self._lib.omr_device_build_bitmaps(self._h, fa, len(filters), out_tensor.data_ptr(), stream))

    def device_search(self, d_queries, k: int, d_rows, d_dist, d_count, d_bitmaps=None, n_groups: int = 0,
                      d_query_group=None, d_max_distance=None, stream: int = 0) -> None:
        nq = d_queries.shape[0]
        p = lambda t: None if t is None else t.data_ptr()
        L.check(self._lib.omr_device_search(self._h, d_queries.data_ptr(), nq, p(d_bitmaps), n_groups, p(d_query_group),
                                            p(d_max_distance), k, d_rows.data_ptr(), d_dist.data_ptr(),
                                            d_count.data_ptr(), stream))

    def device_search_begin(self, d_queries, k: int, d_rows, d_dist, d_count, d_bitmaps=None, n_groups: int = 0,
                            d_query_group=None, d_max_distance=None, stream: int = 0) -> int:
        """omr_device_search_begin: enqueue a batch, return its ticket at once (at most two in flight)."""
        nq = d_queries.shape[0]
        p = lambda t: None if t is None else t.data_ptr()
        ticket = C.c_int32(0)
        L.check(self._lib.omr_device_search_begin(self._h, d_queries.data_ptr(), nq, p(d_bitmaps), n_groups,
                                                  p(d_query_group), p(d_max_distance), k, d_rows.data_ptr(),
                                                  d_dist.data_ptr(), d_count.data_ptr(), stream, C.byref(ticket)))
        return ticket.value

    def device_search_finish(self, ticket: int) -> None:
        L.check(self._lib.omr_device_search_finish(self._h, ticket))

    def device_merge_topk(self, d_in_rows, d_in_dist, n_shards: int, nq: int, k: int, d_rows, d_dist, d_count,
                          stream: int = 0, shard_stride: int = 0) -> None:
        L.check(self._lib.omr_device_merge_topk(self._h, d_in_rows.data_ptr(), d_in_dist.data_ptr(), n_shards, nq, k,
                                                shard_stride, d_rows.data_ptr(), d_dist.data_ptr(),
                                                d_count.data_ptr(), stream))


def exact_topk_noalloc(d_rows, d_query, k: int, d_out_rows, d_out_dist, d_out_count, stream: int = 0) -> None:
    """omr_exact_topk_noalloc: exact top-k of one query over a caller-owned fp32 device matrix with no
    device allocation (for hosts that cannot call cudaMalloc while serving).  Arguments are device
    tensors the caller already owns; at most NOALLOC_MAX_ROWS rows and NOALLOC_MAX_K results, more is
    refused with RecallError.  Calls on one GPU share a fixed scratch array: do not overlap them."""
    n, pitch = d_rows.shape[0], d_rows.stride(0)
    L.check(L.load().omr_exact_topk_noalloc(d_rows.data_ptr(), n, d_query.shape[-1], pitch, d_query.data_ptr(), k,
                                            d_out_rows.data_ptr(), d_out_dist.data_ptr(), d_out_count.data_ptr(),
                                            stream))
