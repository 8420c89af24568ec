"""CPU oracle for the recall hot path -- TEST INFRASTRUCTURE ONLY.

Only ``tests/``, ``__graft_entry__.smoke()`` and the ``cpu_baseline`` /
``--impl reference`` legs of ``bench.py`` may import this module.  The product
package ``omnia_b200`` never does; it fails loudly without its CUDA library.

The arithmetic is in ``oracle.c`` (plain C, built by ``oracle/Makefile``); this
file is the ctypes face plus the few pieces that are easier to state in numpy:

* ``cosine_ranked``  -- the cosine arm of the hybrid query, reference
  ``internal/memory/store.go:240-272`` (cosine_ann -> cosine -> cosine_ranked).
* ``numpy_topk``     -- BLAS-backed sanity variant.

Parity status: ordering / threshold / scope semantics pinned by the reference's
own tests (see tests/test_oracle_kat.py); numeric scores and self-join pair
sets are "parity unpinned" because the reference has no test that checks them.
"""
from __future__ import annotations

import ctypes as C
import hashlib
import os
import subprocess
from dataclasses import dataclass, field
from pathlib import Path
from typing import Optional, Sequence

import numpy as np

_HERE = Path(__file__).resolve().parent
_BUILD = _HERE / "_build"
_LIB = _BUILD / "liboracle.so"

USER_ANY, USER_NULL, USER_EQ, USER_NULL_OR_EQ = 0, 1, 2, 3
FLAG_FORGOTTEN, FLAG_SUPERSEDED, FLAG_HAS_EMBEDDING, FLAG_MUTABLE = 1, 2, 4, 8
NO_EXPIRY = np.iinfo(np.int64).max

VARIANT_F32_SEQ, VARIANT_F64, VARIANT_F32_VEC = 0, 1, 2


def _cpu_signature() -> str:
    """-march=native objects must not travel between different CPUs."""
    try:
        txt = Path("/proc/cpuinfo").read_text()
        keep = [ln for ln in txt.splitlines() if ln.startswith(("model name", "flags"))][:2]
        return hashlib.sha1("\n".join(keep).encode()).hexdigest()
    except OSError:
        return "unknown"


def build(force: bool = False) -> Path:
    """Builds liboracle.so (once per CPU model).  Several test processes may get here together (the multi-process
    tests on a box that has not built it yet), so the whole step holds a file lock."""
    import fcntl
    stamp = _BUILD / "cpu.sig"
    sig = _cpu_signature()
    with open(_HERE / ".build.lock", "w") as lock:
        fcntl.flock(lock, fcntl.LOCK_EX)
        try:
            if force or not _LIB.exists() or not stamp.exists() or stamp.read_text() != sig:
                subprocess.run(["make", "-C", str(_HERE), "clean"], check=True, capture_output=True)
            subprocess.run(["make", "-C", str(_HERE)], check=True, capture_output=True)
            _BUILD.mkdir(exist_ok=True)
            stamp.write_text(sig)
        finally:
            fcntl.flock(lock, fcntl.LOCK_UN)
    return _LIB


_lib = None


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(str(_LIB))
        L.omo_cosine_distance_f32.restype = C.c_double
        L.omo_cosine_distance_f64.restype = C.c_double
        L.omo_cosine_distance_vec.restype = C.c_double
        for f in (L.omo_cosine_distance_f32, L.omo_cosine_distance_f64, L.omo_cosine_distance_vec):
            f.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
        L.omo_search_batch.restype = None
        L.omo_search_batch.argtypes = [
            C.c_void_p, C.c_int64, C.c_int, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int,
            C.c_int, C.c_double, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
        L.omo_eligible.restype = None
        L.omo_eligible.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p]
        L.omo_selfjoin.restype = C.c_int64
        L.omo_selfjoin.argtypes = [
            C.c_void_p, C.c_int64, C.c_int, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
            C.c_float, C.c_int, C.c_int, C.c_int, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p]
        L.omo_max_threads.restype = C.c_int
        _lib = L
    return _lib


def max_threads() -> int:
    return int(lib().omo_max_threads())


def _f32(a) -> np.ndarray:
    return np.ascontiguousarray(a, dtype=np.float32)


def cosine_distance(a, b, variant: int = VARIANT_F32_SEQ) -> float:
    a, b = _f32(a), _f32(b)
    if a.shape != b.shape or a.ndim != 1:
        # pgvector raises "different vector dimensions"; the reference test
        # hnsw_ef_search_test.go:126-144 requires that to surface as an error.
        raise ValueError(f"different vector dimensions {a.shape[-1]} and {b.shape[-1]}")
    fn = {0: lib().omo_cosine_distance_f32, 1: lib().omo_cosine_distance_f64,
          2: lib().omo_cosine_distance_vec}[variant]
    return float(fn(a.ctypes.data, b.ctypes.data, a.shape[0]))


# ---------------------------------------------------------------- row metadata

class _CRows(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in
                ("workspace", "user", "agent", "kind", "purpose", "confidence", "flags", "valid_until_us")]


class _CFilter(C.Structure):
    _fields_ = [("workspace", C.c_uint32), ("user", C.c_uint32), ("user_mode", C.c_int32),
                ("agent", C.c_uint32), ("agent_mode", C.c_int32),
                ("kinds", C.c_void_p), ("n_kinds", C.c_int32),
                ("purposes", C.c_void_p), ("n_purposes", C.c_int32),
                ("min_confidence", C.c_float), ("now_us", C.c_int64)]


@dataclass
class RowMeta:
    """Struct-of-arrays row metadata with interned ids (0 = SQL NULL for user/agent)."""
    workspace: np.ndarray
    user: np.ndarray
    agent: np.ndarray
    kind: np.ndarray
    purpose: np.ndarray
    confidence: np.ndarray
    flags: np.ndarray
    valid_until_us: np.ndarray

    @staticmethod
    def uniform(n: int, workspace: int = 1) -> "RowMeta":
        return RowMeta(
            workspace=np.full(n, workspace, np.uint32), user=np.zeros(n, np.uint32),
            agent=np.zeros(n, np.uint32), kind=np.zeros(n, np.uint16), purpose=np.zeros(n, np.uint16),
            confidence=np.ones(n, np.float32),
            flags=np.full(n, FLAG_HAS_EMBEDDING | FLAG_MUTABLE, np.uint8),
            valid_until_us=np.full(n, NO_EXPIRY, np.int64))

    def __len__(self) -> int:
        return len(self.workspace)

    def _c(self) -> _CRows:
        self.workspace = np.ascontiguousarray(self.workspace, np.uint32)
        self.user = np.ascontiguousarray(self.user, np.uint32)
        self.agent = np.ascontiguousarray(self.agent, np.uint32)
        self.kind = np.ascontiguousarray(self.kind, np.uint16)
        self.purpose = np.ascontiguousarray(self.purpose, np.uint16)
        self.confidence = np.ascontiguousarray(self.confidence, np.float32)
        self.flags = np.ascontiguousarray(self.flags, np.uint8)
        self.valid_until_us = np.ascontiguousarray(self.valid_until_us, np.int64)
        return _CRows(*[getattr(self, n).ctypes.data for n, _ in _CRows._fields_])


@dataclass
class Filter:
    """Reference row a8: store.go:250-258, store_read.go:271-278 (single-tier, USER_ANY/USER_EQ)
    and retrieve_multi_tier.go:313-328 (multi-tier, USER_NULL/USER_NULL_OR_EQ)."""
    workspace: int
    user: int = 0
    user_mode: int = USER_ANY
    agent: int = 0
    agent_mode: int = USER_ANY
    kinds: Sequence[int] = field(default_factory=tuple)
    purposes: Sequence[int] = field(default_factory=tuple)
    min_confidence: float = 0.0
    now_us: int = 0


def eligible(meta: RowMeta, flt: Filter) -> np.ndarray:
    n = len(meta)
    out = np.zeros(n, np.uint8)
    kinds = np.ascontiguousarray(list(flt.kinds), np.uint16)
    purposes = np.ascontiguousarray(list(flt.purposes), np.uint16)
    cf = _CFilter(flt.workspace, flt.user, flt.user_mode, flt.agent, flt.agent_mode,
                  kinds.ctypes.data if len(kinds) else None, len(kinds),
                  purposes.ctypes.data if len(purposes) else None, len(purposes),
                  float(flt.min_confidence), int(flt.now_us))
    rows = meta._c()
    lib().omo_eligible(C.byref(rows), n, C.byref(cf), out.ctypes.data)
    return out


# ---------------------------------------------------------------- recall

def search(rows, queries, k: int, eligible_mask: Optional[np.ndarray] = None,
           query_group: Optional[np.ndarray] = None, max_distance: float = float("inf"),
           variant: int = VARIANT_F32_SEQ, threads: int = 0):
    """ORDER BY d ASC LIMIT k per query (store_read.go:266-280).

    rows [n, dim] fp32, queries [nq, dim] fp32, eligible_mask [n] or [groups, n] bytes.
    Returns (row_idx [nq,k] int64 (-1 pad), distance [nq,k] float64 (inf pad), count [nq])."""
    rows, queries = _f32(rows), _f32(queries)
    if queries.ndim == 1:
        queries = queries[None, :]
    n, dim = rows.shape
    if queries.shape[1] != dim:
        raise ValueError(f"different vector dimensions {dim} and {queries.shape[1]}")
    nq = queries.shape[0]
    el = None
    if eligible_mask is not None:
        el = np.ascontiguousarray(eligible_mask, np.uint8).reshape(-1, n)
    qg = None if query_group is None else np.ascontiguousarray(query_group, np.int32)
    out_rows = np.empty((nq, k), np.int64)
    out_dist = np.empty((nq, k), np.float64)
    out_cnt = np.empty(nq, np.int32)
    lib().omo_search_batch(rows.ctypes.data, n, dim, dim, None if el is None else el.ctypes.data,
                           None if qg is None else qg.ctypes.data, queries.ctypes.data, nq, k,
                           float(max_distance), variant, threads, out_rows.ctypes.data,
                           out_dist.ctypes.data, out_cnt.ctypes.data)
    return out_rows, out_dist, out_cnt


def find_similar_observations(rows, meta: RowMeta, flt: Filter, query, k: int = 5,
                              min_similarity: float = 0.85, variant: int = VARIANT_F32_SEQ):
    """Reference a3, store_read.go:243-306.  Returns [(row, similarity)] most similar first.
    maxDistance := 1.0 - minSimilarity in float64 (store_read.go:260); k <= 0 -> 5 (:254-256)."""
    if k <= 0:
        k = 5
    el = eligible(meta, flt)
    r, d, c = search(rows, query, k, el, max_distance=1.0 - float(min_similarity), variant=variant)
    return [(int(r[0, j]), 1.0 - float(d[0, j])) for j in range(int(c[0]))]


def cosine_ranked(rows, meta: RowMeta, entity: np.ndarray, flt: Filter, query, fanout: int,
                  variant: int = VARIANT_F32_SEQ):
    """Cosine arm of the hybrid query, store.go:240-272.

    cosine_ann: top (4*fanout) eligible observations by distance.
    cosine:     DISTINCT ON (entity_id) keeping each entity's nearest observation.
    cosine_ranked: row_number() OVER (ORDER BY cos_dist), LIMIT fanout.
    Returns [(entity, cos_dist, cos_rn)]; ties ordered by (distance, row)."""
    el = eligible(meta, flt)
    r, d, c = search(rows, query, 4 * fanout, el, variant=variant)
    best = {}
    for j in range(int(c[0])):
        e = int(entity[r[0, j]])
        if e not in best:  # rows arrive nearest-first
            best[e] = (float(d[0, j]), int(r[0, j]))
    ranked = sorted(best.items(), key=lambda kv: (kv[1][0], kv[1][1]))[:fanout]
    return [(e, dist, i + 1) for i, (e, (dist, _)) in enumerate(ranked)]


def rrf_fuse(fts_ranked: Sequence[int], cos_ranked: Sequence[int], k: float = 60.0):
    """store.go:272-282: rrf = 1/(60+fts_rn) + 1/(60+cos_rn) over a FULL OUTER JOIN."""
    score = {}
    for i, e in enumerate(fts_ranked):
        score[e] = score.get(e, 0.0) + 1.0 / (k + i + 1)
    for i, e in enumerate(cos_ranked):
        score[e] = score.get(e, 0.0) + 1.0 / (k + i + 1)
    return score


def numpy_topk(rows, queries, k: int, eligible_mask: Optional[np.ndarray] = None):
    """sgemm + argpartition; fp32 BLAS order, sanity only."""
    rows, queries = _f32(rows), _f32(queries)
    rn = np.sqrt(np.einsum("ij,ij->i", rows, rows, dtype=np.float64))
    qn = np.sqrt(np.einsum("ij,ij->i", queries, queries, dtype=np.float64))
    s = (queries @ rows.T).astype(np.float64) / (qn[:, None] * rn[None, :])
    d = 1.0 - np.clip(s, -1.0, 1.0)
    if eligible_mask is not None:
        d = np.where(np.asarray(eligible_mask, bool)[None, :], d, np.inf)
    kk = min(k, d.shape[1])
    idx = np.argpartition(d, kk - 1, axis=1)[:, :kk]
    dd = np.take_along_axis(d, idx, axis=1)
    o = np.lexsort((idx, dd), axis=1)
    return np.take_along_axis(idx, o, axis=1), np.take_along_axis(dd, o, axis=1)


# ---------------------------------------------------------------- self-join

def selfjoin(rows, member: Optional[np.ndarray], kind: np.ndarray, flags: np.ndarray, ids16: np.ndarray,
             floor: float, require_same_kind: bool = True, require_mutable: bool = True,
             variant: int = VARIANT_F32_SEQ, limit: Optional[int] = None):
    """Reference a11, consolidation/prefilter.go:112-138.

    Returns (canonical_row, dupe_row, similarity) sorted by similarity DESC (ties: by rows),
    truncated to ``limit`` (MaxBucketsPerPass) when given.  ``floor`` is rounded to float32
    first, as PreFilterOptions.SimilarityFloor is a Go float32 (prefilter.go:18)."""
    rows = _f32(rows)
    n, dim = rows.shape
    ids16 = np.ascontiguousarray(ids16, np.uint8).reshape(n, 16)
    kind = np.ascontiguousarray(kind, np.uint16)
    flags = np.ascontiguousarray(flags, np.uint8)
    mem = None if member is None else np.ascontiguousarray(member, np.uint8)
    cap = 1 << 16
    while True:
        a = np.empty(cap, np.int64)
        b = np.empty(cap, np.int64)
        s = np.empty(cap, np.float64)
        got = lib().omo_selfjoin(rows.ctypes.data, n, dim, dim, None if mem is None else mem.ctypes.data,
                                 kind.ctypes.data, flags.ctypes.data, ids16.ctypes.data,
                                 C.c_float(np.float32(floor)), int(require_same_kind), int(require_mutable),
                                 variant, cap, a.ctypes.data, b.ctypes.data, s.ctypes.data)
        if got >= 0:
            break
        cap = int(-got)
    a, b, s = a[:got], b[:got], s[:got]
    o = np.lexsort((b, a, -s))
    a, b, s = a[o], b[o], s[o]
    if limit is not None:
        a, b, s = a[:limit], b[:limit], s[:limit]
    return a, b, s


# ---------------------------------------------------------------- projection input (PCA)

def pca_reduce(m, k: int) -> np.ndarray:
    """PCAReduce, ee/pkg/memory/projection/pca.go:28-67, in float64.

    The decomposition itself lives in a dependency that is not under /root/reference
    (gonum.org/v1/gonum v0.17.0, go.mod:67; stat.PC.PrincipalComponents): it copies the data, subtracts
    the column means, takes a thin SVD of the centred copy and returns the right singular vectors
    (nFeatures x min(nSamples, nFeatures)) ordered by decreasing singular value.  The reference then
    multiplies the ORIGINAL, uncentred matrix by the first k vectors (pca.go:60).  Restated here with
    numpy's SVD; the sign of each component is the SVD implementation's choice, so comparisons with
    this function are made up to a sign per column.  Parity unpinned: the reference's tests check
    shapes and clamping only (pca_test.go:17-99), and those are what tests/ transcribes."""
    m = np.ascontiguousarray(m, np.float64)
    if m.ndim != 2 or m.shape[0] == 0 or m.shape[1] == 0:
        raise ValueError(f"projection: empty input matrix ({m.shape[0] if m.ndim == 2 else 0}x"
                         f"{m.shape[1] if m.ndim == 2 else 0})")
    rows, cols = m.shape
    k = max(1, min(k, cols, rows))                       # pca.go:33-42
    if rows < 2:
        return m[:, :k].copy()                           # firstKCols, pca.go:46-48, 70-83
    centred = m - m.mean(axis=0)
    _, _, vt = np.linalg.svd(centred, full_matrices=False)
    k = min(k, vt.shape[0])                              # pca.go:57-59
    return m @ vt[:k].T
