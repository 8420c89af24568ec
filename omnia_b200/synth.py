"""Seeded synthetic corpora for tests and bench.py.

The reference's seeder (demos/memory-seeder/generate.go:178-267) produces text only; vectors
come from whatever embedding provider memory-api is wired to.  These generators stand in for
that provider with the shapes and tie structure the survey calls for.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Optional, Tuple

import numpy as np

from ._lib import NO_EXPIRY, ROW_FORGOTTEN, ROW_HAS_EMBEDDING, ROW_MUTABLE, ROW_SUPERSEDED

CORPUS_SEED, QUERY_SEED, META_SEED = 1, 2, 3   # corpus seed 1 = the seeder default (main.go:106)


def _normalize(x: np.ndarray) -> np.ndarray:
    n = np.sqrt(np.einsum("ij,ij->i", x, x, dtype=np.float64))
    return (x / np.maximum(n, 1e-30)[:, None]).astype(np.float32)


CHUNK = 1 << 16


def gauss_unit_chunk(c: int, dim: int, seed: int = CORPUS_SEED) -> np.ndarray:
    """Rows [c*CHUNK, (c+1)*CHUNK) of the gauss-unit corpus.  Every chunk has its own stream, so a
    rank of a sharded run can produce exactly its rows without holding the whole corpus."""
    rng = np.random.default_rng([seed, c])
    return _normalize(rng.standard_normal((CHUNK, dim), np.float32))


def gauss_unit_local(n_local: int, dim: int, seed: int, rank: int, world: int) -> np.ndarray:
    """A shard of a gauss-unit corpus generated WITHOUT producing the other shards' rows: rank r of
    `world` draws its n_local rows from its own streams.  Same distribution as gauss_unit, but not the
    same rows; for throughput runs on stores too large to generate world times over (bench.py says so)."""
    from concurrent.futures import ThreadPoolExecutor
    import os

    def one(c: int) -> np.ndarray:
        rng = np.random.default_rng([seed, 1_000_003 + 7919 * world + rank, c])
        m = min(CHUNK, n_local - c * CHUNK)
        return _normalize(rng.standard_normal((m, dim), np.float32))

    n_chunks = (n_local + CHUNK - 1) // CHUNK
    workers = min(n_chunks, max(1, (os.cpu_count() or 8) // max(1, min(world, 8))), 16)
    with ThreadPoolExecutor(max_workers=workers) as ex:
        return np.ascontiguousarray(np.concatenate(list(ex.map(one, range(n_chunks))), axis=0))


def gauss_unit(n: int, dim: int, seed: int = CORPUS_SEED, rank: int = 0, world: int = 1) -> np.ndarray:
    """(i) L2-normalised standard-normal rows; with world > 1, the rows rank::world of the
    n-row corpus (round-robin row sharding)."""
    def one(c: int) -> np.ndarray:
        x = gauss_unit_chunk(c, dim, seed)[: min(CHUNK, n - c * CHUNK)]
        if world > 1:
            x = x[(rank - c * CHUNK) % world::world]
        return x

    n_chunks = (n + CHUNK - 1) // CHUNK
    if n_chunks == 1:
        return np.ascontiguousarray(one(0))
    from concurrent.futures import ThreadPoolExecutor
    import os
    with ThreadPoolExecutor(max_workers=min(n_chunks, max(1, (os.cpu_count() or 8) // max(1, min(world, 8))), 16)) as ex:
        parts = list(ex.map(one, range(n_chunks)))   # numpy's generators release the GIL
    return np.ascontiguousarray(np.concatenate(parts, axis=0))


def queries_from(corpus: np.ndarray, nq: int, noise: float = 0.1, seed: int = QUERY_SEED) -> np.ndarray:
    """Queries = corpus rows + noise * gaussian, renormalised."""
    rng = np.random.default_rng(seed)
    idx = rng.integers(0, corpus.shape[0], nq)
    q = corpus[idx] + noise * rng.standard_normal((nq, corpus.shape[1]), np.float32) / np.sqrt(corpus.shape[1])
    return _normalize(q)


def clustered(n: int, dim: int, seed: int = CORPUS_SEED, topics: int = 6, variants: int = 160,
              sigma: float = 0.02, dup_fraction: float = 0.3) -> np.ndarray:
    """(ii) 6 topic centroids x 160 phrase variants (mirrors generate.go:71-175) + noise, with
    exact duplicates kept so that ties are exercised."""
    rng = np.random.default_rng(seed)
    cent = _normalize(rng.standard_normal((topics, dim), np.float32))
    base = np.repeat(cent, variants, axis=0) + 0.35 * _normalize(
        rng.standard_normal((topics * variants, dim), np.float32))
    base = _normalize(base)                       # 960 distinct "sentences"
    pick = rng.integers(0, base.shape[0], n)
    out = base[pick].copy()
    noisy = np.flatnonzero(rng.random(n) >= dup_fraction)   # the rest stay exact duplicates
    out[noisy] += sigma * _normalize(rng.standard_normal((noisy.size, dim), np.float32))
    out[noisy] = _normalize(out[noisy])
    return out


def low_entropy(n: int, dim: int, seed: int = CORPUS_SEED, clusters: int = 16, per_cluster: int = 300
                ) -> Tuple[np.ndarray, np.ndarray]:
    """(v) vectors whose bf16 rounding errors are COHERENT: few-valued and constant rows in the style of the
    reference's own test vectors (`repeatFloat(0.1, 1536)` and one-hot mixtures,
    internal/memory/store_test.go:912-933).  Every element of a constant row rounds the same way, so the
    approximate score of two such rows is off by up to ~2 * 2^-9 near cosine 1 instead of averaging out.
    Layout: `clusters` centres (centre 0 is the constant 0.1 vector, the others are +-0.1 / +-0.2 sign
    patterns); around each, `per_cluster` near-parallel members = centre * scale + a one-hot twist whose size
    puts the member's cosine to the centre in [0.995, 0.9999]; the rest are few-valued background rows.
    Members of one cluster are spread over consecutive rows AND far apart, so some share a sub-bucket.
    Returns (corpus, queries): one query per centre (the centre itself with a small one-hot twist)."""
    rng = np.random.default_rng([seed, 77])
    cent = rng.choice(np.array([0.1, -0.1, 0.2, -0.2], np.float32), (clusters, dim))
    cent[0] = 0.1
    out = rng.choice(np.array([0.1, -0.1, 0.3], np.float32), (n, dim))     # background: three values only
    per = min(per_cluster, max(1, n // (2 * clusters)))
    for c in range(clusters):
        norm = float(np.sqrt((cent[c].astype(np.float64) ** 2).sum()))
        start = int(rng.integers(0, max(1, n - per)))
        rows = np.concatenate([np.arange(start, start + per // 2),                       # neighbours in one sub-bucket
                               rng.choice(n, per - per // 2, replace=False)]) % n       # and scattered ones
        scale = rng.choice(np.array([0.5, 1.0, 2.0, 3.0], np.float32), rows.size)
        a = rng.uniform(0.015, 0.1, rows.size).astype(np.float32) * norm     # twist size relative to |centre|
        m = cent[c][None, :] * scale[:, None]
        m[np.arange(rows.size), rng.integers(0, dim, rows.size)] += a * scale
        out[rows] = m
    q = cent.copy()
    q[np.arange(clusters), rng.integers(0, dim, clusters)] *= 1.02
    return np.ascontiguousarray(out, np.float32), np.ascontiguousarray(q, np.float32)


def planted_dups(n: int, dim: int, seed: int = CORPUS_SEED, n_pairs: int = 64,
                 sims: Tuple[float, ...] = (0.90, 0.918, 0.922, 0.95, 0.999)) -> Tuple[np.ndarray, np.ndarray]:
    """(iii) gauss-unit corpus with pairs planted at chosen cosines (none within 1e-6 of
    (double)(float)0.92).  Returns (corpus, planted [(i, j, sim)])."""
    rng = np.random.default_rng(seed + 100)
    x = gauss_unit(n, dim, seed)
    planted = []
    idx = rng.permutation(n)[: 2 * n_pairs].reshape(n_pairs, 2)
    for p, (i, j) in enumerate(idx):
        s = sims[p % len(sims)]
        a = x[i].astype(np.float64)
        r = rng.standard_normal(dim)
        r -= r.dot(a) * a
        r /= np.linalg.norm(r)
        x[j] = (s * a + np.sqrt(1 - s * s) * r).astype(np.float32)
        planted.append((int(i), int(j), float(s)))
    return x, np.array(planted, dtype=np.float64)


@dataclass
class Metadata:
    n_workspaces: int
    workspace_idx: np.ndarray
    user_idx: np.ndarray
    agent_idx: np.ndarray
    kind_idx: np.ndarray
    purpose_idx: np.ndarray
    confidence: np.ndarray
    flags: np.ndarray
    valid_until_us: np.ndarray
    now_us: int

    def tables(self, n_users: int = 64, n_agents: int = 8, n_kinds: int = 8, n_purposes: int = 4):
        return dict(workspaces=[f"ws-{i:04d}" for i in range(self.n_workspaces)],
                    users=[f"user-{i}" for i in range(n_users)], agents=[f"agent-{i}" for i in range(n_agents)],
                    kinds=[f"kind-{i}" for i in range(n_kinds)], purposes=[f"purpose-{i}" for i in range(n_purposes)])


def metadata(n: int, seed: int = META_SEED, n_workspaces: int = 1000, zipf_a: float = 1.2, n_users: int = 64,
             n_agents: int = 8, n_kinds: int = 8, n_purposes: int = 4) -> Metadata:
    """(iv) workspaces Zipf-distributed, 10 % forgotten, 5 % superseded, 1 % expired,
    confidence U(0.3, 1), 8 kinds."""
    rng = np.random.default_rng(seed)
    ws = (rng.zipf(zipf_a, n) - 1) % n_workspaces
    now = 1_700_000_000_000_000
    flags = np.full(n, ROW_HAS_EMBEDDING | ROW_MUTABLE, np.uint8)
    flags[rng.random(n) < 0.10] |= ROW_FORGOTTEN
    flags[rng.random(n) < 0.05] |= ROW_SUPERSEDED
    flags[rng.random(n) < 0.02] &= ~np.uint8(ROW_MUTABLE)
    valid = np.full(n, NO_EXPIRY, np.int64)
    exp = rng.random(n) < 0.01
    valid[exp] = now - rng.integers(1, 10**9, int(exp.sum()))
    fut = rng.random(n) < 0.01
    valid[fut] = now + rng.integers(1, 10**9, int(fut.sum()))
    user = rng.integers(-1, n_users, n).astype(np.int32)      # -1 = NULL
    agent = rng.integers(-1, n_agents, n).astype(np.int32)
    return Metadata(n_workspaces, ws.astype(np.uint32), user, agent,
                    rng.integers(0, n_kinds, n).astype(np.uint16), rng.integers(0, n_purposes, n).astype(np.uint16),
                    rng.uniform(0.3, 1.0, n).astype(np.float32), flags, valid, now)
