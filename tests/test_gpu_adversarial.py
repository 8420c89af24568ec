"""Adversarial inputs for the bf16 stage's exactness proof (What's weak 1).

The approximate scorer rounds unit vectors to bf16.  On Gaussian data the per-element rounding errors are
independent and the score error is ~1e-4; on constant or few-valued vectors -- the reference's own test
shapes, internal/memory/store_test.go:912-933 -- every element rounds the same way and the error is coherent,
up to ~4e-3 near cosine 1.  With the deterministic margin (the default) ids and order must still be the
oracle's; a statistical margin would pass wrong answers here without noticing."""
import numpy as np
import pytest

from omnia_b200 import GPUMemoryStore, FilterSpec, synth
from oracle import oracle as O
from helpers import assert_topk_matches

pytestmark = pytest.mark.gpu


def _bf16_score_error(corpus, queries):
    """max |bf16-stage cosine - exact cosine| over the query x row pairs of a small sample (numpy)."""
    def unit_bf16(x):
        u = (x / np.sqrt((x.astype(np.float64) ** 2).sum(1))[:, None]).astype(np.float32)
        b = u.view(np.uint32)
        r = ((b + 0x7FFF + ((b >> 16) & 1)) & 0xFFFF0000).astype(np.uint32)   # round to nearest even
        return r.view(np.float32), u
    rb, ru = unit_bf16(corpus)
    qb, qu = unit_bf16(queries)
    return float(np.abs(qb.astype(np.float64) @ rb.astype(np.float64).T - qu.astype(np.float64) @ ru.astype(np.float64).T).max())


@pytest.mark.parametrize("path", [0, 2, 4])
@pytest.mark.parametrize("n,dim", [(125_000, 128), (250_000, 128), (1_000_000, 128)])
def test_low_entropy_vectors_keep_ids_and_order(path, n, dim):
    corpus, centres = synth.low_entropy(n, dim)
    # 9 copies of every centre query with growing one-hot twists: 144 queries = two query tiles, path 0 -> tensor cores
    rng = np.random.default_rng(5)
    queries = np.repeat(centres, 9, axis=0)
    queries[np.arange(queries.shape[0]), rng.integers(0, dim, queries.shape[0])] *= rng.uniform(1.0, 1.3, queries.shape[0]).astype(np.float32)
    if n == 125_000 and path == 0:
        err = _bf16_score_error(corpus[:4000], queries[:18])
        assert err > 1.5e-3, f"this corpus is meant to break a 1e-3 margin, coherent error only {err:.2e}"
    with GPUMemoryStore(dim, n, force_path=path) as s:
        s.bulk_load(corpus)
        g = s.search(queries, [FilterSpec("default")], 10)
        st = s.stats()
    r = O.search(corpus, queries, 10, variant=O.VARIANT_F64)
    assert_topk_matches(g[0], g[1], g[2], r[0], r[1], r[2], corpus, queries, what=f"low-entropy n={n} path={path}")
    print(f"low-entropy n={n} path={path}: fallback_queries={st.fallback_queries} rescued_rows={st.rescued_rows} "
          f"listed={st.listed_entries}")


@pytest.mark.parametrize("path", [1, 2])
@pytest.mark.parametrize("dim", [64, 100])
def test_hidden_near_duplicates_in_one_sub_bucket_at_small_dim(path, dim):
    # at small dim the bf16 error is larger relative to a fixed statistical margin; three or more
    # near-duplicates of the query inside one sub-bucket hide behind its two keys (top-2 scorers)
    n = 40_000
    rng = np.random.default_rng(dim)
    corpus = synth.gauss_unit(n, dim, seed=3)
    nq = 24 if path == 2 else 6
    queries = synth.queries_from(corpus, nq, noise=0.0, seed=8)
    for i in range(nq):   # five near-duplicates of query i on consecutive rows (same 32-row sub-bucket)
        base = 1000 * (i + 1)
        for j in range(5):
            v = queries[i] + (2e-3 * (j + 1)) * rng.standard_normal(dim).astype(np.float32) / np.sqrt(dim)
            corpus[base + j] = v / np.linalg.norm(v)
    with GPUMemoryStore(dim, n, force_path=path) as s:
        s.bulk_load(corpus)
        g = s.search(queries, [FilterSpec("default")], 4)
    r = O.search(corpus, queries, 4, variant=O.VARIANT_F64)
    assert_topk_matches(g[0], g[1], g[2], r[0], r[1], r[2], corpus, queries, what=f"hidden dups dim={dim} path={path}")
