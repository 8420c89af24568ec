"""CUDA recall vs the CPU oracle on seeded synthetic corpora.

Every case runs through the C ABI (GPUMemoryStore.search -> omr_search).  force_path pins the
scorer: 1 = HBM-streaming, 2 = tcgen05 tensor-core, 3 = exact scan; 0 = the engine's own choice."""
import numpy as np
import pytest

from omnia_b200 import GPUMemoryStore, FilterSpec, synth
from omnia_b200 import MATCH_NULL_OR_EQ, MATCH_NULL_ONLY
from oracle import oracle as O
from helpers import assert_topk_matches

pytestmark = pytest.mark.gpu


def run_case(corpus, queries, k, path, sub_bucket=0, max_distance=float("nan"), rescore_width=0):
    n, dim = corpus.shape
    with GPUMemoryStore(dim, n, force_path=path, sub_bucket=sub_bucket, rescore_width=rescore_width) as s:
        s.bulk_load(corpus)
        g_rows, g_dist, g_cnt, _ = s.search(queries, [FilterSpec("default", max_distance=max_distance)], k)
        st = s.stats()
    md = float("inf") if np.isnan(max_distance) else max_distance
    r_rows, r_dist, r_cnt = O.search(corpus, queries, k, max_distance=md, variant=O.VARIANT_F64)
    assert_topk_matches(g_rows, g_dist, g_cnt, r_rows, r_dist, r_cnt, corpus, queries,
                        what=f"path={path} n={n} dim={dim} k={k}")
    return st


@pytest.mark.parametrize("path", [1, 2, 3, 4])
@pytest.mark.parametrize("n,dim,nq,k", [
    (10_000, 1536, 64, 10),      # BASELINE config C1
    (5_000, 384, 5, 10),
    (3_001, 100, 9, 5),          # ragged: n not a multiple of 256, dim not a multiple of 64
    (300, 768, 130, 100),        # k close to n, two query tiles
    (77, 64, 3, 100),            # k > n
    (20_000, 128, 40, 200),      # K2 = 512: the unfused select / rescore / finalize chain
])
def test_gauss_corpus_matches_oracle(path, n, dim, nq, k):
    corpus = synth.gauss_unit(n, dim)
    queries = synth.queries_from(corpus, nq)
    st = run_case(corpus, queries, k, path)
    assert st.last_path == path
    if path != 3 and n >= 3000 and k <= 10:
        assert st.fallback_queries == 0, "random data must certify without the exact fallback"


@pytest.mark.parametrize("path", [0, 1, 2, 3])
@pytest.mark.parametrize("n,dim,nq,k", [
    (1_500, 2000, 3, 10),        # the widest indexed column (embedding_schema.go:64-68)
    (5_000, 8, 129, 1),          # one 16-byte row; k = 1; a ragged second query tile
    (4_097, 1, 4, 3),            # dimension 1: every cosine is +1 or -1, ties everywhere
    (6_000, 64, 2, 4096),        # the largest k the ABI accepts, k < n
    (9_000, 32, 2049, 3),        # 17 query tiles, the last with one query
    (256, 96, 7, 256),           # exactly one row tile, k = n
    (257, 96, 7, 10),            # one row in the second tile
])
def test_limits_of_the_abi(path, n, dim, nq, k):
    if path == 1 and nq > 64:
        pytest.skip("the streaming scorer takes 8 queries per pass; large batches belong to the tensor-core path")
    corpus = synth.gauss_unit(n, dim, seed=n + dim)
    queries = synth.queries_from(corpus, nq)
    run_case(corpus, queries, k, path)


@pytest.mark.parametrize("path", [0, 1, 2])
def test_rows_and_queries_of_any_magnitude(path):
    # cosine ignores magnitudes; pgvector divides by sqrt(na * nb) in double, so 1e-12 .. 1e12 scales are exact
    n, dim, nq, k = 6_000, 384, 24, 10
    rng = np.random.default_rng(9)
    corpus = synth.gauss_unit(n, dim, seed=21)
    queries = synth.queries_from(corpus, nq)
    corpus = (corpus * (10.0 ** rng.uniform(-12, 12, (n, 1)))).astype(np.float32)
    queries = (queries * (10.0 ** rng.uniform(-12, 12, (nq, 1)))).astype(np.float32)
    run_case(corpus, queries, k, path)


@pytest.mark.parametrize("path", [1, 2, 4])
@pytest.mark.parametrize("S", [32, 64, 128, 256])
def test_every_sub_bucket_size(path, S):
    corpus = synth.gauss_unit(4_000, 256)
    queries = synth.queries_from(corpus, 12)
    run_case(corpus, queries, 10, path, sub_bucket=S)


@pytest.mark.parametrize("path", [1, 2])
def test_against_pgvector_faithful_fp32_oracle(path):
    # scores within 1e-5 relative of the fp32 reference, ids equal except ties within 1e-6
    corpus = synth.gauss_unit(8_000, 1536)
    queries = synth.queries_from(corpus, 16)
    with GPUMemoryStore(1536, 8_000, force_path=path) as s:
        s.bulk_load(corpus)
        g = s.search(queries, [FilterSpec("default")], 10)
    r = O.search(corpus, queries, 10, variant=O.VARIANT_F32_SEQ)
    assert_topk_matches(g[0], g[1], g[2], r[0], r[1], r[2], what="fp32-seq")


@pytest.mark.parametrize("path", [1, 2, 4])
def test_clustered_corpus_with_exact_duplicates(path):
    # tie-heavy data: suspect sub-buckets must be rescued (or fall back) and still be exact
    corpus = synth.clustered(20_000, 384)
    queries = synth.queries_from(corpus, 24, noise=0.05)
    st = run_case(corpus, queries, 10, path)
    assert st.rescued_rows + st.fallback_queries > 0, "this corpus is built to trip the certificate"


@pytest.mark.parametrize("path", [1, 2])
def test_distance_cutoff(path):
    corpus = synth.gauss_unit(6_000, 128)
    queries = synth.queries_from(corpus, 10, noise=0.5)
    run_case(corpus, queries, 20, path, max_distance=0.6)
    run_case(corpus, queries, 20, path, max_distance=0.05)


@pytest.mark.parametrize("path", [1, 2])
def test_negative_similarities_rank_correctly(path):
    # a store where every row is anti-correlated with the query: keys of negative scores must order
    rng = np.random.default_rng(5)
    q = synth.gauss_unit(1, 64, seed=9)
    corpus = (-q * rng.uniform(0.2, 1.0, (900, 1)) + 0.3 * synth.gauss_unit(900, 64, seed=10)).astype(np.float32)
    run_case(corpus, np.repeat(q, 3, 0), 10, path)


@pytest.mark.parametrize("path", [0, 1, 2, 4])
def test_filters_and_mixed_groups_in_one_batch(path):
    n, dim, nq = 30_000, 256, 40 if path != 1 else 8
    corpus = synth.gauss_unit(n, dim)
    queries = synth.queries_from(corpus, nq)
    md = synth.metadata(n, n_workspaces=5, zipf_a=1.5)
    tabs = md.tables()
    with GPUMemoryStore(dim, n, force_path=path) as s:
        s.bulk_load(corpus, workspace_idx=md.workspace_idx, user_idx=md.user_idx, agent_idx=md.agent_idx,
                    kind_idx=md.kind_idx, purpose_idx=md.purpose_idx, confidence=md.confidence, flags=md.flags,
                    valid_until_us=md.valid_until_us, **tabs)
        specs = [
            (FilterSpec("ws-0000", now_us=md.now_us), O.Filter(workspace=1, now_us=md.now_us)),
            (FilterSpec("ws-0001", "user-3", None, None, None, min_confidence=0.5, now_us=md.now_us),
             O.Filter(workspace=2, user=4, user_mode=O.USER_EQ, min_confidence=0.5, now_us=md.now_us)),
            (FilterSpec("ws-0000", "user-7", MATCH_NULL_OR_EQ, None, MATCH_NULL_ONLY, kinds=["kind-1", "kind-2"],
                        now_us=md.now_us),
             O.Filter(workspace=1, user=8, user_mode=O.USER_NULL_OR_EQ, agent_mode=O.USER_NULL, kinds=(1, 2),
                      now_us=md.now_us)),
            (FilterSpec("ws-0002", purposes=["purpose-0"], max_distance=0.95, now_us=md.now_us),
             O.Filter(workspace=3, purposes=(0,), now_us=md.now_us)),
            (FilterSpec("no-such-workspace", now_us=md.now_us), O.Filter(workspace=999, now_us=md.now_us)),
        ]
        qf = np.arange(nq, dtype=np.int32) % len(specs)
        g_rows, g_dist, g_cnt, _ = s.search(queries, [a for a, _ in specs], 10, qf)
    # oracle ids: interned in table order starting at 1 (0 = NULL); kinds/purposes from 0
    meta = O.RowMeta(md.workspace_idx + 1, (md.user_idx + 1).astype(np.uint32), (md.agent_idx + 1).astype(np.uint32),
                     md.kind_idx, md.purpose_idx, md.confidence, md.flags, md.valid_until_us)
    el = np.stack([O.eligible(meta, b) for _, b in specs])
    assert el[0].sum() > 100 and el[4].sum() == 0
    for gi, (_, _) in enumerate(specs):
        sel = np.flatnonzero(qf == gi)
        mdist = 0.95 if gi == 3 else float("inf")
        r = O.search(corpus, queries[sel], 10, el[gi], max_distance=mdist, variant=O.VARIANT_F64)
        assert_topk_matches(g_rows[sel], g_dist[sel], g_cnt[sel], r[0], r[1], r[2], what=f"filter {gi}")


def test_selective_filter_with_few_eligible_rows():
    # far fewer eligible rows than k: nothing may be lost and nothing foreign may leak
    n, dim = 10_000, 128
    corpus = synth.gauss_unit(n, dim)
    ws = np.zeros(n, np.uint32); ws[::997] = 1
    with GPUMemoryStore(dim, n) as s:
        s.bulk_load(corpus, workspaces=["big", "tiny"], workspace_idx=ws)
        q = synth.queries_from(corpus, 20)
        g = s.search(q, [FilterSpec("tiny")], 50)
    el = (ws == 1).astype(np.uint8)
    r = O.search(corpus, q, 50, el, variant=O.VARIANT_F64)
    assert int(r[2][0]) == int(el.sum()) < 50
    assert_topk_matches(g[0], g[1], g[2], r[0], r[1], r[2], what="tiny workspace")


def test_incremental_upserts_equal_bulk_load():
    import uuid
    dim = 64
    corpus = synth.gauss_unit(600, dim)
    q = synth.queries_from(corpus, 4)
    with GPUMemoryStore(dim, 1000) as a, GPUMemoryStore(dim, 1000) as b:
        a.bulk_load(corpus)
        b.bulk_load(corpus[:300])
        for i in range(300, 600):
            b.upsert(i + 10**6, i + 10**6, {"workspace_id": "default"}, corpus[i])
        ra = a.search(q, [FilterSpec("default")], 10)
        rb = b.search(q, [FilterSpec("default")], 10)
    assert np.array_equal(ra[0], rb[0]) and np.array_equal(ra[1], rb[1])


def test_empty_store_and_empty_batch():
    with GPUMemoryStore(32, 100) as s:
        g = s.search(np.ones((2, 32), np.float32), [FilterSpec("default")], 5)
        assert g[2].tolist() == [0, 0] and (g[0] == -1).all()


@pytest.mark.parametrize("dim", [1536])
def test_full_size_store_properties(dim):
    """BASELINE config C2 at full size (1M x 1536, 1024 queries, k=10): size-independent
    properties plus an oracle check on a query sample."""
    n, nq, k = 1_000_000, 1024, 10
    corpus = synth.gauss_unit(n, dim)
    queries = synth.queries_from(corpus, nq)
    rng = np.random.default_rng(synth.QUERY_SEED)
    src = rng.integers(0, n, nq)                      # the rows queries_from() perturbed
    with GPUMemoryStore(dim, n) as s:
        s.bulk_load(corpus)
        rows, dist, cnt, _ = s.search(queries, [FilterSpec("default")], k)
        st = s.stats()
        # exact copies of stored rows must come back first at distance ~0
        rows2, dist2, _, _ = s.search(corpus[src[:128]], [FilterSpec("default")], k)
    assert st.last_path == 4          # 1024 queries: the CTA-pair tensor-core scorer
    assert (cnt == k).all() and (np.diff(dist, axis=1) >= 0).all()
    assert (rows[:, 0] == src).mean() > 0.99          # the perturbed source row is the nearest
    assert (rows2[:, 0] == src[:128]).all() and (dist2[:, 0] < 1e-12).all()
    sample = np.arange(0, nq, 64)
    r = O.search(corpus, queries[sample], k, variant=O.VARIANT_F64)
    assert_topk_matches(rows[sample], dist[sample], cnt[sample], r[0], r[1], r[2], what="1M x 1536 sample")
    assert st.fallback_queries == 0
