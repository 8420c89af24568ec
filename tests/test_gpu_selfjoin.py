"""Consolidation near-duplicate self-join (reference a11) vs the oracle: identical pair sets at
the stated threshold, uuid-ordered canonical/dupe assignment, predicates, ORDER BY / LIMIT."""
import numpy as np
import pytest

from omnia_b200 import GPUMemoryStore, synth, ROW_HAS_EMBEDDING, ROW_MUTABLE
from omnia_b200.store import id_bytes
from oracle import oracle as O

pytestmark = pytest.mark.gpu


def gpu_pairs(store, ws, floor, limit=0, **kw):
    out = store.run_entity_duplicate_candidates(ws, floor, limit, **kw)
    return [(p.canonical_row, p.dupe_row, p.similarity) for p in out], out


@pytest.mark.parametrize("path", [0, 2])          # 0 = the CTA-pair (cta_group::2) join kernel, 2 = the 1-CTA kernel
@pytest.mark.parametrize("n,dim", [(4096, 256), (3000, 1536), (700, 100), (13_000, 128)])
def test_planted_duplicates_pair_set_is_identical(n, dim, path):
    corpus, planted = synth.planted_dups(n, dim, n_pairs=40)
    rng = np.random.default_rng(11)
    ids = rng.integers(0, 256, (n, 16)).astype(np.uint8)          # random uuids: order unrelated to rows
    kind = rng.integers(0, 2, n).astype(np.uint16)
    flags = np.full(n, ROW_HAS_EMBEDDING | ROW_MUTABLE, np.uint8)
    flags[rng.random(n) < 0.05] &= ~np.uint8(ROW_MUTABLE)
    with GPUMemoryStore(dim, n, force_path=path) as s:
        s.bulk_load(corpus, observation_ids=ids, entity_ids=ids, kinds=["a", "b"], kind_idx=kind, flags=flags)
        for floor in (0.92, 0.85):
            got, _ = gpu_pairs(s, "default", floor)
            a, b, sim = O.selfjoin(corpus, None, kind, flags, ids, floor, variant=O.VARIANT_F64)
            want = {(int(x), int(y)) for x, y in zip(a, b)}
            assert {(c, d) for c, d, _ in got} == want, f"floor {floor}"
            assert len(want) > 0
            gs = {(c, d): v for c, d, v in got}
            for x, y, v in zip(a, b, sim):
                assert abs(gs[(int(x), int(y))] - v) < 1e-9
            assert all(got[i][2] >= got[i + 1][2] for i in range(len(got) - 1))   # ORDER BY similarity DESC
        # predicates off -> a superset; LIMIT truncates after the ordering
        all_pairs, _ = gpu_pairs(s, "default", 0.92, require_same_kind=False, require_mutable=False)
        a2, b2, _ = O.selfjoin(corpus, None, kind, flags, ids, 0.92, False, False, variant=O.VARIANT_F64)
        assert {(c, d) for c, d, _ in all_pairs} == {(int(x), int(y)) for x, y in zip(a2, b2)}
        top3, objs = gpu_pairs(s, "default", 0.92, limit=3, require_same_kind=False, require_mutable=False)
        assert top3 == all_pairs[:3]
        assert objs[0].key.startswith("dupe:") and bytes(id_bytes(objs[0].canonical_id)) < bytes(id_bytes(objs[0].dupe_id))


def test_selfjoin_is_scoped_to_one_workspace_and_float32_floor():
    dim = 64
    base = synth.gauss_unit(1, dim)[0]
    def mix(c, seed):
        r = synth.gauss_unit(1, dim, seed=seed)[0].astype(np.float64)
        a = base.astype(np.float64); r -= r.dot(a) * a; r /= np.linalg.norm(r)
        return (c * a + np.sqrt(1 - c * c) * r).astype(np.float32)
    corpus = np.stack([base, mix(0.99, 2), mix(0.99, 3), base, mix(0.5, 4)])
    ws = np.array([0, 0, 1, 1, 0], np.uint32)
    with GPUMemoryStore(dim, 16) as s:
        s.bulk_load(corpus, workspaces=["w0", "w1"], workspace_idx=ws)
        p0, _ = gpu_pairs(s, "w0", 0.9)
        p1, _ = gpu_pairs(s, "w1", 0.9)
        assert {(c, d) for c, d, _ in p0} == {(0, 1)} and {(c, d) for c, d, _ in p1} == {(2, 3)}
        assert gpu_pairs(s, "nope", 0.9)[0] == []
        # (double)(float)0.99 = 0.99000000953...: a pair at exactly 0.99 (to fp rounding) may fall either
        # side, but the oracle and the engine must agree
        a, b, _ = O.selfjoin(corpus[:2], None, np.zeros(2, np.uint16),
                             np.full(2, ROW_HAS_EMBEDDING | ROW_MUTABLE, np.uint8),
                             np.array([[0] * 15 + [0], [0] * 15 + [1]], np.uint8), 0.99, variant=O.VARIANT_F64)
        g, _ = gpu_pairs(s, "w0", 0.99)
        assert len(g) == len(a)


@pytest.mark.parametrize("parts", [2, 3, 8])
def test_block_partitions_of_the_pair_space_union_to_the_whole_join(parts):
    """The multi-GPU form: G ranks with a replicated matrix each scan one interleaved
    partition of the tile pairs; here the G parts run one after another on one GPU."""
    n, dim = 5000, 128
    corpus, _ = synth.planted_dups(n, dim, n_pairs=60)
    with GPUMemoryStore(dim, n) as s:
        s.bulk_load(corpus)
        whole, _ = gpu_pairs(s, "default", 0.9)
        union = []
        for p in range(parts):
            got, _ = gpu_pairs(s, "default", 0.9, part_index=p, part_count=parts)
            union += got
        assert len(union) == len(whole) and sorted(union) == sorted(whole)
        assert len(whole) >= 30
        top5 = sorted(union, key=lambda t: (-t[2], t[0], t[1]))[:5]
        assert top5 == whole[:5]


@pytest.mark.parametrize("path", [0, 2])
@pytest.mark.parametrize("sb", [1, 2, 5])
def test_super_block_walk_covers_every_tile_pair_once(monkeypatch, sb, path):
    # 2900 rows = 12 B tiles (the last one ragged); super-block edges of 1, 2 and 5 tiles give 12, 6 and 3
    # block columns, with full, diagonal and ragged super-blocks.  Every planted pair must appear exactly once.
    monkeypatch.setenv("OMR_SELFJOIN_SB", str(sb))
    n, dim = 2900, 192
    corpus, planted = synth.planted_dups(n, dim, n_pairs=120, sims=(0.93, 0.97, 0.999))
    with GPUMemoryStore(dim, n, force_path=path) as s:
        s.bulk_load(corpus)
        pairs = s.run_entity_duplicate_candidates("default", 0.92, 0, require_same_kind=False)
    got = [(min(p.canonical_row, p.dupe_row), max(p.canonical_row, p.dupe_row)) for p in pairs]
    want = sorted((min(int(i), int(j)), max(int(i), int(j))) for i, j, _ in planted)
    assert sorted(got) == want and len(set(got)) == len(got)


# ---- ORDER BY similarity DESC LIMIT n on the device (consolidation/prefilter.go:129-130; production call: floor 0.85,
# ---- limit 100, ee/pkg/memory/consolidation/worker.go:432-441)

def _rows_ids(n):
    from helpers import ids_from_rows
    return ids_from_rows(n)


@pytest.mark.parametrize("path", [0, 2])
def test_limit_selects_the_oracles_top_n_in_order(path):
    n, dim, limit = 13_000, 128, 100
    corpus, _ = synth.planted_dups(n, dim, n_pairs=400)
    kind = (np.arange(n) % 3 == 0).astype(np.uint16)
    flags = np.full(n, ROW_HAS_EMBEDDING | ROW_MUTABLE, np.uint8)
    with GPUMemoryStore(dim, n, force_path=path) as s:
        s.bulk_load(corpus, kinds=["a", "b"], kind_idx=kind, flags=flags)
        for same_kind in (False, True):
            got, _ = gpu_pairs(s, "default", 0.85, limit=limit, require_same_kind=same_kind)
            a, b, sim = O.selfjoin(corpus, None, kind, flags, _rows_ids(n), 0.85, same_kind, True, variant=O.VARIANT_F64,
                                   limit=limit)
            assert len(got) == len(a) == limit
            assert [(c, d) for c, d, _ in got] == [(int(x), int(y)) for x, y in zip(a, b)]
            assert max(abs(g[2] - v) for g, v in zip(got, sim)) < 1e-9
        st = s.stats()
    assert st.join_candidates > 0


def test_clustered_corpus_limit_100_equals_the_oracle_up_to_ties():
    """The seeder-shaped corpus (6 topics x 160 sentences, 30 % exact duplicates): every same-topic pair clears a
    0.85 floor, and thousands of exact-duplicate pairs tie at similarity 1 to the last bit or two.  The engine's
    top 100 must be the oracle's top 100 up to the identity of pairs whose similarities agree within 1e-9."""
    n, dim, limit = 8_000, 384, 100
    corpus = synth.clustered(n, dim)
    kind = np.zeros(n, np.uint16)
    flags = np.full(n, ROW_HAS_EMBEDDING | ROW_MUTABLE, np.uint8)
    with GPUMemoryStore(dim, n) as s:
        s.bulk_load(corpus)
        got, _ = gpu_pairs(s, "default", 0.85, limit=limit)
        st = s.stats()
    a, b, sim = O.selfjoin(corpus, None, kind, flags, _rows_ids(n), 0.85, variant=O.VARIANT_F64)
    assert len(a) > 3_000_000, "this corpus is meant to put millions of pairs above the floor"
    assert st.join_candidates < len(a) // 4, "the running cut must keep the candidate list far below every pair above the floor"
    key = a * n + b                                   # numpy lookup: a Python dict of 1e7 tuples takes a minute
    order = np.argsort(key)
    key, sim_by_key = key[order], sim[order]
    assert len(got) == limit and len({(c, d) for c, d, _ in got}) == limit
    for i, (c, d, v) in enumerate(got):
        j = int(np.searchsorted(key, c * n + d))
        assert j < len(key) and key[j] == c * n + d, "a returned pair is not a pair of the join"
        assert abs(sim_by_key[j] - v) < 1e-9, "a returned pair's similarity is off"
        assert abs(v - sim[i]) < 1e-9, f"rank {i}: similarity {v} vs the oracle's {sim[i]}"
    # nothing strictly better than the last returned pair may be missing
    better = np.flatnonzero(sim > got[-1][2] + 1e-9)
    have = {(c, d) for c, d, _ in got}
    assert all((int(a[i]), int(b[i])) in have for i in better)


def test_clustered_200k_limit_100_structural():
    """SURVEY.md section 8(d)(ii) at a size the O(n^2) oracle cannot reach: the 100 best pairs of a 200k-row
    seeder-shaped corpus are exact duplicates (similarity 1 up to the last bits), distinct, canonical before dupe."""
    n, dim = 200_000, 384
    corpus = synth.clustered(n, dim)
    with GPUMemoryStore(dim, n) as s:
        s.bulk_load(corpus)
        got, _ = gpu_pairs(s, "default", 0.85, limit=100)
        st = s.stats()
    assert len(got) == 100 and len({(c, d) for c, d, _ in got}) == 100
    for c, d, v in got:
        assert c < d and v > 1.0 - 1e-12 and np.array_equal(corpus[c], corpus[d])
    assert all(got[i][2] >= got[i + 1][2] for i in range(99))
    # every same-topic pair clears the floor (~3.3e9 of them); the running cut keeps the list near the duplicate clusters
    assert st.join_candidates < 200_000_000


def test_candidate_list_regrows_when_it_is_too_short(monkeypatch):
    n, dim = 5000, 128
    corpus, _ = synth.planted_dups(n, dim, n_pairs=120)
    with GPUMemoryStore(dim, n) as s:
        s.bulk_load(corpus)
        want, _ = gpu_pairs(s, "default", 0.9)
    monkeypatch.setenv("OMR_SELFJOIN_PCAP", "16")      # first pass counts, second pass has room
    with GPUMemoryStore(dim, n) as s:
        s.bulk_load(corpus)
        got, _ = gpu_pairs(s, "default", 0.9)
        top7, _ = gpu_pairs(s, "default", 0.9, limit=7)
    assert len(want) > 16 and got == want and top7 == want[:7]


def test_cancel_flag_stops_the_join(monkeypatch):
    import ctypes
    from omnia_b200 import RecallError
    from omnia_b200 import _lib as L
    monkeypatch.setenv("OMR_SELFJOIN_CHUNK", "64")     # many chunks, so the flag is looked at
    n, dim = 6000, 128
    corpus, _ = synth.planted_dups(n, dim, n_pairs=20)
    with GPUMemoryStore(dim, n) as s:
        s.bulk_load(corpus)
        flag = ctypes.c_int32(1)
        with pytest.raises(RecallError, match="cancelled") as e:
            s.run_entity_duplicate_candidates("default", 0.9, 0, cancel=flag)
        assert e.value.code == L.E_CANCELLED
        flag.value = 0
        assert len(s.run_entity_duplicate_candidates("default", 0.9, 0, cancel=flag)) > 0


def test_recall_answers_while_a_join_runs():
    import threading
    from omnia_b200 import FilterSpec
    n, dim = 60_000, 256
    corpus, _ = synth.planted_dups(n, dim, n_pairs=50)
    queries = synth.queries_from(corpus, 64)
    with GPUMemoryStore(dim, n) as s:
        s.bulk_load(corpus)
        want = s.search_rows(queries, [FilterSpec("default")], 10)
        pairs0 = s.run_entity_duplicate_candidates("default", 0.9, 0)
        res = {}

        def join():
            res["pairs"] = [s.run_entity_duplicate_candidates("default", 0.9, 0) for _ in range(3)]

        t = threading.Thread(target=join)
        t.start()
        answers = [s.search_rows(queries, [FilterSpec("default")], 10) for _ in range(20)]
        t.join()
    for g in answers:
        assert all(np.array_equal(x, y) for x, y in zip(want, g))
    assert all([(p.canonical_row, p.dupe_row) for p in ps] == [(p.canonical_row, p.dupe_row) for p in pairs0]
               for ps in res["pairs"])
