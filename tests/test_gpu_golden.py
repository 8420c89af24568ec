"""The CUDA engine against the committed golden vectors (tests/golden/recall_golden.json)."""
import uuid

import numpy as np
import pytest

from omnia_b200 import GPUMemoryStore, FilterSpec, SCOPE_WORKSPACE_ID
from omnia_b200 import synth
from helpers import assert_topk_matches
from test_golden import GOLDEN, build_vec, golden_answers, golden_inputs, sha

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("path", [0, 1, 2, 3])
def test_closed_form_cases_through_the_engine(path):
    for c in GOLDEN["closed_form"]:
        with GPUMemoryStore(c["dim"], 256, force_path=path) as s:
            scope = {SCOPE_WORKSPACE_ID: "w"}
            oid = uuid.uuid4()
            s.upsert(oid, uuid.uuid4(), scope, build_vec(c["b"], c["dim"]))
            got = s.find_similar_observations(scope, build_vec(c["a"], c["dim"]), 5, -1.0)
            assert len(got) == 1 and got[0].observation_id == str(oid)
            assert abs((1.0 - got[0].similarity) - c["distance"]) <= max(c["atol"], 1e-12), c["ref"]


@pytest.mark.parametrize("path", [0, 1, 2, 3])
@pytest.mark.parametrize("case", GOLDEN["cases"], ids=lambda c: c["name"])
def test_engine_matches_the_frozen_answers(case, path):
    corpus, queries = golden_inputs(case)
    md = float("nan") if case["max_distance"] is None else case["max_distance"]
    with GPUMemoryStore(case["dim"], case["n"], force_path=path) as s:
        s.bulk_load(corpus)
        rows, dist, cnt, _ = s.search(queries, [FilterSpec("default", max_distance=md)], case["k"])
    g_rows, g_dist, g_cnt = golden_answers(case)
    assert_topk_matches(rows, dist, cnt, g_rows, g_dist, g_cnt, corpus, queries, what=f"{case['name']} path={path}")


def test_engine_matches_the_frozen_pair_set():
    sj = GOLDEN["selfjoin"]
    corpus, _ = synth.planted_dups(sj["n"], sj["dim"], n_pairs=sj["n_pairs"])
    if sha(corpus) != sj["corpus_sha256"]:
        pytest.skip("numpy's generator produced different bytes than when the golden file was made")
    with GPUMemoryStore(sj["dim"], sj["n"]) as s:
        s.bulk_load(corpus)
        pairs = s.run_entity_duplicate_candidates("default", sj["floor"], 0, require_same_kind=False)
    got = sorted((min(p.canonical_row, p.dupe_row), max(p.canonical_row, p.dupe_row)) for p in pairs)
    assert got == sorted((min(a, b), max(a, b)) for a, b in sj["pairs"])
    want = {(min(a, b), max(a, b)): float.fromhex(h) for (a, b), h in zip(sj["pairs"], sj["sim_hex"])}
    for p in pairs:
        assert abs(p.similarity - want[(min(p.canonical_row, p.dupe_row), max(p.canonical_row, p.dupe_row))]) <= 1e-9
