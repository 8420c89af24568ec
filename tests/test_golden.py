"""The committed golden vectors (tests/golden/recall_golden.json, made by tests/golden/make_golden.py)
against the CPU oracle: closed-form cases of the reference's own tests for every oracle variant, and the
frozen answers on seeded inputs.  The CUDA engine meets the same file in tests/test_gpu_golden.py."""
import hashlib
import json
import os

import numpy as np
import pytest

from omnia_b200 import synth
from oracle import oracle as O
from helpers import assert_topk_matches, ids_from_rows

GOLDEN = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "recall_golden.json")))


def build_vec(spec, dim):
    v = np.zeros(dim, np.float32)
    if spec[0] == "const":
        v[:] = spec[1]
    elif spec[0] == "onehot":
        v[spec[1]] = 1.0
    elif spec[0] == "twohot":
        v[spec[1]] = 1.0
        v[spec[2]] = 1.0
    return v


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def golden_inputs(case):
    corpus = synth.gauss_unit(case["n"], case["dim"])
    queries = synth.queries_from(corpus, case["nq"])
    if sha(corpus) != case["corpus_sha256"] or sha(queries) != case["queries_sha256"]:
        pytest.skip("numpy's generator produced different bytes than when the golden file was made")
    return corpus, queries


def golden_answers(case):
    rows = np.array(case["rows"], np.int64)
    dist = np.array([[float.fromhex(h) for h in r] for r in case["dist_hex"]], np.float64)
    return rows, dist, np.array(case["count"], np.int32)


@pytest.mark.parametrize("variant", [O.VARIANT_F32_SEQ, O.VARIANT_F64, O.VARIANT_F32_VEC])
def test_closed_form_cases_of_the_reference_tests(variant):
    for c in GOLDEN["closed_form"]:
        a, b = build_vec(c["a"], c["dim"]), build_vec(c["b"], c["dim"])
        d = O.cosine_distance(a, b, variant)
        tol = max(c["atol"], 1e-12) if variant == O.VARIANT_F64 else max(c["atol"], 1e-6)
        assert abs(d - c["distance"]) <= tol, (c["ref"], d, c["distance"])


@pytest.mark.parametrize("case", GOLDEN["cases"], ids=lambda c: c["name"])
def test_oracle_reproduces_the_frozen_answers(case):
    corpus, queries = golden_inputs(case)
    md = float("inf") if case["max_distance"] is None else case["max_distance"]
    r_rows, r_dist, r_cnt = O.search(corpus, queries, case["k"], max_distance=md, variant=O.VARIANT_F64)
    g_rows, g_dist, g_cnt = golden_answers(case)
    assert_topk_matches(r_rows, r_dist, r_cnt, g_rows, g_dist, g_cnt, tie=1e-10, rtol=1e-11, what=case["name"])


def test_oracle_reproduces_the_frozen_pair_set():
    sj = GOLDEN["selfjoin"]
    corpus, _ = synth.planted_dups(sj["n"], sj["dim"], n_pairs=sj["n_pairs"])
    if sha(corpus) != sj["corpus_sha256"]:
        pytest.skip("numpy's generator produced different bytes than when the golden file was made")
    n = sj["n"]
    a, b, s = O.selfjoin(corpus, None, (np.arange(n) % 3).astype(np.uint16), np.full(n, 12, np.uint8),
                         ids_from_rows(n), sj["floor"], require_same_kind=False, variant=O.VARIANT_F64)
    assert [[int(x), int(y)] for x, y in zip(a, b)] == sj["pairs"]
    want = np.array([float.fromhex(h) for h in sj["sim_hex"]])
    np.testing.assert_allclose(s, want, rtol=0, atol=1e-12)
