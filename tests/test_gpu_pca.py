"""Memory Galaxy projection input: PCAReduce over resident rows against the
oracle's float64 restatement, plus the reference's own shape / clamping cases
(ee/pkg/memory/projection/pca_test.go)."""
import uuid

import numpy as np
import pytest

from omnia_b200 import GPUMemoryStore, FilterSpec, RecallError, SCOPE_WORKSPACE_ID, SCOPE_VIRTUAL_USER_ID
from omnia_b200 import synth
from oracle import oracle as O

pytestmark = pytest.mark.gpu
WS = "ws-pca"


def load(vectors, **kw):
    s = GPUMemoryStore(vectors.shape[1], max(vectors.shape[0], 256), **kw)
    s.bulk_load(vectors, workspaces=[WS], workspace_idx=np.zeros(vectors.shape[0], np.int32))
    return s


def assert_same_up_to_sign(got, want, tol):
    assert got.shape == want.shape
    for c in range(want.shape[1]):
        d = min(np.abs(got[:, c] - want[:, c]).max(), np.abs(got[:, c] + want[:, c]).max())
        assert d <= tol, (c, d)


def low_rank(n, dim, rank, seed):
    """Data with a clearly separated spectrum, so every leading component is well defined."""
    rng = np.random.default_rng(seed)
    basis = np.linalg.qr(rng.standard_normal((dim, rank)))[0]
    scale = np.linspace(3.0, 1.0, rank) ** 2
    x = (rng.standard_normal((n, rank)) * scale) @ basis.T + 0.01 * rng.standard_normal((n, dim)) + 0.05
    return x.astype(np.float32)


@pytest.mark.parametrize("n,dim,k", [(300, 1536, 50), (2000, 384, 50), (257, 200, 7), (64, 768, 3)])
def test_moments_and_projection_match_the_oracle(n, dim, k):
    x = low_rank(n, dim, 12, seed=n + dim)
    with load(x) as s:
        rows = np.arange(n, dtype=np.int64)
        mean, cov = s.pca_moments(rows)
        x64 = x.astype(np.float64)
        np.testing.assert_allclose(mean, x64.mean(axis=0), rtol=0, atol=1e-14)
        want_cov = np.cov(x64, rowvar=False)
        assert np.abs(cov - want_cov).max() <= 1e-12 * max(1.0, np.abs(want_cov).max())
        assert np.array_equal(cov, cov.T)
        again = s.pca_moments(rows)[1]
        assert np.array_equal(cov, again), "fixed-order reduction: same input, same bits"
        got = s.pca_reduce(rows, k)
        want = O.pca_reduce(x64, k)
        kk = min(k, 12)                       # components inside the noise floor have no stable direction
        assert got.shape == (n, min(k, n, dim))
        assert_same_up_to_sign(got[:, :kk], want[:, :kk], 1e-8)
        # the whole k-dimensional image agrees as a subspace: same Gram matrix of the centred scores
        g = got - got.mean(axis=0)
        w = want - want.mean(axis=0)
        assert abs(np.sum(g * g) - np.sum(w * w)) <= 1e-9 * np.sum(w * w)


def test_projection_uses_the_uncentred_rows_and_a_row_subset():
    x = low_rank(500, 256, 6, seed=5)
    with load(x) as s:
        rows = np.arange(0, 500, 3, dtype=np.int64)[::-1].copy()      # any order, any subset
        got = s.pca_reduce(rows, 4)
        want = O.pca_reduce(x[rows].astype(np.float64), 4)
        assert_same_up_to_sign(got, want, 1e-9)
        basis = np.eye(256, 5)
        np.testing.assert_array_equal(s.pca_project(rows, basis), x[rows, :5].astype(np.float64))


def test_reference_shape_and_clamp_cases():
    # pca_test.go:17-31 Shape
    data = (np.arange(6 * 8) * 7 % 13 / 13.0).reshape(6, 8).astype(np.float32)
    with load(data) as s:
        assert s.pca_reduce(np.arange(6), 3).shape == (6, 3)
    # :33-42 ClampsToInputDim
    with load(np.array([[1, 0], [0, 1], [1, 1], [0, 0]], np.float32)) as s:
        assert s.pca_reduce(np.arange(4), 50).shape[1] <= 2
    # :49-71 MoreComponentsThanSamples (#1588): 6 memories, 1536-dim embeddings, 50 requested
    data = (np.arange(6 * 1536) * 7 % 97 / 97.0).reshape(6, 1536).astype(np.float32)
    with load(data) as s:
        out = s.pca_reduce(np.arange(6), 50)
        assert out.shape[0] == 6 and 1 <= out.shape[1] <= 6
        want = O.pca_reduce(data.astype(np.float64), 50)
        assert_same_up_to_sign(out[:, :5], want[:, :5], 1e-8)   # the centred data has rank 5
    # :76-99 DegenerateSets: a single sample is truncated to its first k features, an empty set is an error
    one = (np.arange(64) / 64.0).astype(np.float32)[None, :]
    with load(one) as s:
        out = s.pca_reduce(np.arange(1), 50)
        assert out.shape == (1, 1) and out[0, 0] == float(one[0, 0])
        np.testing.assert_array_equal(O.pca_reduce(one.astype(np.float64), 50), out)
        with pytest.raises(RecallError, match="empty input matrix"):
            s.pca_reduce(np.zeros(0, np.int64), 50)
        with pytest.raises(RecallError, match="outside the store"):
            s.pca_project(np.array([5], np.int64), np.eye(64, 2))


def test_projection_rows_follow_load_projection_inputs():
    # projection_store.go:33-62: one row per live entity in scope, newest active observation
    dim = 64
    with GPUMemoryStore(dim, 256) as s:
        def vec(i):
            v = np.zeros(dim, np.float32); v[i] = 1.0
            return v
        alice = {SCOPE_WORKSPACE_ID: WS, SCOPE_VIRTUAL_USER_ID: "alice"}
        bob = {SCOPE_WORKSPACE_ID: WS, SCOPE_VIRTUAL_USER_ID: "bob"}
        e = [uuid.uuid4() for _ in range(4)]
        o = [uuid.uuid4() for _ in range(6)]
        s.upsert(o[0], e[0], alice, vec(0))        # row 0
        s.upsert(o[1], e[0], alice, vec(1))        # row 1: newer observation of the same entity
        s.upsert(o[2], e[1], alice, vec(2))        # row 2
        s.upsert(o[3], e[2], bob, vec(3))          # row 3
        s.upsert(o[4], e[3], alice, vec(4))        # row 4
        s.upsert(o[5], e[3], alice, vec(5))        # row 5
        f = FilterSpec(WS, now_us=0)
        assert s.projection_rows(f).tolist() == [1, 2, 3, 5]
        assert s.projection_rows(FilterSpec(WS, "alice", now_us=0)).tolist() == [1, 2, 5]
        s.supersede(o[5])                          # the older observation becomes the entity's active one
        assert s.projection_rows(f).tolist() == [1, 2, 3, 4]
        s.forget(e[1])
        assert s.projection_rows(f).tolist() == [1, 3, 4]
        assert s.projection_rows(f, cap=2).tolist() == [1, 3]
        assert s.projection_rows(FilterSpec("elsewhere", now_us=0)).tolist() == []


def test_galaxy_sized_input_on_the_seeder_like_corpus():
    # default Cap = 2000 entities (internal/memory/projection/projection.go:85), PCADims = 50, D = 1536
    x = synth.clustered(2000, 1536, seed=1)
    with load(x) as s:
        rows = s.projection_rows(FilterSpec(WS, now_us=0))
        assert rows.shape[0] == 2000
        got = s.pca_reduce(rows, 50)
        want = O.pca_reduce(x.astype(np.float64), 50)
        assert got.shape == want.shape == (2000, 50)
        # six topic centroids: the five leading components separate them and are well conditioned
        assert_same_up_to_sign(got[:, :5], want[:, :5], 1e-8)
        g = got - got.mean(axis=0)
        w = want - want.mean(axis=0)
        assert abs(np.sum(g * g) - np.sum(w * w)) <= 1e-9 * np.sum(w * w)
