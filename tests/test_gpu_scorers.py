"""Each approximate scorer in isolation: the packed top-2 keys it emits per (query, sub-bucket)
against a numpy restatement of the same bf16 arithmetic.  This is the test that localises a
tcgen05 / TMA descriptor mistake to a tile, a row or a K block."""
import numpy as np
import pytest
import torch

from omnia_b200 import GPUMemoryStore, synth

pytestmark = pytest.mark.gpu


def bf16_round(x):
    return torch.from_numpy(np.ascontiguousarray(x)).to(torch.bfloat16).to(torch.float32).numpy()


def key_to_score_col(keys):
    k = keys.astype(np.int64)
    col = k & 0xFF
    o = (k & ~0xFF).astype(np.int32)
    b = o ^ ((o >> 31) & 0x7FFFFFFF)
    return b.view(np.float32), col


@pytest.mark.parametrize("path", [1, 2, 4])
@pytest.mark.parametrize("n,dim,nq,S", [(2048, 128, 8, 256), (5000, 1536, 8, 32), (1000, 200, 5, 64),
                                        (256 * 40, 384, 8, 128)])
def test_scorer_keys_match_bf16_arithmetic(path, n, dim, nq, S):
    if path != 1:
        nq = max(nq, 8)
    corpus = synth.gauss_unit(n, dim)
    queries = synth.queries_from(corpus, nq)
    with GPUMemoryStore(dim, n) as s:
        s.bulk_load(corpus)
        keys = s.debug_candidate_keys(queries, path, S)
    rb, qb = bf16_round(corpus), bf16_round(queries)     # inputs are already unit-norm
    scores = qb.astype(np.float64) @ rb.astype(np.float64).T
    n_sb = keys.shape[1] // 2
    pad = n_sb * S - n
    sc = np.pad(scores, ((0, 0), (0, pad)), constant_values=-np.inf).reshape(nq, n_sb, S)
    order = np.argsort(-sc, axis=2, kind="stable")
    for rank in (0, 1):
        got_s, got_c = key_to_score_col(keys[:, rank::2])
        want_c = order[:, :, rank]
        want_s = np.take_along_axis(sc, want_c[:, :, None], 2)[:, :, 0]
        valid = np.isfinite(want_s)
        assert (keys[:, rank::2][~valid] == np.iinfo(np.int32).min).all()
        err = np.abs(got_s[valid] - want_s[valid])
        assert err.max() < 2e-4, f"rank {rank}: score error {err.max():.3e} at {np.argwhere(np.abs(np.where(valid, got_s - want_s, 0)) > 2e-4)[:5]}"
        # the column may differ only when two scores tie within the key's precision
        mism = valid & (got_c != want_c)
        if mism.any():
            alt = np.take_along_axis(sc, np.clip(got_c, 0, S - 1)[:, :, None], 2)[:, :, 0]
            assert np.abs(alt[mism] - want_s[mism]).max() < 2e-4
