"""The completeness argument of the emission mode as an executable model, no GPU.

The model follows the device protocol step by step -- seed threshold from the top-2-per-32-row keys of the first
tiles, tiles scored in an arbitrary interleaving with the threshold read once per tile, a refresh from the newest
list entries every `trigger` appends, the finishing band filter and exact rescore -- with approximate scores that
are off by up to eps in the WORST direction for each row (answer rows pushed down, the rest pushed up).  The claim
under test: whatever the interleaving, the rows that survive to the exact rescore contain the exact top-k, so the
answer equals the exact ORDER BY ... LIMIT k."""
import numpy as np
import pytest

TILE = 256


def kth_largest(v, k):
    return np.partition(v, len(v) - k)[len(v) - k]


def run_model(exact, k, eps, seed_tiles, trigger, rng, adversarial=True):
    n = len(exact)
    order_exact = np.lexsort((np.arange(n), -exact))          # ORDER BY score DESC, row
    answer = order_exact[:k]
    err = rng.uniform(-eps, eps, n)
    if adversarial:
        err[:] = eps                                            # everything pushed up ...
        err[answer] = -eps                                      # ... except the answer rows, pushed down
    approx = (exact + err).astype(np.float32)
    band = np.float32(2 * eps)

    # seed: top-2 of every 32-row sub-bucket of the first tiles; k distinct rows score at least the k-th largest key
    m = min(n, seed_tiles * TILE)
    keys = []
    for s in range(0, m, 32):
        part = np.sort(approx[s:s + 32])[::-1][:2]
        keys.extend(part.tolist())
    keys = np.asarray(keys, np.float32)
    thr = -np.inf if len(keys) < k else kth_largest(keys, k) - band

    # score and emit: tiles in an arbitrary order, threshold read once per tile, refresh every `trigger` appends
    listed = []
    since = 0
    tiles = rng.permutation((n + TILE - 1) // TILE)
    for t in tiles:
        t_thr = thr                                             # begin_tile
        lo, hi = t * TILE, min(n, (t + 1) * TILE)
        rows = np.nonzero(approx[lo:hi] >= t_thr)[0] + lo
        for r in rows:
            listed.append(r)
            since += 1
            if since >= trigger or len(listed) == k:
                since = 0
                newest = np.asarray(listed[-768:])
                if len(newest) >= k:
                    thr = max(thr, kth_largest(approx[newest], k) - band)   # atomicMax: only ever raised
    listed = np.asarray(listed, np.int64)
    assert len(np.unique(listed)) == len(listed)

    # finish: T from the list, band filter, exact rescore, ORDER BY (score DESC, row) LIMIT k
    assert len(listed) >= k
    T = kth_largest(approx[listed], k)
    keep = listed[approx[listed] >= T - band]
    got = keep[np.lexsort((keep, -exact[keep]))][:k]
    return answer, got, len(listed), len(keep)


@pytest.mark.parametrize("k", [1, 10, 100])
@pytest.mark.parametrize("shape", ["gaussian", "near_ties", "few_values"])
def test_answer_survives_any_interleaving_and_worst_case_error(k, shape):
    rng = np.random.default_rng(1234 + k)
    n, eps = 40_000, 4.5e-3
    if shape == "gaussian":
        exact = rng.normal(0, 0.03, n)
    elif shape == "near_ties":                                 # hundreds of rows within eps of the k-th place
        exact = rng.normal(0, 0.03, n)
        exact[rng.choice(n, 600, replace=False)] = 0.2 + rng.uniform(-eps, eps, 600)
    else:                                                      # low-entropy corpus: a handful of distinct scores
        exact = rng.choice(np.linspace(-0.1, 0.3, 9), n)
    exact = exact.astype(np.float64)
    for trial in range(4):
        answer, got, listed, kept = run_model(exact, k, eps, seed_tiles=max(2, (k + 7) // 8 + 1), trigger=max(128, 4 * k),
                                              rng=rng, adversarial=trial % 2 == 0)
        if shape == "gaussian":
            assert listed < n // 10, "the running threshold is supposed to keep the lists short"
        # ids and order exact except where two exact scores tie (SQL leaves tie order open; the engine uses row order)
        assert np.array_equal(exact[got], exact[answer])
        untied = np.ones(k, bool)
        if k > 1:
            d = np.diff(exact[answer])
            untied[:-1] &= d != 0
            untied[1:] &= d != 0
        kth = exact[answer[-1]]
        untied &= ~((exact[answer] == kth) & (np.count_nonzero(exact == kth) > np.count_nonzero(exact[answer] == kth)))
        assert np.array_equal(got[untied], answer[untied])


def test_a_margin_smaller_than_the_error_does_lose_rows():
    """The guard for the guard: with a band narrower than the real error the model must be able to fail, otherwise the
    test above proves nothing."""
    rng = np.random.default_rng(7)
    n, k, eps = 40_000, 10, 4.5e-3
    exact = rng.normal(0, 0.03, n)
    exact[rng.choice(n, 600, replace=False)] = 0.2 + rng.uniform(-eps, eps, 600)
    lost = 0
    for _ in range(6):
        order = np.lexsort((np.arange(n), -exact))
        answer = order[:k]
        err = np.full(n, eps)
        err[answer] = -eps
        approx = (exact + err).astype(np.float32)
        band = np.float32(2 * (eps / 8))                        # claims a bound eight times too small
        T = kth_largest(approx, k)
        keep = np.nonzero(approx >= T - band)[0]
        got = keep[np.lexsort((keep, -exact[keep]))][:k]
        lost += not np.array_equal(exact[got], exact[answer])
    assert lost > 0
