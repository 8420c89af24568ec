"""N > 1 host logic on CPU (gloo, world_size 2 and 3): round-robin row sharding, the chunk-seeded
corpus every rank can cut its own rows from, the allgather layout [G][nq][k], and the global-row
arithmetic.  The local search and the merge are done by the oracle / numpy here (this file is the
checker); on GPUs the same plumbing carries omr_device_search and omr_device_merge_topk."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from omnia_b200 import synth
from omnia_b200.dist import gather_candidates, shard_of
from oracle import oracle as O

N, DIM, NQ, K = 70_000, 48, 6, 10      # spans two 65536-row corpus chunks


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        shard = synth.gauss_unit(N, DIM, rank=rank, world=world)
        n_local, base, stride = shard_of(N, rank, world)
        assert shard.shape[0] == n_local
        full = synth.gauss_unit(N, DIM)
        queries = synth.queries_from(full, NQ)
        assert np.array_equal(shard, full[rank::world])
        rows, d, _ = O.search(shard, queries, K, variant=O.VARIANT_F64)
        rows = np.where(rows >= 0, base + rows * stride, -1)          # local -> global row ids
        gr, gd = gather_candidates(torch.from_numpy(rows), torch.from_numpy(d))
        assert gr.shape == (world, NQ, K) and torch.equal(gr[rank], torch.from_numpy(rows))
        if rank == 0:
            out.put((gr.numpy(), gd.numpy()))
        dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_sharded_recall_plumbing_over_gloo(world):
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, out)) for r in range(world)]
    for p in procs:
        p.start()
    gr, gd = out.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    # merge G*k -> k by (distance, row), the order merge_topk_kernel uses
    full = synth.gauss_unit(N, DIM)
    queries = synth.queries_from(full, NQ)
    want_r, want_d, _ = O.search(full, queries, K, variant=O.VARIANT_F64)
    for q in range(NQ):
        r, d = gr[:, q, :].reshape(-1), gd[:, q, :].reshape(-1)
        o = np.lexsort((r, d))[:K]
        assert np.array_equal(r[o], want_r[q]) and np.allclose(d[o], want_d[q], rtol=0, atol=1e-12)


def test_shard_of_covers_every_row_once():
    for n in (0, 1, 7, 1000, 65537):
        for world in (1, 2, 3, 8):
            rows = []
            for r in range(world):
                cnt, base, stride = shard_of(n, r, world)
                rows += [base + i * stride for i in range(cnt)]
            assert sorted(rows) == list(range(n))


class _StubJoinStore:
    """Stands in for GPUMemoryStore in the self-join plumbing test: its pairs are the oracle's, cut into parts the
    way the kernel's interleaved work items cut them (any disjoint cover will do for the host logic)."""

    def __init__(self, pairs):
        self.pairs = pairs

    def run_entity_duplicate_candidates(self, workspace_id, similarity_floor, max_buckets_per_pass, part_index=0,
                                        part_count=1, **kw):
        mine = [p for i, p in enumerate(self.pairs) if i % part_count == part_index]
        mine.sort(key=lambda p: (-p.similarity, p.canonical_row, p.dupe_row))
        return mine[:max_buckets_per_pass] if max_buckets_per_pass > 0 else mine

    def selfjoin_rows(self, workspace_id, similarity_floor, limit, part_index=0, part_count=1, **kw):
        mine = self.run_entity_duplicate_candidates(workspace_id, similarity_floor, limit, part_index, part_count)
        return (np.array([p.canonical_row for p in mine], np.int64), np.array([p.dupe_row for p in mine], np.int64),
                np.array([p.similarity for p in mine], np.float64))

    def pairs_from_rows(self, c, d, v):
        from omnia_b200.store import DuplicatePair
        return [DuplicatePair(f"e{int(x)}", f"e{int(y)}", float(z), "", int(x), int(y)) for x, y, z in zip(c, d, v)]


def _join_worker(rank, world, port, limit, out):
    from omnia_b200.dist import sharded_selfjoin
    from omnia_b200.store import DuplicatePair
    from helpers import ids_from_rows
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        n, dim = 1500, 64
        corpus, _ = synth.planted_dups(n, dim, n_pairs=60)
        a, b, s = O.selfjoin(corpus, None, np.zeros(n, np.uint16), np.full(n, 12, np.uint8), ids_from_rows(n), 0.9,
                             variant=O.VARIANT_F64)
        pairs = [DuplicatePair(f"e{int(x)}", f"e{int(y)}", float(v), "", int(x), int(y)) for x, y, v in zip(a, b, s)]
        rng = np.random.default_rng(5)
        rng.shuffle(pairs)                                   # the parts must not rely on any input order
        got = sharded_selfjoin(_StubJoinStore(pairs), "default", 0.9, limit)
        want = sorted(pairs, key=lambda p: (-p.similarity, p.canonical_row, p.dupe_row))
        want = want[:limit] if limit > 0 else want
        assert [(p.canonical_row, p.dupe_row) for p in got] == [(p.canonical_row, p.dupe_row) for p in want]
        if rank == 0:
            out.put(len(got))
        dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,limit", [(2, 0), (3, 7)])
def test_sharded_selfjoin_merges_the_parts_over_gloo(world, limit):
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_join_worker, args=(r, world, port, limit, out)) for r in range(world)]
    for p in procs:
        p.start()
    n = out.get(timeout=120)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    assert n == (7 if limit else n) and n > 0
