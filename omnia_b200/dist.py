"""Row-sharded recall across the GPUs of one box.

One process per GPU.  Rank r holds rows r, r+G, r+2G, ... of the store (round-robin, so every
workspace spreads evenly), answers each query batch with its exact local top-k, and the per-rank
answers are exchanged with ONE allgather of k x (float64 distance, int64 global row) per query over
NCCL/NVLink; every rank then merges G*k -> k with the library's merge kernel.  The global top-k is
exact because each shard returns its full local top-k.

torch.distributed is plumbing here (process group + the collective); the scoring, selection and
merge are the library's kernels.
"""
from __future__ import annotations

from typing import Optional, Tuple

import numpy as np
import torch
import torch.distributed as dist

from .store import FilterSpec, GPUMemoryStore


def shard_of(n_rows: int, rank: int, world: int) -> Tuple[int, int, int]:
    """(rows on this rank, row_base, row_stride) for round-robin sharding of n_rows."""
    return (n_rows - rank + world - 1) // world, rank, world


def gather_candidates(rows: torch.Tensor, dist_: torch.Tensor, group=None) -> Tuple[torch.Tensor, torch.Tensor]:
    """Allgather of per-rank [nq, k] answers into [G, nq, k] (works on any backend; gloo in CPU tests)."""
    world = dist.get_world_size(group)
    out_r = torch.empty((world,) + tuple(rows.shape), dtype=rows.dtype, device=rows.device)
    out_d = torch.empty((world,) + tuple(dist_.shape), dtype=dist_.dtype, device=dist_.device)
    # list-of-views form: every backend accepts it (gloo rejects the stacked into_tensor layout)
    dist.all_gather(list(out_r.unbind(0)), rows.contiguous(), group=group)
    dist.all_gather(list(out_d.unbind(0)), dist_.contiguous(), group=group)
    return out_r, out_d


class ShardedRecall:
    """The local shard plus the exchange step.  With no process group it is a single shard."""

    def __init__(self, dim: int, n_rows_total: int, device: int, group=None, exchange: str = "nccl",
                 nq_max: int = 1024, k_max: int = 128, slots: int = 9, **store_kw):
        self.group = group
        live = dist.is_available() and dist.is_initialized()
        self.rank = dist.get_rank(group) if live else 0
        self.world = dist.get_world_size(group) if live else 1
        n_local, base, stride = shard_of(n_rows_total, self.rank, self.world)
        self.store = GPUMemoryStore(dim, max(1, n_local), device=device, row_base=base, row_stride=stride, **store_kw)
        self.device = torch.device("cuda", device)
        self._buf = {}
        # exchange + merge of a finished batch run on their own stream, so that the next batch's scorer
        # (ordered after the caller's stream at begin()) never waits behind a collective
        self._comm = torch.cuda.Stream(device=self.device) if torch.cuda.is_available() else None
        self._slot_ev = {}   # slot -> event of the last exchange that read / wrote the slot's buffers
        # exchange = "p2p": the library's fused exchange (peer-memory writes + on-device flags) instead of
        # an NCCL allgather; the process group is then only used once, to swap the 64-byte IPC handles
        self.exchange = exchange if self.world > 1 else "none"
        self.slots = slots
        if self.exchange == "p2p":
            self._setup_p2p(nq_max, k_max, slots)

    def _setup_p2p(self, nq_max: int, k_max: int, slots: int) -> None:
        import ctypes as C
        from . import _lib as L
        lib = L.load()
        buf, nbytes = C.c_void_p(), C.c_int64(0)
        L.check(lib.omr_exchange_alloc(self.store.handle, self.world, nq_max, k_max, slots, C.byref(buf), C.byref(nbytes)))
        handle = (C.c_uint8 * 64)()
        L.check(lib.omr_ipc_export(buf, handle))
        gathered = [None] * self.world
        dist.all_gather_object(gathered, bytes(handle), group=self.group)
        blob = (C.c_uint8 * (64 * self.world)).from_buffer_copy(b"".join(gathered))
        L.check(lib.omr_exchange_setup(self.store.handle, self.rank, self.world, blob))
        dist.barrier(group=self.group)   # nobody writes into a peer before every mapping exists

    def _buffers(self, nq: int, k: int, slot: int = 0):
        key = (nq, k, slot)
        if key not in self._buf:
            d = self.device
            # rows and distances of one rank live in ONE buffer [2][nq][k] (int64 / float64 views), so the
            # exchange is a single allgather into [G][2][nq][k]
            local = torch.empty((2, nq, k), dtype=torch.int64, device=d)
            gathered = torch.empty((self.world, 2, nq, k), dtype=torch.int64, device=d)
            self._buf[key] = dict(
                local=local, lr=local[0], ld=local[1].view(torch.float64),
                lc=torch.empty(nq, dtype=torch.int32, device=d), gathered=gathered,
                r=torch.empty((nq, k), dtype=torch.int64, device=d), d=torch.empty((nq, k), dtype=torch.float64, device=d),
                c=torch.empty(nq, dtype=torch.int32, device=d))
        return self._buf[key]

    def search_device(self, d_queries: torch.Tensor, k: int, d_bitmaps=None, d_query_group=None, d_max_distance=None):
        """Global top-k for device-resident queries; returns (rows, dist, count) device tensors."""
        nq = d_queries.shape[0]
        if self.exchange == "p2p":
            self._sync_slot = (getattr(self, "_sync_slot", -1) + 1) % self.slots
            r, d, c, _ = self.finish(self.begin(d_queries, k, self._sync_slot, d_bitmaps, d_query_group, d_max_distance))
            return r, d, c
        b = self._buffers(nq, k)
        stream = torch.cuda.current_stream(self.device).cuda_stream
        self.store.device_search(d_queries, k, b["lr"], b["ld"], b["lc"], d_bitmaps, 0, d_query_group,
                                 d_max_distance, stream)
        if self.world == 1:
            return b["lr"], b["ld"], b["lc"]
        dist.all_gather_into_tensor(b["gathered"], b["local"], group=self.group)
        g = b["gathered"]
        self.store.device_merge_topk(g[:, 0], g[:, 1].view(torch.float64), self.world, nq, k, b["r"], b["d"], b["c"],
                                     stream, shard_stride=2 * nq * k)
        return b["r"], b["d"], b["c"]

    # Pipelined form: begin() enqueues the local search of a batch and returns; finish() waits for it,
    # then exchanges and merges on the communication stream.  With two batches in flight the select /
    # rescore tail, the allgather and the merge of batch i all run underneath the scorer of batch i+1.
    # `slot` names a set of result buffers; rotate through >= 4 slots so a new batch never has to wait
    # for the exchange that last used its buffers.
    def begin(self, d_queries: torch.Tensor, k: int, slot: int, d_bitmaps=None, d_query_group=None,
              d_max_distance=None):
        nq = d_queries.shape[0]
        b = self._buffers(nq, k, slot)
        cur = torch.cuda.current_stream(self.device)
        if self.exchange == "p2p":
            import ctypes as C
            from . import _lib as L
            p = lambda t: None if t is None else t.data_ptr()
            ticket = C.c_int32(0)
            L.check(L.load().omr_device_search_exchange_begin(
                self.store.handle, d_queries.data_ptr(), nq, p(d_bitmaps), 0, p(d_query_group), p(d_max_distance), k,
                slot % self.slots, cur.cuda_stream, C.byref(ticket)))
            return (ticket.value, nq, k, slot)
        if slot in self._slot_ev:               # the slot's previous exchange must be done with its buffers
            cur.wait_event(self._slot_ev[slot])
        stream = cur.cuda_stream
        ticket = self.store.device_search_begin(d_queries, k, b["lr"], b["ld"], b["lc"], d_bitmaps, 0, d_query_group,
                                                d_max_distance, stream)
        return (ticket, nq, k, slot)

    def finish(self, handle):
        """Completes a batch started with begin().  Returns (rows, dist, count, event): with more than
        one shard the exchange and merge are enqueued on the store's communication stream and `event`
        marks their completion (wait on it, or synchronise the device, before reading the tensors)."""
        ticket, nq, k, slot = handle
        b = self._buffers(nq, k, slot)
        cur = torch.cuda.current_stream(self.device)
        if self.exchange == "p2p":
            from . import _lib as L
            L.check(L.load().omr_device_search_exchange_finish(self.store.handle, ticket, b["r"].data_ptr(),
                                                               b["d"].data_ptr(), b["c"].data_ptr()))
            ev = torch.cuda.Event()
            ev.record(cur)
            return b["r"], b["d"], b["c"], ev
        self.store.device_search_finish(ticket)
        if self.world == 1:
            ev = torch.cuda.Event()
            ev.record(cur)
            return b["lr"], b["ld"], b["lc"], ev
        self._comm.wait_stream(cur)            # the local answer is ordered on `cur` by device_search_finish
        with torch.cuda.stream(self._comm):
            dist.all_gather_into_tensor(b["gathered"], b["local"], group=self.group)
            g = b["gathered"]
            self.store.device_merge_topk(g[:, 0], g[:, 1].view(torch.float64), self.world, nq, k, b["r"], b["d"],
                                         b["c"], self._comm.cuda_stream, shard_stride=2 * nq * k)
            ev = torch.cuda.Event()
            ev.record(self._comm)
        self._slot_ev[slot] = ev
        return b["r"], b["d"], b["c"], ev

    def close(self):
        self.store.close()


def sharded_selfjoin(store: GPUMemoryStore, workspace_id: str, similarity_floor: float, max_buckets_per_pass: int = 100,
                     group=None, device=None, **kw):
    """Near-duplicate self-join block-partitioned over the ranks: every rank holds the
    whole matrix (C4's 3 GB bf16 replicate easily) and scans part `rank` of `world` of the tile-pair space; with
    a LIMIT each rank already returns only its n best pairs (omr_selfjoin_rows selects on the device).  The
    variable-length lists are exchanged as tensors -- one all_gather of the counts, one of the payload padded to
    the largest count, (canonical row, dupe row, similarity) -- on the process group's device (NCCL over NVLink
    on GPUs, gloo in the CPU tests), and merged: ORDER BY similarity DESC (ties by rows), LIMIT n.  Every rank
    returns the same list."""
    live = dist.is_available() and dist.is_initialized()
    rank = dist.get_rank(group) if live else 0
    world = dist.get_world_size(group) if live else 1
    c, d, v = store.selfjoin_rows(workspace_id, similarity_floor, max_buckets_per_pass, part_index=rank, part_count=world,
                                  **kw)
    if world > 1:
        if device is None:
            device = torch.device("cuda", torch.cuda.current_device()) if dist.get_backend(group) == "nccl" else torch.device("cpu")
        count = torch.tensor([len(c)], dtype=torch.int64, device=device)
        counts = [torch.zeros_like(count) for _ in range(world)]
        dist.all_gather(counts, count, group=group)
        counts = [int(t.item()) for t in counts]
        width = max(1, max(counts))
        # one int64 payload [3][width]: rows as they are, similarities by their bit pattern
        mine = torch.zeros((3, width), dtype=torch.int64)
        mine[0, :len(c)] = torch.from_numpy(np.ascontiguousarray(c))
        mine[1, :len(d)] = torch.from_numpy(np.ascontiguousarray(d))
        mine[2, :len(v)] = torch.from_numpy(np.ascontiguousarray(v).view(np.int64))
        mine = mine.to(device)
        parts = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(parts, mine, group=group)
        parts = [p.cpu().numpy() for p in parts]
        c = np.concatenate([p[0, :n] for p, n in zip(parts, counts)])
        d = np.concatenate([p[1, :n] for p, n in zip(parts, counts)])
        v = np.concatenate([p[2, :n].view(np.float64) for p, n in zip(parts, counts)])
        order = np.lexsort((d, c, -v))
        if max_buckets_per_pass > 0:
            order = order[:max_buckets_per_pass]
        c, d, v = c[order], d[order], v[order]
    return store.pairs_from_rows(c, d, v)
