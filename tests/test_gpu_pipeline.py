"""omr_device_search_begin / _finish: two batches in flight must give exactly the answers of the
synchronous entry point, including batches that need the exact fallback or the unfused chain."""
import numpy as np
import pytest
import torch

from omnia_b200 import GPUMemoryStore, RecallError, synth

pytestmark = pytest.mark.gpu


def outs(nq, k):
    return (torch.empty((nq, k), dtype=torch.int64, device="cuda"), torch.empty((nq, k), dtype=torch.float64, device="cuda"),
            torch.empty(nq, dtype=torch.int32, device="cuda"))


@pytest.mark.parametrize("n,dim,nq,k,path", [(30_000, 256, 200, 10, 0), (30_000, 256, 5, 10, 0), (8_000, 128, 40, 200, 0),
                                             (5_000, 64, 16, 10, 3), (20_000, 384, 130, 100, 2)])
def test_pipelined_equals_synchronous(n, dim, nq, k, path):
    corpus = synth.clustered(n, dim) if k == 10 and path == 0 else synth.gauss_unit(n, dim)
    batches = [torch.from_numpy(synth.queries_from(corpus, nq, seed=s)).cuda() for s in range(5)]
    with GPUMemoryStore(dim, n, force_path=path) as s:
        s.bulk_load(corpus)
        want = []
        for b in batches:
            o = outs(nq, k)
            s.device_search(b, k, *o)
            torch.cuda.synchronize()
            want.append([t.cpu().numpy().copy() for t in o])
        got, o2 = [], [outs(nq, k), outs(nq, k)]
        stream = torch.cuda.current_stream().cuda_stream
        pending = None
        for i, b in enumerate(batches):
            t = s.device_search_begin(b, k, *o2[i & 1], stream=stream)
            if pending is not None:
                s.device_search_finish(pending[0])
                torch.cuda.synchronize()
                got.append([x.cpu().numpy().copy() for x in o2[pending[1]]])
            pending = (t, i & 1)
        s.device_search_finish(pending[0])
        torch.cuda.synchronize()
        got.append([x.cpu().numpy().copy() for x in o2[pending[1]]])
        for w, g in zip(want, got):
            assert np.array_equal(w[2], g[2]) and np.array_equal(w[0], g[0]) and np.array_equal(w[1], g[1])
        # a third begin without a finish is refused, and a stale ticket too
        keep = [outs(nq, k) for _ in range(5)]          # output buffers must outlive their batch
        tickets = [s.device_search_begin(batches[i % 5], k, *keep[i], stream=stream) for i in range(4)]
        with pytest.raises(RecallError, match="already in flight"):
            s.device_search_begin(batches[2], k, *keep[4], stream=stream)
        for t in tickets:
            s.device_search_finish(t)
        with pytest.raises(RecallError, match="not in flight"):
            s.device_search_finish(tickets[0])


def test_host_buffer_pipeline_equals_synchronous_call():
    """omr_search_rows_begin/_finish with filters, from pageable and from page-locked query buffers."""
    from omnia_b200 import FilterSpec
    n, dim, nq, k = 40_000, 128, 96, 10
    corpus = synth.gauss_unit(n, dim)
    ws = (np.arange(n) % 2).astype(np.uint32)
    batches = [synth.queries_from(corpus, nq, seed=s) for s in range(6)]
    pinned = [torch.from_numpy(b).pin_memory().numpy() for b in batches]
    flt = [FilterSpec("w0"), FilterSpec("w1", max_distance=0.8)]
    qf = (np.arange(nq) % 2).astype(np.int32)
    with GPUMemoryStore(dim, n) as s:
        s.bulk_load(corpus, workspaces=["w0", "w1"], workspace_idx=ws)
        want = [s.search_rows(b, flt, k, qf) for b in batches]
        for src in (batches, pinned):
            outs = [(np.empty((nq, k), np.int64), np.empty((nq, k), np.float64), np.empty(nq, np.int32)) for _ in src]
            pend = []
            for i, b in enumerate(src):
                pend.append((s.search_rows_begin(b, flt, k, qf), i))
                if len(pend) == 3:
                    tkt, j = pend.pop(0)
                    s.search_rows_finish(tkt, outs[j])
            for tkt, j in pend:
                s.search_rows_finish(tkt, outs[j])
            for w, g in zip(want, outs):
                assert all(np.array_equal(a, b) for a, b in zip(w, g))
        with pytest.raises(RecallError, match="different vector dimensions"):
            s.search_rows_begin(np.zeros((2, 3), np.float32), flt, k)
        with pytest.raises(RecallError, match="workspace_id is required"):
            s.search_rows_begin(batches[0], [FilterSpec("")], k)
