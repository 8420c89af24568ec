"""Row-sharded recall on ONE GPU: G shard stores emulate G ranks, their local answers are
concatenated the way an allgather would deliver them and merged by omr_device_merge_topk."""
import numpy as np
import pytest
import torch

from omnia_b200 import GPUMemoryStore, synth
from oracle import oracle as O
from helpers import assert_topk_matches

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("G,k,nq", [(2, 10, 16), (4, 100, 130), (8, 10, 3)])
def test_sharded_local_topk_plus_merge_equals_global(G, k, nq):
    n, dim = 12_000, 128
    corpus = synth.gauss_unit(n, dim)
    queries = synth.queries_from(corpus, nq)
    dq = torch.from_numpy(queries).cuda()
    rows_all = torch.empty((G, nq, k), dtype=torch.int64, device="cuda")
    dist_all = torch.empty((G, nq, k), dtype=torch.float64, device="cuda")
    cnt = torch.empty(nq, dtype=torch.int32, device="cuda")
    stores = []
    for g in range(G):                       # round-robin row sharding: global = g + local * G
        s = GPUMemoryStore(dim, n // G + 1, row_base=g, row_stride=G)
        s.bulk_load(corpus[g::G])
        s.device_search(dq, k, rows_all[g], dist_all[g], cnt)
        stores.append(s)
    out_r = torch.empty((nq, k), dtype=torch.int64, device="cuda")
    out_d = torch.empty((nq, k), dtype=torch.float64, device="cuda")
    stores[0].device_merge_topk(rows_all, dist_all, G, nq, k, out_r, out_d, cnt)
    torch.cuda.synchronize()
    r = O.search(corpus, queries, k, variant=O.VARIANT_F64)
    assert_topk_matches(out_r.cpu().numpy(), out_d.cpu().numpy(), cnt.cpu().numpy(), r[0], r[1], r[2], what=f"G={G}")
    for s in stores:
        s.close()
