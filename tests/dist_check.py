"""torchrun target: row-sharded recall on real GPUs vs the oracle (used by tests/test_gpu_multi.py).
    torchrun --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P tests/dist_check.py"""
import os
import sys
from pathlib import Path

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from omnia_b200 import synth                      # noqa: E402
from omnia_b200.dist import ShardedRecall          # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    n, dim = 50_000, 256
    full = synth.gauss_unit(n, dim)
    ok = True
    for nq, k in ((4, 10), (200, 10), (64, 100)):
        queries = synth.queries_from(full, nq, seed=nq)
        sr = ShardedRecall(dim, n, local, exchange=os.environ.get("EXCHANGE", "nccl"), nq_max=256, k_max=128)
        sr.store.bulk_load(full[rank::world])
        dq = torch.from_numpy(queries).cuda()
        r, d, c = sr.search_device(dq, k)
        torch.cuda.synchronize()
        # the pipelined form must give the same answer
        hs = [sr.begin(dq, k, s) for s in range(3)]
        outs = [sr.finish(h) for h in hs]
        torch.cuda.synchronize()
        for o in outs:
            assert torch.equal(o[0], r) and torch.equal(o[1], d) and torch.equal(o[2], c), "pipelined answer differs"
        r, d, c = r.cpu().numpy(), d.cpu().numpy(), c.cpu().numpy()
        if rank == 0:
            from oracle import oracle as O
            wr, wd, wc = O.search(full, queries, k, variant=O.VARIANT_F64)
            same = r == wr
            good = np.array_equal(c, wc) and np.allclose(d, wd, rtol=0, atol=1e-9) and \
                (same.all() or np.abs(d[~same] - wd[~same]).max() <= 1e-6)
            print(f"world={world} nq={nq} k={k}: {'ok' if good else 'MISMATCH'}", flush=True)
            ok &= bool(good)
        sr.close()
    flag = torch.tensor([1 if ok else 0], device="cuda")
    dist.broadcast(flag, 0)
    dist.destroy_process_group()
    sys.exit(0 if flag.item() == 1 else 1)


if __name__ == "__main__":
    main()
