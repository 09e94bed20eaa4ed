"""The N > 1 path on CPU: two gloo ranks, each with the statistics of "its" share of a search, reduced by
blt_b200.dist.reduce_results exactly as bench.py and closest_sharded do after the GPU search."""
import os
import sys

import numpy as np
import torch.multiprocessing as mp

from tests.conftest import ROOT


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist

    from blt_b200.dist import reduce_results

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    n = 5
    local = dict(rc=0, found=(rank == 1), sol=np.arange(n, dtype=np.int64) * (rank + 1), niter=10 + rank, nprune=20 + 2 * rank,
                 nodes=32 + 3 * rank, pivots=100 * (rank + 1), donated=rank, hist=np.arange(n + 1, dtype=np.uint64) + rank,
                 kernel_ms=1.5 + rank)
    out = reduce_results(local, n)
    q.put((rank, out["niter"], out["nprune"], out["nodes"], out["pivots"], list(map(int, out["hist"])), out["kernel_ms"],
           out["found"], list(map(int, out["sol"])), out["rc"]))
    dist.destroy_process_group()


def test_two_rank_reduction():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 400
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in range(2))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, niter, nprune, nodes, pivots, hist, ms, found, sol, rc in res:
        assert (niter, nprune, nodes, pivots) == (21, 42, 67, 300)
        assert hist == [1, 3, 5, 7, 9, 11]
        assert ms == 2.5 and found and rc == 0
        assert sol == [0, 2, 4, 6, 8]  # rank 1 is the only finder: every rank ends with its point
