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


def _unsat_worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist

    from blt_b200.dist import reduce_results

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    n = 4
    # nobody found a point; rank 0 recorded the leaf with path (0, 2, 1, 300), rank 1 the earlier (0, 1, 7, 0); an LP
    # failure on rank 1 must survive (no point was found)
    keys = [np.array([0, 2, 1, 300]), np.array([0, 1, 7, 0])]
    local = dict(rc=(2 if rank == 1 else 0), found=False, sol=np.full(n, 10 + rank, dtype=np.int64), niter=1, nprune=2, nodes=5,
                 pivots=1, donated=0, hist=np.zeros(n + 1, dtype=np.uint64), kernel_ms=1.0, best_key=keys[rank])
    out = reduce_results(local, n)
    q.put((rank, list(map(int, out["sol"])), out["rc"], out["found"]))
    dist.destroy_process_group()


def test_two_rank_unsat_returns_the_depth_first_last_candidate():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29300 + os.getpid() % 150
    procs = [ctx.Process(target=_unsat_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in range(2))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, sol, rc, found in res:
        assert sol == [10] * 4 and rc == 2 and not found  # rank 0's leaf comes later in depth-first order


def _batch_worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist

    from blt_b200.dist import gather_batch

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    count, n = 7, 3
    lo, hi = count * rank // world, count * (rank + 1) // world
    # what bltcu_closest_batch leaves on a rank: its own share filled in, placeholders elsewhere
    status = np.full(count, -1, dtype=np.int32)
    status[lo:hi] = 0
    sol = np.zeros((count, n), dtype=np.int64)
    sol[lo:hi] = -(np.arange(lo, hi)[:, None] * 10 + np.arange(n)[None, :])
    niter = np.zeros(count, dtype=np.uint64)
    niter[lo:hi] = np.arange(lo, hi) + 100
    nprune = np.zeros(count, dtype=np.uint64)
    nprune[lo:hi] = 2 * np.arange(lo, hi)
    found = np.zeros(count, dtype=bool)
    found[lo:hi] = (np.arange(lo, hi) % 2 == 0)
    out = gather_batch(dict(status=status, sol=sol, niter=niter, nprune=nprune, found=found, kernel_ms=3.0 + rank), count)
    q.put((rank, out["status"].tolist(), out["sol"].tolist(), out["niter"].tolist(), out["nprune"].tolist(),
           out["found"].tolist(), out["kernel_ms"]))
    dist.destroy_process_group()


def test_two_rank_batch_gather():
    """Config 5 shards whole problems: every rank ends with the results of all of them."""
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29900 + os.getpid() % 90
    procs = [ctx.Process(target=_batch_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in range(2))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, status, sol, niter, nprune, found, ms in res:
        assert status == [0] * 7
        assert sol == [[-(i * 10 + j) for j in range(3)] for i in range(7)]
        assert niter == [100 + i for i in range(7)] and nprune == [2 * i for i in range(7)]
        assert found == [i % 2 == 0 for i in range(7)]
        assert ms == 4.0


def _flag_worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    import ctypes

    import torch.distributed as dist

    from blt_b200.dist import StopFlag

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    flag = StopFlag(name=f"bltcu_test_{port}")
    flag.reset()
    words = (ctypes.c_int * 16).from_address(flag.address)
    size_word, epoch1 = int(words[2]), int(words[7])   # bytes of the block (mailboxes beyond the header), mailbox epoch
    if rank == 1:
        words[0] = 1                      # what the finder's host thread does
    dist.barrier()
    seen = flag.value
    # the on-demand counter: every rank takes three tickets; together they must be 0..5, each once
    mine = []
    for _ in range(3):
        # Python has no atomic fetch-add on shared memory; ranks take turns instead (the C library does it atomically)
        for turn in range(world):
            dist.barrier()
            if turn == rank:
                mine.append(int(words[1]))
                words[1] += 1
    dist.barrier()
    total = int(words[1])
    flag.reset()
    cleared = flag.value
    assert size_word == flag.size == 16 << 20 and int(words[2]) == size_word and int(words[7]) == epoch1 + 1
    q.put((rank, seen, sorted(mine), total, cleared))
    flag.close()
    dist.destroy_process_group()


def test_shared_block_is_visible_to_both_ranks():
    """blt_b200.dist.StopFlag: the block of bltcu_opts.stop_flag in POSIX shared memory (one process per GPU): stop word,
    on-demand counter, and the size and epoch words that describe the mailboxes behind the header."""
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29100 + os.getpid() % 150
    procs = [ctx.Process(target=_flag_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in range(2))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert all(seen == 1 and total == 6 and cleared == 0 for _, seen, _, total, cleared in res)
    assert sorted(res[0][2] + res[1][2]) == [0, 1, 2, 3, 4, 5]
