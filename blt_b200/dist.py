"""One process per GPU: every rank searches its share of the deterministic static split (bltcu_opts.rank/world) and the
counters are summed at the end.  No collective on the data path; torch.distributed is plumbing.

    res = blt_b200.dist.closest_sharded(lat, con)   # inside a torchrun / init_process_group context
"""
from __future__ import annotations

import numpy as np


def reduce_results(local: dict, n: int, group=None, device=None) -> dict:
    """All-reduces the per-rank statistics of one sharded search.  `local` is the dict returned by blt_b200.closest();
    works with the nccl backend (device tensors) and with gloo (CPU tensors, used by the CPU tests)."""
    import torch
    import torch.distributed as dist

    world = dist.get_world_size(group) if dist.is_initialized() else 1
    if world == 1:
        return dict(local)
    dev = device if device is not None else ("cuda" if dist.get_backend(group) == "nccl" else "cpu")
    hist = np.zeros(n + 1, dtype=np.float64)
    hist[: len(local["hist"])] = np.asarray(local["hist"], dtype=np.float64)
    sums = torch.tensor([float(local["niter"]), float(local["nprune"]), float(local["nodes"]), float(local["pivots"]),
                         float(local.get("donated", 0))] + hist.tolist(), dtype=torch.float64, device=dev)
    maxs = torch.tensor([float(local["kernel_ms"]), float(local["rc"]), 1.0 if local["found"] else 0.0],
                        dtype=torch.float64, device=dev)
    dist.all_reduce(sums, op=dist.ReduceOp.SUM, group=group)
    dist.all_reduce(maxs, op=dist.ReduceOp.MAX, group=group)
    rank = dist.get_rank(group)
    sol = torch.from_numpy(np.asarray(local["sol"], dtype=np.int64).copy()).to(dev)

    def _bcast(src_rank):
        dist.broadcast(sol, src=dist.get_global_rank(group, src_rank) if group is not None else src_rank, group=group)

    # SAT: the point of the lowest rank that found one.  UNSAT: the reference returns the last leaf it recorded in
    # depth-first order (closest.cc:300-303), so the rank whose candidate has the lexicographically last path key
    # (bltcu_stats.best_key, top level first) provides `sol`; without any recorded leaf every rank keeps zeros.
    finder = torch.tensor([rank if local["found"] else world], dtype=torch.int64, device=dev)
    dist.all_reduce(finder, op=dist.ReduceOp.MIN, group=group)
    if int(finder.item()) < world:
        _bcast(int(finder.item()))
    else:
        key = local.get("best_key")
        mine = torch.full((n + 1,), -1, dtype=torch.int64, device=dev)
        if key is not None:
            mine[0] = 1
            mine[1:] = torch.as_tensor(np.asarray(key, dtype=np.int64), device=dev)
        allk = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(allk, mine, group=group)
        cands = [(tuple(int(x) for x in k[1:].tolist()), r) for r, k in enumerate(allk) if int(k[0].item()) == 1]
        if cands:
            _bcast(max(cands)[1])
        else:
            sol.zero_()
    s = sums.cpu().numpy()
    out = dict(local)
    out.update(niter=int(s[0]), nprune=int(s[1]), nodes=int(s[2]), pivots=int(s[3]), donated=int(s[4]),
               hist=s[5:].astype(np.uint64), kernel_ms=float(maxs[0].item()), rc=int(maxs[1].item()),
               found=bool(maxs[2].item() > 0), sol=sol.cpu().numpy())
    # an LP failure in one rank's share is irrelevant once a point was found anywhere (bltcu_closest_ex does the same
    # for its own warps); a hard error on any rank stays an error
    if out["found"] and out["rc"] == 2:
        out["rc"] = 0
    return out


class StopFlag:
    """The shared block of a multi-rank search (bltcu_opts.stop_flag: stop word + on-demand subtree counter) when the
    ranks are processes on one node: 16 ints in a POSIX shared-memory
    page (/dev/shm), mapped by every rank (bltcu_opts.stop_flag); each rank's host thread relays it to and from its GPU
    while the search kernel runs.  The rank that finds a point sets it, the others stop within a millisecond: closest() returns
    at the first point (closest.cc:300-312).  No collective is involved: torch.distributed only provides the barriers
    around reset()."""

    HEADER_INTS = 1024   # the first 4096 bytes: flags and counters (search.cu, struct Share); mailboxes follow

    def __init__(self, name: str | None = None, size_bytes: int = 16 << 20):
        """size_bytes: 4096 gives the legacy block (stop word and on-demand counter only); anything larger also holds
        the mailboxes through which the ranks' host threads pass subtree records between the GPUs (8 records per rank
        when they fit)."""
        import ctypes
        import mmap
        import os

        import torch.distributed as dist

        self.rank = dist.get_rank() if dist.is_initialized() else 0
        self.size = max(mmap.PAGESIZE, (int(size_bytes) + mmap.PAGESIZE - 1) // mmap.PAGESIZE * mmap.PAGESIZE)
        self.path = "/dev/shm/" + (name or f"bltcu_stop_{os.environ.get('MASTER_PORT', '0')}_{os.getuid()}")
        if self.rank == 0:
            fd = os.open(self.path, os.O_CREAT | os.O_RDWR | os.O_TRUNC, 0o600)
            os.ftruncate(fd, self.size)
        if dist.is_initialized():
            dist.barrier()
        if self.rank != 0:
            fd = os.open(self.path, os.O_RDWR)
        self._mm = mmap.mmap(fd, self.size)
        os.close(fd)
        # [0] stop word, [1] on-demand subtree counter, [2] bytes of the block, [7] epoch of the mailbox sequence numbers
        self._words = (ctypes.c_int * self.HEADER_INTS).from_buffer(self._mm)
        self.address = ctypes.addressof(self._words)
        self._epoch = 0

    def reset(self):
        """Zeroes the header and starts a new epoch; collective (every rank calls it before the search starts)."""
        import ctypes

        import torch.distributed as dist

        if dist.is_initialized():
            dist.barrier()
        if self.rank == 0:
            self._epoch = (self._epoch + 1) & 0x7fff
            ctypes.memset(self.address, 0, 4 * self.HEADER_INTS)
            self._words[2] = self.size if self.size > 4096 else 0
            self._words[7] = self._epoch
        if dist.is_initialized():
            dist.barrier()

    @property
    def value(self) -> int:
        return int(self._words[0])

    def close(self):
        import os

        import torch.distributed as dist

        if dist.is_initialized():
            dist.barrier()
        if self.rank == 0 and os.path.exists(self.path):
            os.unlink(self.path)


def closest_sharded(lat, con, split_depth: int = 6, stop: "StopFlag | None" = None, **opts) -> dict:
    """closest() of libblt/closest.h:50-54 over all ranks of the default process group (one GPU each).  With `stop`
    (a StopFlag every rank constructed) the first point found on any rank ends the search on all of them."""
    import torch.distributed as dist

    import blt_b200

    rank = dist.get_rank() if dist.is_initialized() else 0
    world = dist.get_world_size() if dist.is_initialized() else 1
    if stop is not None:
        stop.reset()
        opts["stop_flag"] = stop.address
    local = blt_b200.closest(lat, con, rank=rank, world=world, split_depth=split_depth if world > 1 else 0, **opts)
    return reduce_results(local, np.asarray(lat).shape[0])


def gather_batch(local: dict, count: int, group=None, device=None) -> dict:
    """Every rank has searched its contiguous share [count*rank/world, count*(rank+1)/world) of a batch (the entries of
    the other problems are untouched placeholders); returns the complete per-problem arrays on every rank.  One
    all-reduce per array: the shares are disjoint, so a sum over ranks of the masked arrays is the gather."""
    import torch
    import torch.distributed as dist

    world = dist.get_world_size(group) if dist.is_initialized() else 1
    if world == 1:
        return dict(local)
    rank = dist.get_rank(group)
    dev = device if device is not None else ("cuda" if dist.get_backend(group) == "nccl" else "cpu")
    lo, hi = count * rank // world, count * (rank + 1) // world
    out = dict(local)
    for key in ("status", "sol", "niter", "nprune", "found"):
        a = np.asarray(local[key])
        masked = np.zeros(a.shape, dtype=np.int64)
        masked[lo:hi] = a[lo:hi].astype(np.int64)
        t = torch.from_numpy(masked).to(dev)
        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)
        out[key] = t.cpu().numpy().astype(a.dtype)
    ms = torch.tensor([float(local.get("kernel_ms", 0.0))], dtype=torch.float64, device=dev)
    dist.all_reduce(ms, op=dist.ReduceOp.MAX, group=group)
    out["kernel_ms"] = float(ms.item())
    return out


def closest_batch_sharded(lats, cons, **opts) -> dict:
    """closest_batch() over all ranks of the default process group: rank r searches the r-th contiguous share of the
    problems on its GPU, then the per-problem results are gathered on every rank."""
    import torch.distributed as dist

    import blt_b200

    rank = dist.get_rank() if dist.is_initialized() else 0
    world = dist.get_world_size() if dist.is_initialized() else 1
    local = blt_b200.closest_batch(lats, cons, rank=rank, world=world, **opts)
    return gather_batch(local, len(np.asarray(lats)))
