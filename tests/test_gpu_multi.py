"""Several GPUs on one search: the share split, the shared stop word (closest() returns at the first point,
libblt/closest.cc:300-312) and the one-process multi-GPU form of the seam (bltcu_opts.gpus / BLT_GPUS).

The stop word is testable on ONE device: the ranks of a split run one after the other, the finder first."""
import ctypes

import numpy as np
import pytest

import oracle
from tests.instances import dense_bilp, reduced_search_input

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def gpu():
    import blt_b200
    return blt_b200


def _device_count():
    import torch
    return torch.cuda.device_count()


def _sat_instance():
    for seed in range(20):
        A, C = dense_bilp(20, 160, seed)
        lat, con, _ = reduced_search_input(A, C)
        ref = oracle.closest(lat, con)
        if ref["rc"] == 0 and oracle.check_soln(lat, con, ref["sol"]) and ref["nodes"] > 3000:
            return lat, con
    raise AssertionError("no SAT instance with a tree worth splitting")


def test_stop_word_ends_the_other_ranks(gpu):
    """world = 2 on one device.  Without a stop word each rank searches its whole share; with it, a rank started after
    the finder has set the word returns at once with `stopped` and without a point of its own."""
    lat, con = _sat_instance()
    alone = [gpu.closest(lat, con, rank=r, world=2, split_depth=3) for r in range(2)]
    assert all(a["rc"] == 0 for a in alone) and any(a["found"] for a in alone)
    finder = 0 if alone[0]["found"] else 1
    other = 1 - finder
    word = (ctypes.c_int * 16)()          # plain host memory
    addr = ctypes.addressof(word)
    first = gpu.closest(lat, con, rank=finder, world=2, split_depth=3, stop_flag=addr)
    assert first["rc"] == 0 and first["found"] and oracle.check_soln(lat, con, first["sol"])
    assert word[0] == 1                   # raised by the finder's host thread
    second = gpu.closest(lat, con, rank=other, world=2, split_depth=3, stop_flag=addr)
    assert second["rc"] == 0 and not second["found"] and second["stopped"]
    assert second["nodes"] <= alone[other]["nodes"] // 2 + 64
    # exhaustive mode ignores the stop word, and the on-demand counter (word 1) hands every subtree of the second half
    # to exactly one rank: the two ranks' counts add up to the whole tree whoever takes what
    whole = gpu.closest(lat, con, no_early_exit=1)
    for i in range(16):
        word[i] = 0
    word[0] = 1
    parts = [gpu.closest(lat, con, rank=r, world=2, split_depth=3, stop_flag=addr, no_early_exit=1) for r in (other, finder)]
    assert all(p["rc"] == 0 and not p["stopped"] for p in parts)
    assert sum(p["nodes"] for p in parts) == whole["nodes"] and sum(p["niter"] for p in parts) == whole["niter"]
    assert word[1] > 0


def test_one_process_two_gpus_exhaustive(gpu, golden):
    """bltcu_opts.gpus = 2: host threads, one per device, inside ONE call of the seam; counts, histogram and uhat of
    the 1.8 M-node n = m = 60 instance equal the oracle's."""
    if _device_count() < 2:
        pytest.skip("needs two GPUs")
    z = np.load(f"{golden}/bench_dense60.npz")
    got = gpu.closest(z["lat"], z["con"], gpus=2, timeout_ms=300000)
    assert got["rc"] == 0 and not got["found"] and got["gpus"] == 2
    assert got["niter"] == int(z["niter"]) and got["nprune"] == int(z["nprune"]) and got["nodes"] == int(z["nodes"])
    assert (got["hist"] == z["hist"]).all() and got["pivots"] == int(z["pivots"])
    assert (got["sol"] == z["sol"]).all()


def test_records_shared_between_two_gpus(gpu, golden, monkeypatch):
    """BLTCU_SHARE=1: a GPU that runs dry asks, the other one's warps hand records over through the host threads
    (search.cu, struct Share), and the search ends when every rank is idle and every record sent has arrived.  Whatever
    moves where, counts, histogram, pivots and uhat are those of the oracle, twice in a row."""
    if _device_count() < 2:
        pytest.skip("needs two GPUs")
    monkeypatch.setenv("BLTCU_SHARE", "1")
    z = np.load(f"{golden}/bench_dense60.npz")
    for _ in range(2):
        got = gpu.closest(z["lat"], z["con"], gpus=2, timeout_ms=300000)
        assert got["rc"] == 0 and not got["found"] and got["gpus"] == 2
        assert got["niter"] == int(z["niter"]) and got["nprune"] == int(z["nprune"]) and got["nodes"] == int(z["nodes"])
        assert (got["hist"] == z["hist"]).all() and got["pivots"] == int(z["pivots"])
        assert (got["sol"] == z["sol"]).all()
    lat, con = _sat_instance()
    got = gpu.closest(lat, con, gpus=2)
    assert got["rc"] == 0 and got["found"] and oracle.check_soln(lat, con, got["sol"])


def test_one_process_two_gpus_sat_and_blt_check(gpu, monkeypatch):
    """A SAT instance over two GPUs returns a valid point; BLT_GPUS makes blt_check() do the same behind blt.h."""
    if _device_count() < 2:
        pytest.skip("needs two GPUs")
    lat, con = _sat_instance()
    got = gpu.closest(lat, con, gpus=2)
    assert got["rc"] == 0 and got["found"] and oracle.check_soln(lat, con, got["sol"])
    from blt_b200 import api
    from oracle import context
    monkeypatch.setenv("BLT_GPUS", "2")
    for seed, w in ((0, 110), (2, 160)):
        A, C = dense_bilp(20, w, seed)
        ref = context.check_bilp(A, C)
        ctx = api.Context()
        v = [ctx.var(str(j)) for j in range(20)]
        for i in range(20):
            e = ctx.const(0)
            for j in range(20):
                e = e + int(A[i, j]) * v[j]
            assert ctx.assume(int(C[i, 0]), e, int(C[i, 1])) == 0
        rc = ctx.check()
        assert rc == ref["rc"]
        if rc == 10:
            x = np.array([ctx.model(str(j))[1] for j in range(20)], dtype=object)
            zz = A.astype(object) @ x
            assert all(int(C[i, 0]) <= int(zz[i]) <= int(C[i, 1]) for i in range(20))
        else:
            st = ctx.last_stats()
            assert (st["niter"], st["nprune"]) == (ref["niter"], ref["nprune"])
        ctx.free()
