"""The allocation-free search entry point, bltcu_closest_noalloc (include/bltcu.h).

It exists for callers that may not touch the device allocator: a plugin living inside an application that manages
the whole GPU heap itself, a control loop that cannot tolerate the device-wide synchronisation a cudaMalloc can
cause, or a process that has already given every free byte of HBM to its own arena.  Its kernel scratch is a fixed
20 MiB array inside the library, shared by all calls, so the second call starts on top of the first call's leftovers.
That is exactly what this file checks: through the C ABI, two calls in a row, each compared with the regular entry
point (bltcu_closest_ex) on the same input.  Counts and the returned vector must be equal bit for bit on exhaustive
(UNSAT) runs; on SAT runs the decision must agree and the returned point must lie in the box.

Kept in a file that pytest collects last, so a failure here cannot hide the result of any other test.
"""
import numpy as np
import pytest

import oracle
from tests.instances import dense_bilp, reduced_search_input

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def gpu():
    import blt_b200
    return blt_b200


def _same_as_regular(gpu, lat, con):
    ref = gpu.closest(lat, con)
    got = gpu.closest_noalloc(lat, con)
    orc = oracle.closest(lat, con)   # the CPU restatement of closest(): the parity reference of both entry points
    assert got["rc"] == ref["rc"] == orc["rc"] == 0
    assert got["found"] == ref["found"] == oracle.check_soln(lat, con, orc["sol"])
    if got["found"]:
        assert oracle.check_soln(lat, con, got["sol"])
    else:
        assert got["niter"] == ref["niter"] == orc["niter"] and got["nprune"] == ref["nprune"] == orc["nprune"]
        assert (got["sol"] == ref["sol"]).all() and (got["sol"] == orc["sol"]).all()
    return got


def test_two_calls_in_a_row_match_regular_entry_point(gpu):
    A, C = dense_bilp(18, 100, 3)
    big, con_big, _ = reduced_search_input(A, C)       # UNSAT, a few thousand nodes
    A, C = dense_bilp(12, 70, 1)
    small, con_small, _ = reduced_search_input(A, C)   # different shape: the leftovers of call 1 do not fit call 2
    first = _same_as_regular(gpu, big, con_big)
    second = _same_as_regular(gpu, small, con_small)
    third = _same_as_regular(gpu, big, con_big)
    assert not first["found"] and (first["niter"], first["nprune"]) == (third["niter"], third["nprune"])
    assert first["niter"] > 100 and second["rc"] == 0


def test_hard_instance_at_the_size_limit(gpu):
    """156,563 nodes at the largest admitted shape (20 x 20): every warp is busy and its stack of suspended sibling
    loops is deep.  The stacks inside the fixed array are sized for the worst depth-first path, so the search must not
    depend on the small shared record pool (it used to: the pool could run dry with BLT_STATUS_ERROR)."""
    A, C = dense_bilp(20, 140, 48)
    lat, con, _ = reduced_search_input(A, C)
    for _ in range(2):
        got = _same_as_regular(gpu, lat, con)
        assert not got["found"] and got["niter"] + got["nprune"] + 2 == 156563


def test_sat_instance_twice(gpu):
    A, C = dense_bilp(20, 160, 2)
    lat, con, _ = reduced_search_input(A, C)
    for _ in range(2):
        got = _same_as_regular(gpu, lat, con)
        assert got["found"]


def test_rejects_sizes_beyond_the_fixed_scratch(gpu):
    n = gpu.search.NOALLOC_MAX_DIM + 1
    lat = np.eye(n, dtype=np.int64) * 3
    con = np.array([[1, 2]] * n, dtype=np.int64)
    assert gpu.closest_noalloc(lat, con)["rc"] == gpu.BLT_STATUS_INPUT_ERROR
    assert gpu.closest(lat, con)["rc"] == gpu.BLT_STATUS_OK  # the regular entry point takes it
