"""Parity of the batch form of the search seam (bltcu_closest_batch) with the oracle.

Every problem of a batch is searched in the reference's own depth-first order (idle warps help with subtrees; a
shared problem that turns out SAT is searched again by one warp alone), so the bar is bit-exact for SAT problems too:
status, decision, niter, nprune and the returned vector equal the sequential oracle's.
"""
import os

import numpy as np
import pytest

import oracle
from tests.conftest import GOLDEN
from tests.instances import dense_bilp, reduced_search_input

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def gpu():
    import blt_b200
    return blt_b200


@pytest.fixture(scope="module")
def fixture():
    return np.load(os.path.join(GOLDEN, "batch20.npz"))


def _check_against(fx, got, sel):
    assert (got["status"] == fx["rc"][sel]).all()
    assert (got["found"] == fx["found"][sel]).all()
    assert (got["niter"] == fx["niter"][sel]).all()
    assert (got["nprune"] == fx["nprune"][sel]).all()
    assert (got["sol"] == fx["sol"][sel]).all()


def test_batch_matches_golden_fixture(gpu, fixture):
    """400 instances of config 5 (n = m = 20, 61 of them SAT), expected values from the oracle (make_batch_fixture.py)."""
    got = gpu.closest_batch(fixture["lat"], fixture["con"])
    assert got["rc"] == 0
    _check_against(fixture, got, slice(None))
    assert int(fixture["found"].sum()) == 61 and got["found"].sum() == 61
    # every model is re-checked in exact integers
    for i in np.nonzero(got["found"])[0]:
        assert oracle.check_soln(fixture["lat"][i], fixture["con"][i], got["sol"][i])
    # the summed statistics are the sums: one initial solve per problem on top of niter + nprune (+1 where the top
    # level ended on a prune or a find, which the reference does not count either)
    assert got["niter"].sum() == int(fixture["niter"].sum())
    assert got["nodes"] >= int(fixture["niter"].sum() + fixture["nprune"].sum()) + 400


def test_batch_is_independent_of_the_launch_shape(gpu, fixture):
    sel = slice(0, 96)
    for kw in (dict(ctas=1, warps_per_cta=1), dict(ctas=3, warps_per_cta=5), dict(ctas=148, warps_per_cta=2)):
        got = gpu.closest_batch(fixture["lat"][sel], fixture["con"][sel], **kw)
        assert got["rc"] == 0
        _check_against(fixture, got, sel)


def test_batch_equals_single_searches(gpu, fixture):
    """UNSAT problems: the batch numbers are those of the one-problem entry point (which uses the whole GPU)."""
    idx = [i for i in range(60) if not fixture["found"][i]][:12]
    got = gpu.closest_batch(fixture["lat"][idx], fixture["con"][idx])
    for j, i in enumerate(idx):
        one = gpu.closest(fixture["lat"][i], fixture["con"][i])
        assert one["rc"] == got["status"][j] and not one["found"]
        assert one["niter"] == got["niter"][j] and one["nprune"] == got["nprune"][j]
        assert (one["sol"] == got["sol"][j]).all()


def test_batch_shards_by_rank(gpu, fixture):
    """world = 3: every rank writes its own contiguous share and nothing else."""
    sel = slice(100, 177)
    count = 77
    seen = np.zeros(count, dtype=int)
    for rank in range(3):
        got = gpu.closest_batch(fixture["lat"][sel], fixture["con"][sel], rank=rank, world=3)
        lo, hi = count * rank // 3, count * (rank + 1) // 3
        assert (got["status"][:lo] == -1).all() and (got["status"][hi:] == -1).all()
        part = {k: got[k][lo:hi] for k in ("status", "found", "niter", "nprune", "sol")}
        _check_against(fixture, part, slice(100 + lo, 100 + hi))
        seen[lo:hi] += 1
    assert (seen == 1).all()


@pytest.mark.parametrize("n,m,w", [(8, 8, 40), (10, 14, 150), (12, 12, 90), (5, 5, 400)])
def test_batch_other_shapes_live_oracle(gpu, n, m, w):
    """Shapes other than 20 x 20, including the over-determined twin (more rows than unknowns); expected values
    computed by the oracle in the test."""
    lats, cons = [], []
    for seed in range(24):
        A, C = dense_bilp(n, w, 500 + seed, m=m)
        lat, con, _ = reduced_search_input(A, C)
        if lat.shape != (n, m):
            continue
        lats.append(lat)
        cons.append(con)
    assert len(lats) >= 12
    got = gpu.closest_batch(np.array(lats), np.array(cons))
    assert got["rc"] == 0
    nsat = 0
    for i, (lat, con) in enumerate(zip(lats, cons)):
        ref = oracle.closest(lat, con)
        sat = ref["rc"] == 0 and oracle.check_soln(lat, con, ref["sol"])
        nsat += int(sat)
        assert got["status"][i] == ref["rc"]
        assert bool(got["found"][i]) == sat
        assert got["niter"][i] == ref["niter"] and got["nprune"][i] == ref["nprune"]
        assert (got["sol"][i] == ref["sol"]).all()


def test_batch_edge_cases(gpu):
    lat = np.zeros((0, 4, 4), dtype=np.int64)
    con = np.zeros((0, 4, 2), dtype=np.int64)
    got = gpu.closest_batch(lat, con)
    assert got["rc"] == 0 and got["status"].shape == (0,)
    # m < n is an input error for the whole call (closest.cc:173-177)
    got = gpu.closest_batch(np.ones((2, 4, 3), dtype=np.int64), np.ones((2, 3, 2), dtype=np.int64))
    assert got["rc"] == gpu.BLT_STATUS_INPUT_ERROR
    # a problem whose basis rows are dependent gets its own LP_FAIL, its neighbours are unaffected
    good = np.array([[1, 0], [-1, 2]], dtype=np.int64)
    bad = np.array([[1, 0], [1, 0]], dtype=np.int64)
    box = np.array([[2, 3], [2, 3]], dtype=np.int64)
    got = gpu.closest_batch(np.array([good, bad, good]), np.array([box, box, box]))
    assert got["rc"] == 0
    assert got["status"][0] == 0 and got["status"][2] == 0 and got["status"][1] == gpu.BLT_STATUS_LP_FAIL
    assert oracle.closest(bad, box)["rc"] == gpu.BLT_STATUS_LP_FAIL
    assert got["found"][0] and got["found"][2] and not got["found"][1]
    ref = oracle.closest(good, box)
    assert (got["sol"][0] == ref["sol"]).all() and got["niter"][0] == ref["niter"] and got["nprune"][0] == ref["nprune"]


def test_batch_of_large_problems_is_shared_between_warps(gpu):
    """Three copies of the 1.8 M-node n = m = 60 instance: far more warps than problems, so almost all of the work is
    done by helpers on handed-over subtrees -- counts, histogram total and the returned vector still equal the
    oracle's stored values."""
    z = np.load(os.path.join(GOLDEN, "bench_dense60.npz"))
    lats = np.stack([z["lat"]] * 3)
    cons = np.stack([z["con"]] * 3)
    got = gpu.closest_batch(lats, cons, timeout_ms=300000)
    assert got["rc"] == 0 and (got["status"] == 0).all() and not got["found"].any()
    assert (got["niter"] == int(z["niter"])).all() and (got["nprune"] == int(z["nprune"])).all()
    assert (got["sol"] == z["sol"]).all()
    assert got["nodes"] == 3 * int(z["nodes"]) and got["pivots"] == 3 * int(z["pivots"])
    assert got["donated"] == 3  # all three were shared
