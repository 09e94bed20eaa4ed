"""Parity of the CUDA search (through the C ABI of include/bltcu.h) with the CPU oracle.

Bit-exact bar: status, SAT/UNSAT decision, niter, nprune, the per-level node histogram, the pivot count and the
returned vector (uhat) on exhaustive runs; decision and model validity on SAT runs (a parallel search may find a
different lattice point first).  Floating point: the layer distance of the reference's LP
known answers within the reference's own 1e-4 relative tolerance (DIST_EPS, libblt/glpk_wrapper.h:13), and
bit-for-bit against the oracle.
"""
import numpy as np
import pytest

import oracle
from oracle import bltio, context
from tests.instances import dense_bilp, reduced_search_input, subset_sum

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def gpu():
    import blt_b200
    return blt_b200


def _compare_exhaustive(gpu, lat, con, **opts):
    ref = oracle.closest(lat, con)
    got = gpu.closest(lat, con, **opts)
    assert got["rc"] == ref["rc"]
    sat_ref = ref["rc"] == 0 and oracle.check_soln(lat, con, ref["sol"])
    sat_got = got["rc"] == 0 and got["found"]
    assert sat_ref == sat_got
    if sat_got:
        assert oracle.check_soln(lat, con, got["sol"]) and gpu.check_soln(lat, con, got["sol"])
    else:
        assert got["niter"] == ref["niter"] and got["nprune"] == ref["nprune"]
        assert got["nodes"] == ref["nodes"]
        assert (got["hist"] == ref["hist"]).all()
        assert got["pivots"] == ref["pivots"]
        assert (got["sol"] == ref["sol"]).all()
    return ref, got


# ---- reference LP known answers on the device (libblt/test_glpk.cc) ------------------------------------
def test_infdist_kats(gpu):
    cases = [([[2, 1], [1, 2]], [0, 1], 1, [1, 1], 0.66667), ([[3, 0], [0, 3]], [0, 1], 1, [1, 0], 3.0),
             ([[3, 0], [0, 3]], [0, 1], 1, [0, 0], 3.0), ([[3, 0], [0, 3]], [0, 1], 1, [0, 3], 0.0),
             ([[1, 0], [-1, 2]], [0, 0], 2, [2.5, 2.5], 0.0)]
    for B, u, k, p, want in cases:
        rc, d, x = gpu.infdist(B, u, k, p)
        rc2, d2, x2 = oracle.infdist(B, u, k, p)
        assert rc == 0 and rc2 == 0
        assert abs(d - want) <= max(abs(d), abs(want)) * 1e-4 + 1e-12  # the reference's tolerance
        assert d == d2 and (x == x2).all()                              # bit-exact against the oracle
    rc, d, x = gpu.infdist([[1, 0], [-1, 2]], [0, 0], 2, [2.5, 2.5])
    assert abs(x[0] - 3.75) < 1e-9 and abs(x[1] - 1.25) < 1e-9


def test_infdist_random_bitexact(gpu):
    rng = np.random.default_rng(3)
    for _ in range(60):
        n = int(rng.integers(2, 14))
        m = n + int(rng.integers(0, 4))
        B = rng.integers(-30, 31, size=(n, m)).astype(float)
        if np.linalg.matrix_rank(B) < n:
            continue
        k = int(rng.integers(1, n + 1)) if m > n else int(rng.integers(1, n))
        u = rng.integers(-5, 6, size=n)
        p = rng.integers(-60, 61, size=m) + 0.5 * rng.integers(0, 2, size=m)
        rc, d, x = gpu.infdist(B, u, k, p)
        rc2, d2, x2 = oracle.infdist(B, u, k, p)
        assert rc == rc2 == 0 and d == d2 and (x == x2).all()


def test_check_soln(gpu):  # libblt/test_closest.cc:90-114
    A, C = [[1, 0], [-1, 2]], [[2, 3], [2, 3]]
    assert not gpu.check_soln(A, C, [15, -5])
    assert gpu.check_soln(A, C, [3, 1]) and gpu.check_soln(A, C, [4, 1])
    assert not gpu.check_soln(A, C, [5, 1])


def test_bad_input(gpu):  # libblt/test_closest.cc:33-45
    A = np.zeros((2, 2), dtype=np.int64)
    assert gpu.closest(A, np.zeros((1, 2), dtype=np.int64))["rc"] == gpu.BLT_STATUS_INPUT_ERROR
    assert gpu.closest(A, np.zeros((2, 3), dtype=np.int64))["rc"] == gpu.BLT_STATUS_INPUT_ERROR
    assert gpu.closest(np.zeros((3, 2), dtype=np.int64), np.zeros((2, 2), dtype=np.int64))["rc"] == gpu.BLT_STATUS_INPUT_ERROR


def test_closest_state_init(gpu):  # libblt/test_closest.cc:51-63
    _compare_exhaustive(gpu, np.array([[1, 0], [-1, 2]]), np.array([[2, 3], [2, 3]]))


def test_tiny_dims(gpu):
    _compare_exhaustive(gpu, np.array([[3]]), np.array([[1, 2]]))      # n = 1, UNSAT
    _compare_exhaustive(gpu, np.array([[3]]), np.array([[2, 4]]))      # n = 1, SAT
    _compare_exhaustive(gpu, np.array([[3, 1]]), np.array([[1, 2], [0, 5]]))  # n = 1 < m


def test_fixtures_sat(gpu, golden):  # libblt/test_closest.cc:139-156, test_regression.cc:84-106
    M = bltio.read_boost(f"{golden}/test_4x4.lat")
    C = bltio.read_boost(f"{golden}/test_4x4.con")
    _compare_exhaustive(gpu, M, C)
    for name in ("test_95", "test_50", "test_50_8"):
        lat = bltio.read_boost(f"{golden}/{name}.lat")
        con = bltio.read_boost(f"{golden}/{name}.con")
        r = context.solve_raw(lat, con)
        got = gpu.closest(r["lat"], con)
        assert got["rc"] == 0 and got["found"] and oracle.check_soln(r["lat"], con, got["sol"])


@pytest.mark.parametrize("n,w,seed", [(6, 40, 0), (10, 60, 1), (14, 80, 2), (20, 60, 0), (20, 110, 0), (20, 110, 5),
                                       (24, 100, 3), (30, 100, 0)])
def test_dense_unsat_counts(gpu, n, w, seed):
    A, C = dense_bilp(n, w, seed)
    lat, con, _ = reduced_search_input(A, C)
    _compare_exhaustive(gpu, lat, con)


def test_overdetermined(gpu):  # n < m: 12 variables, 20 rows
    A, C = dense_bilp(12, 120, 4, m=20)
    lat, con, _ = reduced_search_input(A, C)
    assert lat.shape == (12, 20)
    _compare_exhaustive(gpu, lat, con)


def test_counts_independent_of_schedule(gpu):
    """One warp, one CTA, static splits of several depths and the default dynamic schedule all enumerate the same
    tree."""
    A, C = dense_bilp(22, 105, 11)
    lat, con, _ = reduced_search_input(A, C)
    ref, got = _compare_exhaustive(gpu, lat, con)
    assert ref["rc"] == 0 and not got["found"], "instance must be UNSAT for this test"
    for opts in (dict(ctas=1, warps_per_cta=1), dict(ctas=3, warps_per_cta=2), dict(split_depth=1),
                 dict(split_depth=4), dict(split_depth=8), dict(donate_min_level=2), dict(donate_min_level=15)):
        _compare_exhaustive(gpu, lat, con, **opts)


def test_sharded_ranks_sum(gpu):
    """world = 3: the three shards' counts add up to the single-rank counts (no collective on the data path)."""
    A, C = dense_bilp(22, 105, 11)
    lat, con, _ = reduced_search_input(A, C)
    ref = oracle.closest(lat, con)
    tot = dict(niter=0, nprune=0, nodes=0)
    for r in range(3):
        got = gpu.closest(lat, con, rank=r, world=3, split_depth=5)
        assert got["rc"] == 0 and not got["found"]
        for key in tot:
            tot[key] += got[key]
    assert tot["niter"] == ref["niter"] and tot["nprune"] == ref["nprune"] and tot["nodes"] == ref["nodes"]


def test_sat_dense(gpu):
    for seed in range(4):
        A, C = dense_bilp(20, 160, seed)
        lat, con, _ = reduced_search_input(A, C)
        _compare_exhaustive(gpu, lat, con)


def test_subset_sum_sat(gpu):  # config 3 shape, small
    A, C, xs = subset_sum(16, 20, 0)
    lat, con, U = reduced_search_input(A, C)
    ref = oracle.closest(lat, con)
    got = gpu.closest(lat, con)
    assert got["rc"] == ref["rc"] == 0
    assert got["found"] == oracle.check_soln(lat, con, ref["sol"])
    if got["found"]:
        assert oracle.check_soln(lat, con, got["sol"])


@pytest.mark.parametrize("fixture", ["bench_dense60.npz", "bench_dense60_w120.npz", "bench_dense60_w130.npz", "bench_dense60_w140.npz"])
def test_bench_workload_full_size(gpu, golden, fixture):
    """BASELINE.json's metric shape (n = m = 60) at full size (1.8 M and 6.3 M nodes): counts, histogram, pivots and uhat
    equal the values the oracle produced when the fixture was made (tests/golden/make_bench_fixture.py)."""
    z = np.load(f"{golden}/{fixture}")
    got = gpu.closest(z["lat"], z["con"], timeout_ms=300000)
    assert got["rc"] == 0 and not got["found"]
    assert got["niter"] == int(z["niter"]) and got["nprune"] == int(z["nprune"]) and got["nodes"] == int(z["nodes"])
    assert (got["hist"] == z["hist"]).all() and got["pivots"] == int(z["pivots"])
    assert (got["sol"] == z["sol"]).all()
    # 2 ranks' shares add up (world-size-2 run on one device, one rank after the other)
    tot = 0
    for r in range(2):
        g = gpu.closest(z["lat"], z["con"], rank=r, world=2, split_depth=6, timeout_ms=300000)
        assert g["rc"] == 0
        tot += g["nodes"]
    assert tot == int(z["nodes"])


@pytest.mark.parametrize("name", ["test_95", "test_50_8"])
def test_first_dive_reaches_the_babai_point(gpu, golden, name):
    """The reference finds these two JPEG instances on its first descent (65 solves).  The first descent of the GPU search
    is flagged to go down after the first explored child of every level, so a single warp needs exactly as many."""
    z = np.load(f"{golden}/jpeg_reduced.npz")
    lat, con = z[f"lat_{name}"], z[f"con_{name}"]
    got = gpu.closest(lat, con, ctas=1, warps_per_cta=1)
    assert got["rc"] == 0 and got["found"] and oracle.check_soln(lat, con, got["sol"])
    assert got["nodes"] == int(z[f"nodes_{name}"]) == 65
    assert (got["sol"] == z[f"sol_{name}"]).all()
    full = gpu.closest(lat, con)
    assert full["rc"] == 0 and full["found"] and oracle.check_soln(lat, con, full["sol"])


def test_sound_mode_keeps_layers_inside_the_safety_margin(gpu):
    """bltcu_opts.sound (the reference's -y mode, closest.cc:258-267, without Yices).  Two layers of this instance lie at
    distance 29,999,999.5 and 30,000,000.5 from the centre with radius 29,996,985.5: beyond the 1e-4 band by less
    than one part in 10^6.  The plain search prunes them (5 solves), the sound search keeps them (11 solves, 2 layers
    counted as in-band); both equal the oracle run the same way.  On an ordinary instance the modes coincide."""
    lat = np.array([[1000000000, 0], [0, 30000000]], dtype=np.int64)
    con = np.array([[500000000 - 29996985, 500000000 + 29996986], [-29996985, 29996986]], dtype=np.int64)
    for sound in (0, 1):
        ref = oracle.closest(lat, con, sound=bool(sound))
        got = gpu.closest(lat, con, sound=sound)
        assert got["rc"] == ref["rc"] == 0 and not got["found"]
        assert (got["niter"], got["nprune"], got["nodes"], got["inband"]) == (ref["niter"], ref["nprune"], ref["nodes"], ref["inband"])
        assert (got["nodes"], got["inband"]) == ((11, 2) if sound else (5, 0))
    A, C = dense_bilp(20, 110, 5)
    lat, con, _ = reduced_search_input(A, C)
    ref = oracle.closest(lat, con, sound=True)
    got = gpu.closest(lat, con, sound=1)
    assert (got["niter"], got["nprune"], got["pivots"], got["inband"]) == (ref["niter"], ref["nprune"], ref["pivots"], 0)
