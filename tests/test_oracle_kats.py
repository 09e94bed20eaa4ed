"""Pins the CPU oracle against every known answer the reference's own tests hold for the hot path.

Each test cites the reference test it restates (paths relative to /root/reference).  Nothing here needs a GPU or
reads /root/reference at run time: inputs are inline or under tests/golden/.
"""
from fractions import Fraction as Fr

import numpy as np
import pytest

import oracle
from oracle import bltio, context
from oracle.lll import lll

DIST_EPS = 1e-4


def approx_eq(a, b, eps=DIST_EPS):  # util.cc:17-20
    return abs(a - b) <= max(abs(a), abs(b)) * eps


# ---- libblt/test_glpk.cc ------------------------------------------------------------------------------
def test_glpk_tilted():  # test_glpk.cc:35-55
    rc, d, _ = oracle.infdist([[2, 1], [1, 2]], [0, 1], 1, [1, 1])
    assert rc == 0 and approx_eq(d, 0.66667)


def test_glpk_horiz():  # test_glpk.cc:72-104
    B = [[3, 0], [0, 3]]
    for p, want in (([1, 0], 3.0), ([0, 0], 3.0), ([0, 3], 0.0)):
        rc, d, _ = oracle.infdist(B, [0, 1], 1, p)
        assert rc == 0 and approx_eq(d, want)


def test_glpk_return():  # test_glpk.cc:106-133 (A(0,0) is 1 when infdist runs)
    rc, d, x = oracle.infdist([[1, 0], [-1, 2]], [0, 0], 2, [2.5, 2.5])
    assert rc == 0 and approx_eq(d, 0.0) or abs(d) < 1e-9
    assert approx_eq(x[0], 3.75) and approx_eq(x[1], 1.25)


@pytest.mark.parametrize("p1,want", [(6.25, 2), (-3.25, -1), (-4.5, -2)])
def test_closest_init_square_lattice(p1, want):  # test_glpk.cc:136-169; C round() is half away from zero
    rc, d, x = oracle.infdist([[3, 0], [0, 3]], [0, 0], 2, [0.0, p1])
    assert rc == 0 and abs(d) < 1e-9
    r = int(np.sign(x[1]) * np.floor(abs(x[1]) + 0.5))
    assert r == want


# ---- libblt/test_yices.cc: exact prune sets hold for any correct L-inf layer distance -------------------
def test_yprune_2d():  # test_yices.cc:37-86
    B, p = [[8, 4], [4, 8]], [4, 4]
    rc, d, _ = oracle.infdist(B, [0, 1], 1, p)
    assert rc == 0 and d > 2 * (1 + DIST_EPS) and d <= 3 * (1 + 1e-9)
    for u1 in range(-10, 11):
        rc, d, _ = oracle.infdist(B, [0, u1], 1, p)
        assert rc == 0
        assert (d <= 4 * (1 + 1e-9)) == (u1 in (0, 1))


def test_yprune_3d():  # test_yices.cc:97-123
    B = [[1, 0, 3], [0, 2, 1], [4, 4, -4]]
    for u2 in range(-10, 11):
        rc, d, _ = oracle.infdist(B, [0, 0, u2], 2, [0, 0, 0])
        assert rc == 0
        assert (d <= 4 * (1 + 1e-9)) == (-1 <= u2 <= 1)


# ---- libblt/test_closest.cc ---------------------------------------------------------------------------
def test_closest_bad_input():  # test_closest.cc:33-45
    A = np.zeros((2, 2), dtype=np.int64)
    assert oracle.closest(A, np.zeros((1, 2), dtype=np.int64))["rc"] == oracle.BLT_STATUS_INPUT_ERROR
    assert oracle.closest(A, np.zeros((2, 3), dtype=np.int64))["rc"] == oracle.BLT_STATUS_INPUT_ERROR


def test_closest_state_init():  # test_closest.cc:51-63
    r = oracle.closest([[1, 0], [-1, 2]], [[2, 3], [2, 3]])
    assert r["rc"] == oracle.BLT_STATUS_OK


def test_check_soln():  # test_closest.cc:90-114
    A, C = [[1, 0], [-1, 2]], [[2, 3], [2, 3]]
    assert not oracle.check_soln(A, C, [15, -5])
    assert oracle.check_soln(A, C, [3, 1])
    assert oracle.check_soln(A, C, [4, 1])
    assert not oracle.check_soln(A, C, [5, 1])


def test_read_lattice(golden):  # test_closest.cc:118-136
    M = bltio.read_boost(f"{golden}/test_95.lat")
    assert M.shape == (64, 64)
    assert int(M.sum()) % (1 << 16) == 61856
    assert bltio.read_boost(f"{golden}/test_95.con").shape == (64, 2)


def test_closest_dct4x4(golden):  # test_closest.cc:139-156 (no LLL there)
    M = bltio.read_boost(f"{golden}/test_4x4.lat")
    C = bltio.read_boost(f"{golden}/test_4x4.con")
    r = oracle.closest(M, C)
    assert r["rc"] == oracle.BLT_STATUS_OK and oracle.check_soln(M, C, r["sol"])


# ---- libblt/test_regression.cc:84-106 (solve_raw: LLL + closest, SAT) -----------------------------------
@pytest.mark.parametrize("name", ["test_95", "test_50", "test_50_8"])
def test_regression_sat(golden, name):
    lat = bltio.read_boost(f"{golden}/{name}.lat")
    con = bltio.read_boost(f"{golden}/{name}.con")
    r = context.solve_raw(lat, con)
    assert r["rc"] == oracle.BLT_CHECK_SAT
    assert r["refactors"] == 0


# ---- libblt/test_ntl.cc:32-44 -------------------------------------------------------------------------
def test_lll_rank_lattice_95(golden):
    N = bltio.read_ntl(f"{golden}/lattice_95.tst")
    assert len(N) == 64 and len(N[0]) == 64
    assert lll(N)[0] == 64
    assert lll([row[:32] for row in N])[0] == 32


# ---- libblt/test_context.cc ---------------------------------------------------------------------------
def _check(rows, nvars):
    """rows: list of (lo, coeffs dict/list over vars, hi, const).  Returns oracle.context.check result."""
    L0 = [[Fr(r[1][v]) for r in rows] for v in range(nvars)]
    C0 = [[Fr(r[0]) - Fr(r[3]), Fr(r[2]) - Fr(r[3])] for r in rows]
    return context.check(L0, C0)


def test_check_sat_2d():  # test_context.cc:107-121
    r = _check([(0, [1, 0], 3, 0), (3, [0, 1], 5, 0)], 2)
    assert r["rc"] == oracle.BLT_CHECK_SAT
    x, y = r["model"]
    assert 0 <= x <= 3 and 3 <= y <= 5


def test_check_sat_3d():  # test_context.cc:123-143
    rows = [(-1, [1, 0, 1], 1, 0), (-3, [0, 1, -1], 5, 0)]
    assert _check(rows, 3)["rc"] == oracle.BLT_CHECK_SAT
    rows.append((-1, [1, 13, 101], 2, 0))
    r = _check(rows, 3)
    assert r["rc"] == oracle.BLT_CHECK_SAT
    u, v, z = r["model"]
    assert -1 <= u + z <= 1 and -3 <= v - z <= 5 and -1 <= u + 13 * v + 101 * z <= 2


def test_check_unsat():  # test_context.cc:145-153
    assert _check([(1, [3, 0], 2, 0), (3, [0, 5], 4, 0)], 2)["rc"] == oracle.BLT_CHECK_UNSAT


def test_check_sat_rat():  # test_context.cc:156-178
    h, t, f = Fr(1, 2), Fr(1, 3), Fr(1, 5)
    rows = [(-h, [h, 0, h], h, 0), (-3 * t, [0, t, -t], 5 * t, 0), (-f, [f, 13 * f, 101 * f], 2 * f, 0)]
    r = _check(rows, 3)
    assert r["rc"] == oracle.BLT_CHECK_SAT
    u, v, z = r["model"]
    assert -1 <= u + z <= 1 and -3 <= v - z <= 5 and -1 <= u + 13 * v + 101 * z <= 2


def test_check_2ip():  # test_context.cc:187-208
    LB = -5
    r = _check([(0, [0, 1], 2, 0), (LB, [1, 0], 3, 0), (LB, [1, -1], 2, 0), (LB, [-2, 1], 0, 0)], 2)
    assert r["rc"] == oracle.BLT_CHECK_SAT
    x, y = r["model"]
    assert 0 <= y <= 2 and x <= 3 and x - y <= 2 and -2 * x + y <= 0


def test_check_2ip_hard():  # test_context.cc:219-242: unique model x=2, y=1
    B = 20
    rows = [(0, [3, -4], B, 0), (-B, [1, -4], 0, 0), (-B, [-2, -1], -4, 0), (-5, [-1, -2], B, 0), (-B, [2, -1], 4, 0)]
    r = _check(rows, 2)
    assert r["rc"] == oracle.BLT_CHECK_SAT and r["model"] == [2, 1]


def test_check_affine():  # test_context.cc:244-261
    r = _check([(1, [1, 0], 2, 3), (100, [1, 1], 125, 6)], 2)
    assert r["rc"] == oracle.BLT_CHECK_SAT
    x, y = r["model"]
    assert 1 <= x + 3 <= 2 and 100 <= x + y + 6 <= 125


# ---- libblt/test_blt.c --------------------------------------------------------------------------------
def test_blt_2d_problem():  # test_blt.c:32-61
    r = _check([(2, [1, -1], 3, 0), (2, [0, 2], 3, 0)], 2)
    assert r["rc"] in (oracle.BLT_CHECK_SAT, oracle.BLT_CHECK_UNSAT)  # the reference only prints it
    if r["rc"] == oracle.BLT_CHECK_SAT:
        x, y = r["model"]
        assert 2 <= x - y <= 3 and 2 <= 2 * y <= 3


def test_tight_parallelogram():  # test_blt.c:82-106: needs genuine 64-bit paths
    r = _check([(0, [4283000000, -3245000001], 999999, 0), (1, [4283000001, -3245000000], 1000000, 0)], 2)
    assert r["rc"] == oracle.BLT_CHECK_UNSAT


def test_tight_parallelogram_simple():  # test_blt.c:126-150
    r = _check([(0, [428300, -324501], 99, 0), (1, [428301, -324500], 100, 0)], 2)
    assert r["rc"] == oracle.BLT_CHECK_UNSAT


# ---- run_blt on the two usable .bilp fixtures ------------------------------------------------
@pytest.mark.parametrize("name", ["pentagon", "test_4x4"])
def test_bilp_fixtures(golden, name):
    A, C = bltio.read_bilp(f"{golden}/{name}.bilp")
    r = context.check_bilp(A, C)
    assert r["rc"] == oracle.BLT_CHECK_SAT
    v = A @ np.array(r["model"], dtype=np.int64)
    assert (C[:, 0] <= v).all() and (v <= C[:, 1]).all()


# ---- test/UnitTests.hs / QCTests.hs style: 0 always feasible => SAT ---------------------------------------
def test_quickcheck_zero_feasible():  # test/QCTests.hs:48-115
    rng = np.random.default_rng(7)
    for _ in range(40):
        nv, nr = int(rng.integers(1, 8)), int(rng.integers(1, 8))
        A = rng.integers(-10, 11, size=(nr, nv))
        if not A.any(axis=0).all():
            continue  # a variable in no row is not "active" in BLT
        b = rng.integers(1, 11, size=(nr, 2))
        C = np.stack([-b[:, 0], b[:, 1]], axis=1)
        r = context.check_bilp(A, C)
        assert r["rc"] == oracle.BLT_CHECK_SAT
        v = A @ np.array(r["model"], dtype=np.int64)
        assert (C[:, 0] <= v).all() and (v <= C[:, 1]).all()


# ---- libblt/test_compare_glpk_yices.cc:47-104 in spirit: the layer distance against an independent LP solver ----------
def test_infdist_agrees_with_an_independent_lp_solver():
    """Random instances in 2 to 12 dimensions: min_x max_j |sum_{i<k} x_i B_ij + sum_{i>=k} u_i B_ij - p_j| from the oracle's
    dual simplex equals the optimum of the same LP handed to SciPy's HiGHS (relative 1e-7), and the returned x attains it."""
    from scipy.optimize import linprog

    rng = np.random.default_rng(2024)
    checked = 0
    for trial in range(150):
        n = int(rng.integers(2, 13))
        m = n + int(rng.integers(0, 4))
        B = rng.integers(-50, 51, size=(n, m)).astype(np.float64)
        if np.linalg.matrix_rank(B) < n:
            continue
        k = int(rng.integers(0, n + 1))
        u = np.zeros(n, dtype=np.int64)
        u[k:] = rng.integers(-5, 6, size=n - k)
        p = rng.integers(-200, 201, size=m).astype(np.float64) / 2.0
        rc, d, x = oracle.infdist(B, u, k, p)
        assert rc == 0
        c = p - u[k:].astype(np.float64) @ B[k:]          # target of the free part
        # variables (x_0..x_{k-1}, t): minimise t subject to |c_j - sum_i x_i B_ij| <= t
        A_ub = np.vstack([np.hstack([-B[:k].T, -np.ones((m, 1))]), np.hstack([B[:k].T, -np.ones((m, 1))])])
        b_ub = np.concatenate([-c, c])
        res = linprog(np.r_[np.zeros(k), 1.0], A_ub=A_ub, b_ub=b_ub, bounds=[(None, None)] * k + [(0, None)], method="highs")
        assert res.status == 0
        assert abs(d - res.fun) <= 1e-7 * max(1.0, abs(res.fun))
        if k:
            assert abs(np.max(np.abs(c - x @ B[:k])) - d) <= 1e-7 * max(1.0, d)
        checked += 1
    assert checked >= 120
