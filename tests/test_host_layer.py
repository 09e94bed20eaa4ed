"""Host side of the boundary on a machine without a GPU: bltcu_lll against the oracle's exact LLL (bit-exact B, U,
rank), the blt.h bookkeeping that never reaches the device, the .bilp parser through run_blt."""
import os
import subprocess

import numpy as np
import pytest

import blt_b200
from blt_b200 import api
from oracle import bltio
from oracle.lll import lll as oracle_lll
from tests.conftest import ROOT

RUN_BLT = os.path.join(ROOT, "blt_b200", "lib", "run_blt")


def _to_i64(M):
    return np.array([[((int(x) + (1 << 63)) % (1 << 64)) - (1 << 63) for x in row] for row in M], dtype=np.int64)


def test_lll_matches_oracle_random():
    rng = np.random.default_rng(0)
    for _ in range(150):
        r, c = int(rng.integers(1, 10)), int(rng.integers(1, 10))
        B = rng.integers(-9, 10, size=(r, c))
        rk, Bo, Uo = blt_b200.lll(B)
        rk2, B2, U2 = oracle_lll(B.tolist())
        assert rk == rk2 and (Bo == _to_i64(B2)).all() and (Uo == _to_i64(U2)).all()


@pytest.mark.parametrize("r,c,a", [(40, 20, 100), (20, 20, 100), (60, 60, 100), (24, 30, 10**9)])
def test_lll_matches_oracle_shapes(r, c, a):  # 40 x 20: config 2, the dependent-row branch
    B = np.random.default_rng(r * 1000 + c).integers(-a, a + 1, size=(r, c))
    rk, Bo, Uo = blt_b200.lll(B)
    rk2, B2, U2 = oracle_lll(B.tolist())
    assert rk == rk2 == min(r, c)
    assert (Bo == _to_i64(B2)).all() and (Uo == _to_i64(U2)).all()   # U narrowed like NTL's to_long()
    assert not Bo[: r - rk].any()                                    # zero rows first (context.cc:176-181)


def test_lll_rank_kats(golden):  # libblt/test_ntl.cc:32-44 (entries fit int64)
    N = np.array(bltio.read_ntl(f"{golden}/lattice_95.tst"), dtype=np.int64)
    assert blt_b200.lll(N)[0] == 64
    assert blt_b200.lll(N[:, :32])[0] == 32


def test_lll_fixture_bases(golden):
    for name in ("test_95", "test_50"):
        M = bltio.read_boost(f"{golden}/{name}.lat")
        rk, Bo, Uo = blt_b200.lll(M)
        rk2, B2, U2 = oracle_lll(M.tolist())
        assert rk == rk2 == 64 and (Bo == _to_i64(B2)).all() and (Uo == _to_i64(U2)).all()


def test_fixed_capacity_lll_equals_arbitrary_precision():
    """The batched GPU kernel runs lll_fixed.h; here the same code runs on the host and must give the basis, the
    transform and the rank of the arbitrary-precision implementation, or say that its numbers did not fit."""
    rng = np.random.default_rng(7)
    cases = []
    for _ in range(120):
        r, c = int(rng.integers(2, 22)), int(rng.integers(2, 22))
        amax = int(rng.choice([3, 100, 10**4, 10**9]))
        cases.append(rng.integers(-amax, amax + 1, size=(r, c)))
    for _ in range(30):  # dependent rows
        r, c = int(rng.integers(3, 14)), int(rng.integers(2, 10))
        M = rng.integers(-50, 51, size=(r, c))
        M[-1] = M[0] + 2 * M[1]
        cases.append(M)
    for _ in range(10):  # config 2: 40 x 20 of rank 20
        cases.append(rng.integers(-100, 101, size=(40, 20)))
    fitted = 0
    for B in cases:
        rk0, B0, U0 = blt_b200.lll_arbitrary(B)
        rkd, Bd, Ud = blt_b200.lll(B)  # the default entry point: whichever arithmetic finished, the same answer
        assert rkd == rk0 and (Bd == B0).all() and (Ud == U0).all()
        for limbs in (16, 32):
            rc, rk1, B1, U1 = blt_b200.lll_fixed_host(B, limbs)
            if rc != blt_b200.BLT_STATUS_OK:
                assert rc == blt_b200.BLT_STATUS_ERROR  # did not fit: said so, nothing truncated
                continue
            fitted += 1
            assert rk1 == rk0 and (B1 == B0).all() and (U1 == U0).all()
    assert fitted >= 250
    # numbers that cannot fit 256 bits are reported, not truncated
    big = rng.integers(-10**17, 10**17, size=(12, 12))
    assert blt_b200.lll_fixed_host(big, 8)[0] == blt_b200.BLT_STATUS_ERROR


def test_api_bookkeeping():  # libblt/test_context.cc:67-104 and blt.h contracts, no device needed
    ctx = api.Context()
    x, y = ctx.var("x"), ctx.var("y")
    assert ctx.assume(0, x + y, 1) == blt_b200.BLT_STATUS_OK
    s1 = ctx.save()
    z = ctx.var("z")
    assert ctx.assume(3, z, 5) == blt_b200.BLT_STATUS_OK
    s2 = ctx.save()
    assert ctx.backtrack(s1) == blt_b200.BLT_STATUS_OK
    assert ctx.backtrack(s2) == blt_b200.BLT_STATUS_ERROR          # no backtracking forward
    assert ctx.assume(2, x, 1) == blt_b200.BLT_CHECK_UNSAT          # crossing bounds (context.cc:113-131)
    assert api.lib().blt_smul(ctx.h, 1, 0, x.h) is None             # d == 0 -> NULL (blt.cc:54-65)
    assert ctx.model("nope")[0] == blt_b200.BLT_STATUS_INPUT_ERROR
    ctx.free()


def test_check_input_errors_before_the_device():
    ctx = api.Context()
    assert ctx.check() == blt_b200.BLT_STATUS_INPUT_ERROR           # nothing assumed (context.cc:146-149)
    x = ctx.var("x")
    assert ctx.assume(3, x, 3) == blt_b200.BLT_STATUS_OK
    assert ctx.check() == blt_b200.BLT_STATUS_INPUT_ERROR           # zero width is rejected (context.cc:465-468)
    ctx.free()


def test_run_blt_cli_errors(tmp_path):
    r = subprocess.run([RUN_BLT, "-v"], capture_output=True, text=True)
    assert r.returncode == 0 and r.stdout.strip() == "BLT version 1.0.0"
    bad = tmp_path / "bad.bilp"
    bad.write_text("1 <= 2 3\n")
    r = subprocess.run([RUN_BLT, str(bad)], capture_output=True, text=True)
    assert r.returncode == 1 and "cannot parse input file" in r.stderr
    r = subprocess.run([RUN_BLT], capture_output=True, text=True)
    assert r.returncode == 1 and "Usage:" in r.stdout


# ---- exact layer-meets-cube decision (the reference's Yices known answers, libblt/test_yices.cc:37-123) -------------
def _cube(p, b):
    return [[pi - b, pi + b] for pi in p]


def test_yprune_2d_known_answers():
    import blt_b200
    B = [[8, 4], [4, 8]]
    assert blt_b200.yprune(B, [0, 1], 1, _cube([4, 4], 2)) == (20, 11)      # farther than 2: prune
    assert blt_b200.yprune(B, [0, 1], 1, _cube([4, 4], 3)) == (20, 10)      # not farther than 3, just barely: keep
    assert blt_b200.yprune(B, [0, 1], 1, _cube([4, 4], 1000)) == (20, 10)
    for u1 in range(-10, 11):
        assert blt_b200.yprune(B, [0, u1], 1, _cube([4, 4], 4)) == (20, 10 if u1 in (0, 1) else 11)


def test_yprune_3d_known_answers():
    import blt_b200
    B = [[1, 0, 3], [0, 2, 1], [4, 4, -4]]
    for u2 in range(-10, 11):
        assert blt_b200.yprune(B, [0, 0, u2], 2, _cube([0, 0, 0], 4)) == (20, 10 if -1 <= u2 <= 1 else 11)


def test_yprune_agrees_with_the_layer_distance_and_knows_its_limits():
    """Property in the spirit of libblt/test_compare_glpk_yices.cc:47-104: exact UNSAT => the oracle's layer distance is
    definitely larger than the radius, exact SAT => it is approximately <= the radius (1e-4, relative)."""
    import numpy as np
    import oracle
    import blt_b200
    rng = np.random.default_rng(7)
    for _ in range(300):
        n = int(rng.integers(2, 5))
        B = rng.integers(-9, 10, size=(n, n))
        if abs(np.linalg.det(B)) < 0.5:
            continue
        k = int(rng.integers(1, n))
        u = rng.integers(-4, 5, size=n)
        p = rng.integers(-20, 21, size=n)
        b = int(rng.integers(1, 12))
        rc, dec = blt_b200.yprune(B, u, k, _cube(list(p), b))
        assert rc == 20 and dec in (10, 11)
        orc, d, _ = oracle.infdist(B.astype(float), u, k, p.astype(float))
        assert orc == 0
        if dec == 11:
            assert d > b * (1 - 1e-9)
        else:
            assert d <= b * (1 + 1e-4)
    big = np.eye(8, dtype=np.int64)
    assert blt_b200.yprune(big, [0] * 8, 7, _cube([0] * 8, 1)) == (21, 12)    # beyond BLTCU_YPRUNE_MAX_K: unknown


def test_boost_matrix_reader_rejects_malformed_files(tmp_path):
    """read_lattice (libblt/parser.cc:43-55) of the product: bad input is BLT_STATUS_INPUT_ERROR before any GPU work."""
    from blt_b200 import api
    good_con = tmp_path / "c.con"
    good_con.write_text("[2,2]((0,1),(0,1))")
    for text in ("", "[2,2]((1,2),(3))", "[2,2]((1,2),(3,4)", "[0,2]()", "[2,2]((1,x),(3,4))"):
        bad = tmp_path / "bad.lat"
        bad.write_text(text)
        assert api.solve_raw_files(str(bad), str(good_con))["rc"] == 3
    lat = tmp_path / "l.lat"
    lat.write_text("[2,3]((1,0,0),(0,1,0))")
    assert api.solve_raw_files(str(lat), str(good_con))["rc"] == 3   # con must be c x 2 with c = 3 (closest.cc:173-177)


def test_fp_decision_pass_of_the_lll_equals_the_exact_algorithm():
    """lll_fp.h (host instantiation): decisions in fp64 against a margin, row operations in exact int64.  Wherever the
    pass completes, basis, transform and rank are the oracle's exact integral LLL bit for bit; dependent generating sets
    are always deferred; bltcu_lll (which tries this pass first) equals the oracle either way."""
    import numpy as np
    import blt_b200
    from oracle.lll import lll as oracle_lll

    def wrap(M):
        return np.array([[((int(x) + 2**63) % 2**64) - 2**63 for x in row] for row in M], dtype=np.int64)

    rng = np.random.default_rng(5)
    done = deferred = 0
    for r, c, amax, cnt in ((20, 20, 100, 60), (12, 12, 50, 40), (30, 30, 100, 8), (60, 60, 100, 1), (10, 14, 1000, 20), (5, 5, 3, 40)):
        for _ in range(cnt):
            B = rng.integers(-amax, amax + 1, size=(r, c))
            ork, oB, oU = oracle_lll([[int(x) for x in row] for row in B])
            d, rk, Bf, Uf = blt_b200.lll_fp_host(B)
            if d:
                deferred += 1
                assert (Bf == B).all()
            else:
                done += 1
                assert rk == ork == r and (Bf == wrap(oB)).all() and (Uf == wrap(oU)).all()
            rk2, B2, U2 = blt_b200.lll(B)
            assert rk2 == ork and (B2 == wrap(oB)).all() and (U2 == wrap(oU)).all()
    assert done > 100 and deferred < done // 4
    for _ in range(3):   # 40 vectors of rank 20: the dependent-row branch belongs to the multiprecision routine
        assert blt_b200.lll_fp_host(rng.integers(-100, 101, size=(40, 20)))[0]
    B = np.array([[1, 2], [2, 4]])                       # dependent rows of a square matrix
    assert blt_b200.lll_fp_host(B)[0]
    B = np.array([[3, 1], [1, 1]]) * 2**40               # inner products beyond 2^53
    assert blt_b200.lll_fp_host(B)[0]
