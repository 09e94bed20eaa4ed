"""The blt.h API and run_blt on the GPU: the reference's own integration tests (libblt/test_context.cc,
libblt/test_blt.c, test/UnitTests.hs) re-expressed against libblt.so, with the oracle's restated Context::check as the
cross-check for decision, model and counters."""
import os
import re
import subprocess

import numpy as np
import pytest

import oracle
from fractions import Fraction as Fr
from oracle import bltio, context
from tests.conftest import ROOT

pytestmark = pytest.mark.gpu
RUN_BLT = os.path.join(ROOT, "blt_b200", "lib", "run_blt")
SAT, UNSAT = 10, 11


@pytest.fixture(scope="module")
def api():
    from blt_b200 import api
    return api


def _system(api, rows, names):
    ctx = api.Context()
    v = [ctx.var(nm) for nm in names]
    for lo, coef, hi, const in rows:
        e = ctx.const(const)
        for c, x in zip(coef, v):
            if c != 0:
                e = e + c * x
        assert ctx.assume(lo, e, hi) == 0
    return ctx


def _oracle(rows, nvars):
    L0 = [[Fr(r[1][v]) for r in rows] for v in range(nvars)]
    C0 = [[Fr(r[0]) - Fr(r[3]), Fr(r[2]) - Fr(r[3])] for r in rows]
    return context.check(L0, C0)


CASES = [
    ("sat_2d", [(0, [1, 0], 3, 0), (3, [0, 1], 5, 0)], "xy", SAT),                                  # test_context.cc:107-121
    ("sat_3d", [(-1, [1, 0, 1], 1, 0), (-3, [0, 1, -1], 5, 0), (-1, [1, 13, 101], 2, 0)], "uvz", SAT),  # :123-143
    ("unsat", [(1, [3, 0], 2, 0), (3, [0, 5], 4, 0)], "ab", UNSAT),                                  # :145-153
    ("pentagon", [(0, [0, 1], 2, 0), (-5, [1, 0], 3, 0), (-5, [1, -1], 2, 0), (-5, [-2, 1], 0, 0)], "xy", SAT),  # :187-208
    ("2ip_hard", [(0, [3, -4], 20, 0), (-20, [1, -4], 0, 0), (-20, [-2, -1], -4, 0), (-5, [-1, -2], 20, 0),
                  (-20, [2, -1], 4, 0)], "xy", SAT),                                                   # :219-242
    ("affine", [(1, [1, 0], 2, 3), (100, [1, 1], 125, 6)], "xy", SAT),                               # :244-261
    ("parallelogram", [(0, [4283000000, -3245000001], 999999, 0), (1, [4283000001, -3245000000], 1000000, 0)], "xy",
     UNSAT),                                                                                           # test_blt.c:82-106
    ("parallelogram_simple", [(0, [428300, -324501], 99, 0), (1, [428301, -324500], 100, 0)], "xy", UNSAT),  # :126-150
]


@pytest.mark.parametrize("name,rows,names,want", CASES, ids=[c[0] for c in CASES])
def test_reference_systems(api, name, rows, names, want):
    ctx = _system(api, rows, names)
    rc = ctx.check()
    assert rc == want
    ref = _oracle(rows, len(names))
    assert ref["rc"] == rc
    if rc == SAT:
        vals = [ctx.model(nm) for nm in names]
        assert all(v[0] == 0 for v in vals)
        x = [v[1] for v in vals]
        for lo, coef, hi, const in rows:
            s = sum(c * xi for c, xi in zip(coef, x)) + const
            assert lo <= s <= hi
        if name == "2ip_hard":
            assert x == [2, 1]  # the unique model
    else:
        st = ctx.last_stats()
        assert (st["niter"], st["nprune"]) == (ref["niter"], ref["nprune"])
    ctx.free()


def test_rational_rows(api):  # test_context.cc:156-178
    h, t, f = Fr(1, 2), Fr(1, 3), Fr(1, 5)
    ctx = api.Context()
    u, v, z = ctx.var("u"), ctx.var("v"), ctx.var("z")
    ctx.assume(-h, h * (u + z), h)
    ctx.assume(-3 * t, t * (v - z), 5 * t)
    ctx.assume(-f, f * (u + 13 * v + 101 * z), 2 * f)
    assert ctx.check() == SAT
    u0, v0, z0 = (ctx.model(n)[1] for n in "uvz")
    assert -1 <= u0 + z0 <= 1 and -3 <= v0 - z0 <= 5 and -1 <= u0 + 13 * v0 + 101 * z0 <= 2
    ctx.free()


def test_incremental_check_and_backtrack(api):  # test_context.cc:123-143 + blt_save/blt_backtrack
    ctx = api.Context()
    u, v, z = ctx.var("u"), ctx.var("v"), ctx.var("z")
    ctx.assume(-1, u + z, 1)
    ctx.assume(-3, v - z, 5)
    assert ctx.check() == SAT
    st = ctx.save()
    ctx.assume(1, 3 * u, 2)       # no integer u
    assert ctx.check() == UNSAT
    assert ctx.backtrack(st) == 0
    assert ctx.check() == SAT
    ctx.free()


def test_sound_mode_request_selects_the_gpu_sound_search(api):  # blt_init(1), run_blt -y; context.cc:210-228, closest.cc:258-267
    """The reference needs libyices.so for this and fails without it; here sound mode is the search kernel's own
    (a prune needs a bound clear of the radius).  Decisions and models must be those of the plain mode."""
    for name, rows, names, want in CASES[:6]:
        ctx = api.Context(1)
        v = [ctx.var(nm) for nm in names]
        for lo, coef, hi, const in rows:
            e = ctx.const(const)
            for c, x in zip(coef, v):
                if c != 0:
                    e = e + c * x
            assert ctx.assume(lo, e, hi) == 0
        assert ctx.check() == want, name
        if want == SAT:
            x = [ctx.model(nm)[1] for nm in names]
            for lo, coef, hi, const in rows:
                assert lo <= sum(c * xi for c, xi in zip(coef, x)) + const <= hi
        ctx.free()


def test_tof(api):  # libblt/test_expr.cc: round(2/7 * (4x + y + 2/7)) -> x:1, y:0, const 0
    ctx = api.Context()
    x, y = ctx.var("x"), ctx.var("y")
    e = (Fr(2, 7) * (4 * x + y + ctx.const(Fr(2, 7)))).tof(1)
    ctx.assume(5, e, 6)   # i.e. 5 <= x <= 6
    ctx.assume(0, y, 0 + 1)
    assert ctx.check() == SAT
    assert 5 <= ctx.model("x")[1] <= 6
    ctx.free()


@pytest.mark.parametrize("name", ["pentagon", "test_4x4"])
def test_run_blt_fixtures(golden, name):  # config 1 of BASELINE.json
    path = f"{golden}/{name}.bilp"
    r = subprocess.run([RUN_BLT, "-m", path], capture_output=True, text=True, timeout=120)
    assert r.returncode == SAT
    lines = r.stdout.strip().splitlines()
    assert re.fullmatch(rf"PROBLEM {re.escape(path)} RUNTIME \S+ niter XXX npru XXX rc XXX check  SAT ", lines[0])
    mt = re.fullmatch(rf"PROBLEM {re.escape(path)} MODEL \((.*)\)", lines[1])
    x = np.array([int(t) for t in mt.group(1).split(",")], dtype=np.int64)
    A, C = bltio.read_bilp(path)
    v = A @ x
    assert (C[:, 0] <= v).all() and (v <= C[:, 1]).all()
    ref = context.check_bilp(A, C)
    assert ref["rc"] == SAT


def test_run_blt_dense_unsat(tmp_path):  # config 2/5 shape: n = 20, counters through BLT_STATS
    from tests.instances import dense_bilp
    A, C = dense_bilp(20, 110, 0)
    p = tmp_path / "d20.bilp"
    p.write_text("\n".join(f"{C[i, 0]} <= {' '.join(str(a) for a in A[i])} <= {C[i, 1]}" for i in range(20)) + "\n")
    r = subprocess.run([RUN_BLT, str(p)], capture_output=True, text=True, timeout=120, env=dict(os.environ, BLT_STATS="1"))
    assert r.returncode == UNSAT and "check  UNSAT " in r.stdout
    ref = context.check_bilp(A, C)
    assert ref["rc"] == UNSAT
    mt = re.search(r"BLT_STATS niter (\d+) nprune (\d+) rank (\d+)", r.stderr)
    assert (int(mt.group(1)), int(mt.group(2)), int(mt.group(3))) == (ref["niter"], ref["nprune"], 20)


def test_check_batch_equals_single_checks(api):
    """blt_check_batch (extension): the reference's own systems plus a pile of config-5-style dense systems, all in one
    call; every decision equals blt_check's and the oracle's, every SAT context has a model that satisfies its rows."""
    systems = [(rows, names, want) for _, rows, names, want in CASES]
    rng = np.random.default_rng(11)
    for i in range(40):  # 8 variables, 8 rows: A uniform on [-100, 100], common width (a mix of SAT and UNSAT)
        A = rng.integers(-100, 101, size=(8, 8))
        c = rng.integers(-10**6, 10**6 + 1, size=8)
        w = int(rng.choice([60, 110, 160]))
        rows = [(int(c[r]), [int(x) for x in A[r]], int(c[r]) + w, 0) for r in range(8)]
        systems.append((rows, "abcdefgh", None))
    ctxs = [_system(api, rows, names) for rows, names, _ in systems]
    got = api.check_batch(ctxs)
    nsat = 0
    for (rows, names, want), ctx, g in zip(systems, ctxs, got):
        ref = _oracle(rows, len(names))
        assert g == ref["rc"]
        if want is not None:
            assert g == want
        single = _system(api, rows, names)
        assert single.check() == g
        single.free()
        if g == SAT:
            nsat += 1
            x = [ctx.model(nm)[1] for nm in names]
            for lo, coef, hi, const in rows:
                val = sum(c * xi for c, xi in zip(coef, x)) + const
                assert lo <= val <= hi
            st = ctx.last_stats()
            assert st["niter"] == ref["niter"] and st["nprune"] == ref["nprune"]  # the batch search is the sequential order
    assert 5 <= nsat < len(systems)
    for ctx in ctxs:
        ctx.free()
    assert api.check_batch([]) == []


def test_zero_is_feasible_means_sat(api):
    """test/QCTests.hs:48-115: random systems up to 12 x 12 whose bounds [-b, -1] ... [1, b] all contain 0 must come
    back SAT with a model inside the bounds -- here all of them in one blt_check_batch call (many shapes at once)."""
    rng = np.random.default_rng(48)
    systems = []
    for _ in range(36):
        nv, nr = int(rng.integers(1, 13)), int(rng.integers(1, 13))
        A = rng.integers(-9, 10, size=(nr, nv))
        for r in range(nr):
            if not A[r].any():
                A[r, int(rng.integers(0, nv))] = 1
        lo = -rng.integers(1, 11, size=nr)
        hi = rng.integers(1, 11, size=nr)
        names = [f"v{i}" for i in range(nv)]
        rows = [(int(lo[r]), [int(x) for x in A[r]], int(hi[r]), 0) for r in range(nr)]
        # every variable must occur in some row, or it is not part of the lattice (context.cc:336-377)
        used = A.any(axis=0)
        names = [nm for nm, u in zip(names, used) if u]
        rows = [(l, [c for c, u in zip(cf, used) if u], h, k) for l, cf, h, k in rows]
        systems.append((rows, names))
    ctxs = [_system(api, rows, names) for rows, names in systems]
    got = api.check_batch(ctxs)
    for (rows, names), ctx, g in zip(systems, ctxs, got):
        assert g == SAT
        x = [ctx.model(nm)[1] for nm in names]
        for lo, coef, hi, const in rows:
            assert lo <= sum(c * xi for c, xi in zip(coef, x)) + const <= hi
        ctx.free()


def test_reference_c_client_unchanged():
    """libblt/test_blt.c, the reference's own plain-C client of blt.h, compiled unchanged against include/blt.h +
    include/codes.h and linked with libblt.so (blt_b200.build.build_ref_client, built where /root/reference exists):
    its three systems (one SAT with a printed model, two UNSAT parallelograms, the first needing genuine 64-bit
    coefficients behind #ifdef ENVIRONMENT_64, test_blt.c:15-18) must pass on the GPU path."""
    exe = os.path.join(ROOT, "oracle", "_ref", "test_blt")
    if not os.path.exists(exe):
        pytest.skip("oracle/_ref/test_blt was not built (no /root/reference at build time)")
    r = subprocess.run([exe], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "Tests pass!" in r.stdout


@pytest.mark.parametrize("name", ["test_4x4", "test_95", "test_50", "test_50_8"])
def test_run_blt_latcon_on_the_reference_fixtures(api, golden, name):
    """The legacy lattice/constraint pairs of test/data through the PRODUCT (reader of the Boost matrix text,
    libblt/parser.cc:43-55; solve_raw, libblt/test_regression.cc:36-61, :84-106): SAT, and the returned coefficients
    give a point of the box when multiplied with the rows of the file."""
    lat_p, con_p = f"{golden}/{name}.lat", f"{golden}/{name}.con"
    r = api.solve_raw_files(lat_p, con_p)
    assert r["rc"] == SAT
    lat = np.array(bltio.read_boost(lat_p), dtype=object)
    con = bltio.read_boost(con_p)
    z = np.array(r["x"], dtype=object) @ lat
    assert all(int(con[i][0]) <= int(z[i]) <= int(con[i][1]) for i in range(len(con)))
    out = subprocess.run([RUN_BLT, "-m", "--latcon", lat_p, con_p], capture_output=True, text=True, timeout=300)
    assert out.returncode == SAT and " SAT " in out.stdout and "MODEL (" in out.stdout
    assert re.search(r"niter XXX npru XXX rc XXX", out.stdout)       # the reference's line, XXX and all (run_blt.cc:173-175)


def test_checked_stretch_names_the_overflow(golden, tmp_path):
    """test_50 as a .bilp (the transposed lattice with its bounds, what libblt/latcon2bilp.cc writes) needs a 73-bit lcm
    in stretch_to_cube.  With BLT_CHECKED_STRETCH=1 the product says so and returns
    BLT_STATUS_INPUT_ERROR (exit 3) instead of wrapping silently."""
    lat = bltio.read_boost(f"{golden}/test_50.lat")
    con = bltio.read_boost(f"{golden}/test_50.con")
    bilp = tmp_path / "test_50.bilp"
    with open(bilp, "w") as f:
        for i in range(len(con)):
            f.write(f"{int(con[i][0])} <= " + " ".join(str(int(lat[j][i])) for j in range(len(lat))) + f" <= {int(con[i][1])}\n")
    env = dict(os.environ, BLT_CHECKED_STRETCH="1")
    out = subprocess.run([RUN_BLT, str(bilp)], capture_output=True, text=True, timeout=300, env=env)
    assert out.returncode == 3 and "does not fit 64 bits" in out.stderr
