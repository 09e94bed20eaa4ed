"""Study (CPU only, oracle): would a two-level probe below the static split predict the size of each subtree well
enough to level the ranks?  Compiles a copy of oracle/blt_oracle.c with -DORACLE_CENSUS into /tmp, runs the search of a
fixture, and deals the subtrees below the explored nodes of one level largest-first by the probe's weights.

    python profiles/census_study.py [fixture.npz] [depth] [near]
"""
import ctypes
import os
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
fx = sys.argv[1] if len(sys.argv) > 1 else "bench_dense60.npz"
depth = int(sys.argv[2]) if len(sys.argv) > 2 else 6
nears = [int(x) for x in sys.argv[3:]] or [1, 2, 3]
so = os.path.join(tempfile.gettempdir(), "libblt_oracle_census.so")
subprocess.check_call(["gcc", "-O2", "-ffp-contract=off", "-fPIC", "-shared", "-DORACLE_CENSUS", "-o", so,
                       os.path.join(ROOT, "oracle", "blt_oracle.c"), "-lm"])
L = ctypes.CDLL(so)
z = np.load(os.path.join(ROOT, "tests", "golden", fx))
lat = np.ascontiguousarray(z["lat"], dtype=np.int64)
con = np.ascontiguousarray(z["con"], dtype=np.int64)
n, m = lat.shape
i64p, u64p = ctypes.POINTER(ctypes.c_int64), ctypes.POINTER(ctypes.c_uint64)
L.oracle_set_census.argtypes = [ctypes.c_int, ctypes.c_int, u64p, u64p, ctypes.c_int]
L.oracle_closest.argtypes = [i64p, i64p, ctypes.c_int, ctypes.c_int, i64p, u64p, u64p, u64p, u64p, ctypes.c_uint64]
cap = 1 << 20


def lpt(weights, sizes, world):
    order = np.argsort(-weights, kind="stable")
    load = np.zeros(world)
    est = np.zeros(world)
    for i in order:
        r = int(np.argmin(est))
        est[r] += weights[i]
        load[r] += sizes[i]
    return load.max() / load.mean()


for near in nears:
    size = np.zeros(cap, dtype=np.uint64)
    probe = np.zeros(cap, dtype=np.uint64)
    L.oracle_set_census(n - depth + 1, near, size.ctypes.data_as(u64p), probe.ctypes.data_as(u64p), cap)
    sol = np.zeros(n, dtype=np.int64)
    ni, npr = ctypes.c_uint64(), ctypes.c_uint64()
    hist = np.zeros(n + 1, dtype=np.uint64)
    st = np.zeros(4, dtype=np.uint64)
    rc = L.oracle_closest(lat.ctypes.data_as(i64p), con.ctypes.data_as(i64p), n, m, sol.ctypes.data_as(i64p),
                          ctypes.byref(ni), ctypes.byref(npr), hist.ctypes.data_as(u64p), st.ctypes.data_as(u64p), 0)
    cnt = L.oracle_census_count()
    s = size[:cnt].astype(float)
    p = probe[:cnt].astype(float)
    assert int(hist.sum()) == int(z["nodes"]), "the census must not change the search"
    ratio = np.log((s + 1) / (p + 1))
    print(f"{fx}: rc {rc}, {cnt} subtrees below depth {depth}, {int(s.sum())} of {int(hist.sum())} nodes in them; probe of {near} "
          f"levels visits {int(p.sum())} nodes ({100 * p.sum() / hist.sum():.2f} %); log(size/probe) std {ratio.std():.3f}; "
          f"rank correlation of the 10 % largest: {np.corrcoef(np.argsort(np.argsort(s))[s >= np.quantile(s, 0.9)], np.argsort(np.argsort(p))[s >= np.quantile(s, 0.9)])[0, 1]:.3f}")
    for world in (4, 8, 16):
        rr = np.array([s[r::world].sum() for r in range(world)])
        print(f"    world {world:2d}: slowest rank / mean: round-robin {rr.max() / rr.mean():.3f}, largest-first by probe {lpt(p + 1, s, world):.3f}, "
              f"by true size {lpt(s, s, world):.3f}")
