"""CPU oracle for the BLT lattice-search hot path.  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this package;
the product (blt_b200, libbltcu.so, libblt.so, run_blt) never does.

  oracle.closest / oracle.infdist / oracle.check_soln   -> ctypes wrappers over oracle/blt_oracle.c
  oracle.lll.lll                                          -> exact integral LLL (pure Python ints)
  oracle.context.check                                    -> Context::check pipeline restated
  oracle.bltio                                            -> .bilp / .lat / .con / NTL-format readers
"""
from __future__ import annotations

import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

BLT_STATUS_OK, BLT_STATUS_ERROR, BLT_STATUS_LP_FAIL, BLT_STATUS_INPUT_ERROR = 0, 1, 2, 3
BLT_CHECK_SAT, BLT_CHECK_UNSAT, BLT_CHECK_UNKNOWN = 10, 11, 12
ORACLE_TRUNCATED = 100


def build(force: bool = False) -> str:
    """Compile oracle/blt_oracle.c into oracle/libblt_oracle.so (gcc, -ffp-contract=off)."""
    so = os.path.join(_HERE, "libblt_oracle.so")
    src = os.path.join(_HERE, "blt_oracle.c")
    if force or not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.check_call(
            ["gcc", "-O2", "-ffp-contract=off", "-fPIC", "-shared", "-o", so, src, "-lm"], cwd=_HERE
        )
    return so


def lib() -> ctypes.CDLL:
    global _LIB
    if _LIB is None:
        L = ctypes.CDLL(build())
        i64p = ctypes.POINTER(ctypes.c_int64)
        u64p = ctypes.POINTER(ctypes.c_uint64)
        f64p = ctypes.POINTER(ctypes.c_double)
        L.oracle_closest.argtypes = [i64p, i64p, ctypes.c_int, ctypes.c_int, i64p, u64p, u64p, u64p, u64p,
                                     ctypes.c_uint64]
        L.oracle_closest.restype = ctypes.c_int
        L.oracle_check_soln.argtypes = [i64p, i64p, ctypes.c_int, ctypes.c_int, i64p]
        L.oracle_check_soln.restype = ctypes.c_int
        L.oracle_infdist.argtypes = [f64p, ctypes.c_int, ctypes.c_int, i64p, ctypes.c_int, f64p, f64p, f64p]
        L.oracle_infdist.restype = ctypes.c_int
        L.oracle_set_sound.argtypes = [ctypes.c_double]
        L.oracle_set_sound.restype = None
        L.oracle_inband.argtypes = []
        L.oracle_inband.restype = ctypes.c_uint64
        _LIB = L
    return _LIB


def _i64(a):
    return np.ascontiguousarray(a, dtype=np.int64)


def _p(a, t):
    return a.ctypes.data_as(ctypes.POINTER(t))


def closest(lat, con, max_nodes: int = 0, sound: bool = False):
    """closest() of libblt/closest.cc:166-327.  Returns dict(rc, sol, niter, nprune, hist, pivots, ...)."""
    lat = _i64(lat)
    con = _i64(con)
    if lat.ndim != 2 or con.ndim != 2 or con.shape[1] != 2 or con.shape[0] != lat.shape[1]:
        return dict(rc=BLT_STATUS_INPUT_ERROR, sol=None, niter=0, nprune=0)
    n, m = lat.shape
    if n == 0 or m == 0 or m < n:
        return dict(rc=BLT_STATUS_INPUT_ERROR, sol=None, niter=0, nprune=0)
    sol = np.zeros(n, dtype=np.int64)
    niter = ctypes.c_uint64(0)
    nprune = ctypes.c_uint64(0)
    hist = np.zeros(n + 1, dtype=np.uint64)
    stats = np.zeros(4, dtype=np.uint64)
    lib().oracle_set_sound(1.0e-6 if sound else 0.0)   # SOUND_MARGIN of blt_b200/csrc/search.cu
    rc = lib().oracle_closest(_p(lat, ctypes.c_int64), _p(con, ctypes.c_int64), n, m, _p(sol, ctypes.c_int64),
                              ctypes.byref(niter), ctypes.byref(nprune), _p(hist, ctypes.c_uint64),
                              _p(stats, ctypes.c_uint64), ctypes.c_uint64(max_nodes))
    return dict(rc=rc, sol=sol, niter=int(niter.value), nprune=int(nprune.value), hist=hist,
                pivots=int(stats[0]), refactors=int(stats[1]), quick=int(stats[2]), colds=int(stats[3]),
                nodes=int(hist.sum()), inband=int(lib().oracle_inband()))


def check_soln(lat, con, u) -> bool:
    """check_soln() of libblt/closest.cc:330-340."""
    lat = _i64(lat)
    con = _i64(con)
    u = _i64(u)
    n, m = lat.shape
    return bool(lib().oracle_check_soln(_p(lat, ctypes.c_int64), _p(con, ctypes.c_int64), n, m,
                                        _p(u, ctypes.c_int64)))


def infdist(B, u, k: int, p):
    """infdist() of libblt/glpk_wrapper.cc:33-70 as a cold solve.  Returns (rc, dist, x[0:k])."""
    B = np.ascontiguousarray(B, dtype=np.float64)
    u = _i64(u)
    p = np.ascontiguousarray(p, dtype=np.float64)
    n, m = B.shape
    x = np.zeros(max(n, 1), dtype=np.float64)
    d = ctypes.c_double(0.0)
    rc = lib().oracle_infdist(_p(B, ctypes.c_double), n, m, _p(u, ctypes.c_int64), k, _p(p, ctypes.c_double),
                              _p(x, ctypes.c_double), ctypes.byref(d))
    return rc, float(d.value), x[:k].copy()
