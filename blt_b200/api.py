"""ctypes mirror of include/blt.h (the BLT C API, reference libblt/blt.h:47-134) on top of blt_b200/lib/libblt.so.

    ctx = Context(); x = ctx.var("x"); y = ctx.var("y")
    ctx.assume(2, x - y, 3); ctx.assume(2, 2 * y, 3); ctx.check(); ctx.model("x")

Same names, argument meaning and return codes as the reference; expressions are owned by the context.
"""
from __future__ import annotations

import ctypes
import os
from fractions import Fraction

from .search import lib as _culib

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


def lib() -> ctypes.CDLL:
    global _LIB
    if _LIB is None:
        _culib()  # libbltcu.so first (libblt.so links against it)
        path = os.path.join(_HERE, "lib", "libblt.so")
        if not os.path.exists(path):
            raise RuntimeError(f"{path} is missing: run __graft_entry__.build()")
        L = ctypes.CDLL(path)
        vp, i64, ci = ctypes.c_void_p, ctypes.c_int64, ctypes.c_int
        L.blt_init.argtypes, L.blt_init.restype = [ci], vp
        L.blt_var.argtypes, L.blt_var.restype = [vp, ctypes.c_char_p], vp
        L.blt_const.argtypes, L.blt_const.restype = [vp, i64, i64], vp
        L.blt_add.argtypes, L.blt_add.restype = [vp, vp, vp], vp
        L.blt_smul.argtypes, L.blt_smul.restype = [vp, i64, i64, vp], vp
        L.blt_tof.argtypes, L.blt_tof.restype = [vp, i64, vp], vp
        L.blt_assume.argtypes, L.blt_assume.restype = [vp, i64, i64, vp, i64, i64], ci
        L.blt_check.argtypes, L.blt_check.restype = [vp], ci
        L.blt_model.argtypes, L.blt_model.restype = [vp, ctypes.c_char_p, ctypes.POINTER(i64)], ci
        L.blt_save.argtypes, L.blt_save.restype = [vp], vp
        L.blt_backtrack.argtypes, L.blt_backtrack.restype = [vp, vp], ci
        L.blt_free.argtypes, L.blt_free.restype = [vp], None
        L.blt_b200_last_stats.argtypes = [vp, ctypes.POINTER(ctypes.c_uint64), ctypes.POINTER(ctypes.c_uint64),
                                          ctypes.POINTER(ci), ctypes.POINTER(ctypes.c_double),
                                          ctypes.POINTER(ctypes.c_double)]
        L.blt_b200_last_stats.restype = ci
        L.blt_check_batch.argtypes, L.blt_check_batch.restype = [ctypes.POINTER(vp), ci, ctypes.POINTER(ci)], ci
        L.blt_b200_solve_raw_files.argtypes = [ctypes.c_char_p, ctypes.c_char_p, ci, ctypes.POINTER(i64), ci, ctypes.POINTER(ci),
                                               ctypes.POINTER(ctypes.c_uint64), ctypes.POINTER(ctypes.c_uint64)]
        L.blt_b200_solve_raw_files.restype = ci
        _LIB = L
    return _LIB


class Expr:
    def __init__(self, ctx: "Context", handle):
        self.ctx, self.h = ctx, handle

    def __add__(self, other):
        other = self.ctx.coerce(other)
        return Expr(self.ctx, lib().blt_add(self.ctx.h, self.h, other.h))

    __radd__ = __add__

    def __mul__(self, s):
        f = Fraction(s)
        return Expr(self.ctx, lib().blt_smul(self.ctx.h, f.numerator, f.denominator, self.h))

    __rmul__ = __mul__

    def __neg__(self):
        return self * -1

    def __sub__(self, other):
        return self + (self.ctx.coerce(other) * -1)

    def tof(self, s: int):
        return Expr(self.ctx, lib().blt_tof(self.ctx.h, s, self.h))


class Context:
    def __init__(self, enable_yices: int = 0):
        self.h = lib().blt_init(enable_yices)

    def coerce(self, v):
        return v if isinstance(v, Expr) else self.const(v)

    def var(self, name: str) -> Expr:
        return Expr(self, lib().blt_var(self.h, name.encode()))

    def const(self, v) -> Expr:
        f = Fraction(v)
        return Expr(self, lib().blt_const(self.h, f.numerator, f.denominator))

    def assume(self, lo, expr: Expr, hi) -> int:
        a, b = Fraction(lo), Fraction(hi)
        return lib().blt_assume(self.h, a.numerator, a.denominator, expr.h, b.numerator, b.denominator)

    def check(self) -> int:
        return lib().blt_check(self.h)

    def model(self, name: str):
        v = ctypes.c_int64(0)
        rc = lib().blt_model(self.h, name.encode(), ctypes.byref(v))
        return rc, int(v.value)

    def save(self):
        return lib().blt_save(self.h)

    def backtrack(self, st) -> int:
        return lib().blt_backtrack(self.h, st)

    def last_stats(self):
        ni, np_, rk = ctypes.c_uint64(0), ctypes.c_uint64(0), ctypes.c_int(0)
        lm, sm = ctypes.c_double(0), ctypes.c_double(0)
        lib().blt_b200_last_stats(self.h, ctypes.byref(ni), ctypes.byref(np_), ctypes.byref(rk), ctypes.byref(lm),
                                  ctypes.byref(sm))
        return dict(niter=int(ni.value), nprune=int(np_.value), rank=int(rk.value), lll_ms=lm.value,
                    search_ms=sm.value)

    def free(self):
        if self.h:
            lib().blt_free(self.h)
            self.h = None


def check_batch(contexts):
    """blt_check_batch(): blt_check() for many contexts at once (one batched LLL and one batched search per shape on
    the GPU).  Returns the list of per-context results (BLT_CHECK_SAT / BLT_CHECK_UNSAT / BLT_STATUS_*)."""
    n = len(contexts)
    hs = (ctypes.c_void_p * n)(*[c.h for c in contexts])
    res = (ctypes.c_int * n)()
    rc = lib().blt_check_batch(hs, n, res)
    if rc != 0:
        raise RuntimeError(f"blt_check_batch failed with status {rc}")
    return list(res)


def solve_raw_files(lat_path: str, con_path: str, sound: bool = False):
    """solve_raw() of libblt/test_regression.cc:36-61 on two Boost-format matrix files (blt_b200_solve_raw_files).
    Returns dict(rc, x, niter, nprune): rc BLT_CHECK_SAT 10 / BLT_CHECK_UNSAT 11 or a status code; x = coefficients of the
    point with respect to the rows of the lattice file."""
    cap = 4096
    x = (ctypes.c_int64 * cap)()
    rows, ni, npr = ctypes.c_int(0), ctypes.c_uint64(0), ctypes.c_uint64(0)
    rc = lib().blt_b200_solve_raw_files(lat_path.encode(), con_path.encode(), 1 if sound else 0, x, cap, ctypes.byref(rows),
                                        ctypes.byref(ni), ctypes.byref(npr))
    return dict(rc=rc, x=[int(v) for v in x[:rows.value]], niter=int(ni.value), nprune=int(npr.value))
