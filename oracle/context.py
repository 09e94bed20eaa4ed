"""Context::check pipeline restated.  TEST INFRASTRUCTURE (part of oracle/).

Follows libblt/context.cc:139-280 (check), :405-444 (convert_to_integral), :455-501 (stretch_to_cube) and the
`solve_raw` helper of libblt/test_regression.cc:36-61.  Arithmetic is done in Python ints; where the reference
would overflow int64 (undefined behaviour there, e.g. the lcm of test_50.bilp's widths) this
restatement raises OverflowError instead of reproducing the undefined result.
"""
from __future__ import annotations

from fractions import Fraction
from math import gcd

import numpy as np

import oracle
from oracle.lll import lll

I64_MAX = (1 << 63) - 1


def _lcm(a: int, b: int) -> int:
    return abs(a * b) // gcd(a, b) if a and b else 0


def _chk(x: int) -> int:
    if not (-I64_MAX - 1 <= x <= I64_MAX):
        raise OverflowError("int64 overflow (undefined behaviour in the reference)")
    return x


def convert_to_integral(L0, C0):
    """L0: vars x rows of Fractions, C0: rows x 2 Fractions -> integer L, C (context.cc:405-444)."""
    nrows, ncols = len(L0), len(L0[0])
    L = [[0] * ncols for _ in range(nrows)]
    C = [[0, 0] for _ in range(ncols)]
    for c in range(ncols):
        mult = 1
        for r in range(nrows):
            mult = _chk(_lcm(mult, Fraction(L0[r][c]).denominator))
        mult = _chk(_lcm(mult, Fraction(C0[c][0]).denominator))
        mult = _chk(_lcm(mult, Fraction(C0[c][1]).denominator))
        for r in range(nrows):
            f = Fraction(L0[r][c])
            L[r][c] = _chk((mult // f.denominator) * f.numerator)
        for t in (0, 1):
            f = Fraction(C0[c][t])
            C[c][t] = _chk((mult // f.denominator) * f.numerator)
    return L, C


def stretch_to_cube(L, C):
    """In place; returns a BLT status (context.cc:455-501)."""
    n = len(C)
    l = 1
    for i in range(n):
        r = C[i][1] - C[i][0]
        if r <= 0:
            return oracle.BLT_STATUS_INPUT_ERROR
        l = _chk(_lcm(l, r))
    for i in range(n):
        mult = l // (C[i][1] - C[i][0])
        C[i][0] = _chk(C[i][0] * mult)
        C[i][1] = _chk(C[i][1] * mult)
        for row in L:
            row[i] = _chk(row[i] * mult)
    return oracle.BLT_STATUS_OK


def check(L0, C0, max_nodes: int = 0):
    """L0: vars x rows coefficient matrix (rows of the lattice = variables), C0: rows x 2 bounds (constants already
    subtracted).  Returns dict(rc, model, niter, nprune, rank, lat, U).  rc is BLT_CHECK_SAT/UNSAT or a status."""
    m = len(L0)
    n = len(L0[0]) if m else 0
    if m == 0 or n == 0:
        return dict(rc=oracle.BLT_STATUS_INPUT_ERROR)
    L, C = convert_to_integral(L0, C0)
    rc = stretch_to_cube(L, C)
    if rc != oracle.BLT_STATUS_OK:
        return dict(rc=rc)
    rank, Lr, U = lll(L)
    if rank == 0:
        return dict(rc=oracle.BLT_STATUS_INPUT_ERROR)
    newL = np.array([[_chk(x) for x in row] for row in Lr[m - rank:]], dtype=np.int64)
    Cn = np.array(C, dtype=np.int64)
    res = oracle.closest(newL, Cn, max_nodes=max_nodes)
    if res["rc"] != oracle.BLT_STATUS_OK:
        return dict(rc=res["rc"], niter=res["niter"], nprune=res["nprune"])
    presoln = [0] * (m - rank) + [int(x) for x in res["sol"]]
    soln = [sum(presoln[i] * U[i][j] for i in range(m)) for j in range(m)]
    sat = oracle.check_soln(newL, Cn, res["sol"])
    return dict(rc=oracle.BLT_CHECK_SAT if sat else oracle.BLT_CHECK_UNSAT, model=soln, niter=res["niter"],
                nprune=res["nprune"], rank=rank, lat=newL, con=Cn, U=U, hist=res["hist"], pivots=res["pivots"],
                sol=res["sol"])


def check_bilp(A, bounds, max_nodes: int = 0):
    """run_blt's framing (libblt/run_blt.cc:151-171): rows `lo <= A[i].x <= hi`, every variable active."""
    A = np.asarray(A)
    L0 = [[Fraction(int(A[i][j])) for i in range(A.shape[0])] for j in range(A.shape[1])]
    C0 = [[Fraction(int(b[0])), Fraction(int(b[1]))] for b in bounds]
    return check(L0, C0, max_nodes=max_nodes)


def solve_raw(lat, con, max_nodes: int = 0):
    """test_regression.cc:36-61: LLL then closest, no stretch; returns dict with rc SAT/UNSAT."""
    rank, Lr, U = lll([[int(x) for x in row] for row in np.asarray(lat)])
    M = np.array(Lr, dtype=np.int64)
    res = oracle.closest(M, con, max_nodes=max_nodes)
    if res["rc"] != oracle.BLT_STATUS_OK:
        return dict(rc=res["rc"], niter=res["niter"], nprune=res["nprune"])
    sat = oracle.check_soln(M, con, res["sol"])
    out = dict(res)
    out["rc"] = oracle.BLT_CHECK_SAT if sat else oracle.BLT_CHECK_UNSAT
    out["lat"] = M
    out["rank"] = rank
    return out
