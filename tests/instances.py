"""Synthetic instance generators shared by the tests and bench.py.

Generators only; the LLL reduction used to prepare search inputs for the parity tests comes from the oracle
(test infrastructure), so this module must not be imported by the product.
"""
from __future__ import annotations

import numpy as np


def dense_bilp(n: int, w: int, seed: int, m: int | None = None, amax: int = 100):
    """Config 2/4/5: A in Z^{m x n} uniform on [-amax, amax], bounds c_i <= A_i x <= c_i + w, c_i uniform on
    [-1e6, 1e6].  Returns (A, C) in run_blt's framing (rows = inequalities)."""
    m = m or n
    rng = np.random.default_rng(seed)
    A = rng.integers(-amax, amax + 1, size=(m, n))
    c = rng.integers(-10**6, 10**6 + 1, size=m)
    return A.astype(np.int64), np.stack([c, c + w], axis=1).astype(np.int64)


def subset_sum(nvars: int, bits: int, seed: int):
    """Config 3: 0 <= x_i <= 1 and 2s-1 <= 2 a.x <= 2s+1 with a hidden solution of nvars/2 ones."""
    rng = np.random.default_rng(seed)
    a = rng.integers(1, 2**bits + 1, size=nvars).astype(np.int64)
    xs = np.zeros(nvars, dtype=np.int64)
    xs[rng.permutation(nvars)[: nvars // 2]] = 1
    s = int(a @ xs)
    A = np.vstack([np.eye(nvars, dtype=np.int64), 2 * a[None, :]])
    C = np.array([[0, 1]] * nvars + [[2 * s - 1, 2 * s + 1]], dtype=np.int64)
    return A, C, xs


def reduced_search_input(A, C):
    """(A, C) -> the (lat, con) pair closest() sees, via the oracle's restatement of Context::check's front half
    (convert/stretch/LLL, context.cc:139-193).  Test-side helper."""
    from fractions import Fraction

    from oracle import context
    from oracle.lll import lll

    A = np.asarray(A)
    L0 = [[Fraction(int(A[i][j])) for i in range(A.shape[0])] for j in range(A.shape[1])]
    C0 = [[Fraction(int(b[0])), Fraction(int(b[1]))] for b in C]
    L, Ci = context.convert_to_integral(L0, C0)
    rc = context.stretch_to_cube(L, Ci)
    assert rc == 0
    rank, Lr, U = lll(L)
    lat = np.array(Lr[len(Lr) - rank:], dtype=np.int64)
    return lat, np.array(Ci, dtype=np.int64), U
