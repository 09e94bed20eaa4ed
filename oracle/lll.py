"""Exact integral LLL, delta = 3/4, with dependent rows.  TEST INFRASTRUCTURE (part of oracle/).

Restates what `LLL(matrix<int64>& B, matrix<int64>& U)` of libblt/ntl_wrapper.cc:37-52 computes, i.e.
`NTL::LLL(det2, B, U)` (NTL 9.11.0, pinned in bootstrap.sh:39-53; source NOT under /root/reference).  NTL's routine
is the all-integer LLL of Cohen, "A Course in Computational Algebraic Number Theory", Algorithm 2.6.7 (the
variant of 2.6.3 that tolerates linearly dependent input): state D[0..s] (Gram sub-determinants), integral
lam[k][j] = D[j]*mu[k][j], P[k] = position of row k among the independent ones (0 = dependent).  Rows found
dependent migrate to the front; on return the first (rows - rank) rows of B are zero, the last `rank` rows are
the reduced basis, and U*B_in = B_out.

PARITY STATUS: unpinned.  The reference's tests pin only the rank (libblt/test_ntl.cc:32-44: lattice_95.tst -> 64,
its first 32 columns -> 32).  The reduced basis entries and U are defined by this restatement (arbitrary
precision Python ints, so exact by construction); tests check the invariants that any correct implementation
must satisfy (U*B_in == B_out, size-reduced, Lovasz with 3/4, zero rows first).
"""
from __future__ import annotations


def _baldiv(a: int, d: int) -> int:
    """Nearest integer to a/d (d > 0), ties towards zero."""
    q, r = divmod(a, d)  # floor division, 0 <= r < d
    r2 = 2 * r
    if r2 > d or (r2 == d and q < 0):
        q += 1
    return q


def _xgcd(a: int, b: int):
    """(g, x, y) with a*x + b*y = g >= 0: classical Euclid on the magnitudes, signs restored afterwards."""
    aneg, bneg = a < 0, b < 0
    u, v = abs(a), abs(b)
    u1, v1, u2, v2 = 1, 0, 0, 1
    while v:
        q, r = divmod(u, v)
        u, v = v, r
        u1, u2 = u2, u1 - q * u2
        v1, v2 = v2, v1 - q * v2
    if aneg:
        u1 = -u1
    if bneg:
        v1 = -v1
    return u, u1, v1


def lll(Bin, a: int = 3, b: int = 4):
    """Returns (rank, B, U) as lists of lists of Python ints.  Bin: r x c integers."""
    B = [[int(x) for x in row] for row in Bin]
    m = len(B)
    n = len(B[0]) if m else 0
    U = [[1 if i == j else 0 for j in range(m)] for i in range(m)]
    # 1-based bookkeeping as in the textbook statement
    P = [0] * (m + 1)
    D = [0] * (m + 1)
    D[0] = 1
    lam = [[0] * (m + 1) for _ in range(m + 1)]  # lam[k][j], j = position index 1..s
    s = 0

    def inner(x, y):
        return sum(p * q for p, q in zip(x, y))

    def incremental_gs(k):
        nonlocal s
        for j in range(1, k):
            posj = P[j]
            if posj == 0:
                continue
            u = inner(B[k - 1], B[j - 1])
            for i in range(1, posj):
                u = (D[i] * u - lam[k][i] * lam[j][i]) // D[i - 1]
            lam[k][posj] = u
        u = inner(B[k - 1], B[k - 1])
        for i in range(1, s + 1):
            u = (D[i] * u - lam[k][i] * lam[k][i]) // D[i - 1]
        if u == 0:
            P[k] = 0
        else:
            s += 1
            P[k] = s
            D[s] = u

    def reduce(k, l):
        pl = P[l]
        if pl == 0:
            return
        if abs(2 * lam[k][pl]) <= D[pl]:
            return
        r = _baldiv(lam[k][pl], D[pl])
        bk, bl = B[k - 1], B[l - 1]
        for t in range(n):
            bk[t] -= r * bl[t]
        uk, ul = U[k - 1], U[l - 1]
        for t in range(m):
            uk[t] -= r * ul[t]
        for j in range(1, l):
            pj = P[j]
            if pj != 0:
                lam[k][pj] -= r * lam[l][pj]
        lam[k][pl] -= r * D[pl]

    def swap(k, max_k):
        """Swap rows k-1 and k (P[k-1] != 0).  Returns 1 if row k-1 needs re-reduction."""
        if P[k] != 0:
            B[k - 2], B[k - 1] = B[k - 1], B[k - 2]
            U[k - 2], U[k - 1] = U[k - 1], U[k - 2]
            for j in range(1, k - 1):
                pj = P[j]
                if pj != 0:
                    lam[k - 1][pj], lam[k][pj] = lam[k][pj], lam[k - 1][pj]
            pk = P[k]
            for i in range(k + 1, max_k + 1):
                t1 = (lam[i][pk - 1] * lam[k][pk - 1] + lam[i][pk] * D[pk - 2]) // D[pk - 1]
                t2 = (lam[i][pk - 1] * D[pk] - lam[i][pk] * lam[k][pk - 1]) // D[pk - 1]
                lam[i][pk - 1] = t1
                lam[i][pk] = t2
            D[pk - 1] = (D[pk] * D[pk - 2] + lam[k][pk - 1] * lam[k][pk - 1]) // D[pk - 1]
            return 0
        pk1 = P[k - 1]
        if lam[k][pk1] != 0:
            e, x, y = _xgcd(lam[k][pk1], D[pk1])
            t1 = lam[k][pk1] // e
            t2 = D[pk1] // e
            t3 = t2
            t2 = -t2
            # (row k-1, row k) <- (t1*row(k-1) + t2*row(k), y*row(k-1) + x*row(k))
            for M, width in ((B, n), (U, m)):
                r1, r2 = M[k - 2], M[k - 1]
                for t in range(width):
                    c1, c2 = r1[t], r2[t]
                    r1[t] = t1 * c1 + t2 * c2
                    r2[t] = y * c1 + x * c2
            for j in range(1, k - 1):
                pj = P[j]
                if pj != 0:
                    c1, c2 = lam[k - 1][pj], lam[k][pj]
                    lam[k - 1][pj] = t1 * c1 + t2 * c2
                    lam[k][pj] = y * c1 + x * c2
            t2 = t2 * t2
            D[pk1] = D[pk1] // t2
            for i in range(k + 1, max_k + 1):
                pi = P[i]
                if pi != 0:
                    D[pi] = D[pi] // t2
                    for j in range(i + 1, max_k + 1):
                        lam[j][pi] = lam[j][pi] // t2
            for i in range(k + 1, max_k + 1):
                lam[i][pk1] = lam[i][pk1] // t3
            P[k - 1], P[k] = P[k], P[k - 1]
            return 1
        B[k - 2], B[k - 1] = B[k - 1], B[k - 2]
        U[k - 2], U[k - 1] = U[k - 1], U[k - 2]
        for j in range(1, k - 1):
            pj = P[j]
            if pj != 0:
                lam[k - 1][pj], lam[k][pj] = lam[k][pj], lam[k - 1][pj]
        P[k - 1], P[k] = P[k], P[k - 1]
        return 0

    k = 1
    max_k = 0
    force_reduce = 1
    while k <= m:
        if k > max_k:
            incremental_gs(k)
            max_k = k
        if k == 1:
            force_reduce = 1
            k += 1
            continue
        if force_reduce:
            for j in range(k - 1, 0, -1):
                reduce(k, j)
        if P[k - 1] != 0 and (
            P[k] == 0
            or a * D[P[k] - 1] * D[P[k] - 1] > b * (D[P[k]] * D[P[k] - 2] + lam[k][P[k] - 1] * lam[k][P[k] - 1])
        ):
            force_reduce = swap(k, max_k)
            k -= 1
        else:
            force_reduce = 1
            k += 1
    return s, B, U
