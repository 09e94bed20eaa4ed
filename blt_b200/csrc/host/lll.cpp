// lll.cpp -- exact integral LLL (delta = 3/4, dependent rows allowed) behind the bltcu_lll() seam.
//
// Replaces LLL(matrix<int64>& B, matrix<int64>& U) of libblt/ntl_wrapper.cc:37-52, i.e. NTL::LLL(det2, B, U):
// the all-integer algorithm of Cohen (Alg. 2.6.7) on Gram sub-determinants D[i] and integral lam[k][j], with rows found
// dependent moved to the front.  On return the first (r - rank) rows of B are zero, the last `rank` rows are the
// reduced basis and U * B_in = B_out; entries are narrowed to int64 the way NTL's to_long() does (low 64 bits).
//
// This is the exact host implementation (own multiprecision integers; no GMP/NTL in the image): fixed-capacity
// arithmetic first (lll_fixed.h, shared with the batched device kernel), arbitrary precision when a number outgrows
// it.  The batched kernel falls back to it; it never calls into oracle/.
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <utility>
#include <vector>

#include "../../../include/bltcu.h"
#include "../lll_fixed.h"
#include "../lll_fp.h"

namespace {

// ---- minimal signed multiprecision integer (sign-magnitude, base 2^32) ---------------------------------------
struct Big {
    std::vector<uint32_t> d;  // little endian, no leading zeros
    bool neg = false;

    Big() {}
    Big(int64_t v) {
        neg = v < 0;
        uint64_t a = neg ? (uint64_t)0 - (uint64_t)v : (uint64_t)v;
        while (a) { d.push_back((uint32_t)a); a >>= 32; }
    }
    bool zero() const { return d.empty(); }
    void trim() { while (!d.empty() && d.back() == 0) d.pop_back(); if (d.empty()) neg = false; }
    int64_t to_i64() const {  // low 64 bits with sign, like NTL's to_long
        uint64_t a = 0;
        if (d.size() > 0) a |= d[0];
        if (d.size() > 1) a |= (uint64_t)d[1] << 32;
        return (int64_t)(neg ? (uint64_t)0 - a : a);
    }
};

int cmp_mag(const Big &a, const Big &b) {
    if (a.d.size() != b.d.size()) return a.d.size() < b.d.size() ? -1 : 1;
    for (size_t i = a.d.size(); i-- > 0;)
        if (a.d[i] != b.d[i]) return a.d[i] < b.d[i] ? -1 : 1;
    return 0;
}
int cmp(const Big &a, const Big &b) {
    if (a.neg != b.neg) return a.neg ? -1 : 1;
    const int c = cmp_mag(a, b);
    return a.neg ? -c : c;
}
void add_mag(const Big &a, const Big &b, Big &r) {
    const Big &x = a.d.size() >= b.d.size() ? a : b, &y = a.d.size() >= b.d.size() ? b : a;
    std::vector<uint32_t> o(x.d.size() + 1);
    uint64_t c = 0;
    for (size_t i = 0; i < x.d.size(); ++i) {
        c += (uint64_t)x.d[i] + (i < y.d.size() ? y.d[i] : 0);
        o[i] = (uint32_t)c; c >>= 32;
    }
    o[x.d.size()] = (uint32_t)c;
    r.d.swap(o);
}
void sub_mag(const Big &a, const Big &b, Big &r) {  // |a| >= |b|
    std::vector<uint32_t> o(a.d.size());
    int64_t c = 0;
    for (size_t i = 0; i < a.d.size(); ++i) {
        c += (int64_t)a.d[i] - (i < b.d.size() ? b.d[i] : 0);
        o[i] = (uint32_t)c; c >>= 32;
    }
    r.d.swap(o);
}
Big operator+(const Big &a, const Big &b) {
    Big r;
    if (a.neg == b.neg) { add_mag(a, b, r); r.neg = a.neg; }
    else {
        const int c = cmp_mag(a, b);
        if (c == 0) return Big();
        if (c > 0) { sub_mag(a, b, r); r.neg = a.neg; } else { sub_mag(b, a, r); r.neg = b.neg; }
    }
    r.trim();
    return r;
}
Big operator-(const Big &a) { Big r = a; if (!r.zero()) r.neg = !r.neg; return r; }
Big operator-(const Big &a, const Big &b) { return a + (-b); }
Big operator*(const Big &a, const Big &b) {
    if (a.zero() || b.zero()) return Big();
    Big r;
    r.d.assign(a.d.size() + b.d.size(), 0);
    for (size_t i = 0; i < a.d.size(); ++i) {
        uint64_t c = 0;
        const uint64_t ai = a.d[i];
        for (size_t j = 0; j < b.d.size(); ++j) {
            c += ai * b.d[j] + r.d[i + j];
            r.d[i + j] = (uint32_t)c; c >>= 32;
        }
        r.d[i + b.d.size()] = (uint32_t)c;
    }
    r.neg = a.neg != b.neg;
    r.trim();
    return r;
}
Big absb_(const Big &a) { Big r = a; r.neg = false; return r; }
// magnitude division (Knuth D): q = |a| / |b|, rem = |a| % |b|
void divmod_mag(const Big &a, const Big &b, Big &q, Big &rem) {
    q = Big(); rem = Big();
    if (cmp_mag(a, b) < 0) { rem.d = a.d; return; }
    if (b.d.size() == 1) {
        const uint64_t dv = b.d[0];
        q.d.assign(a.d.size(), 0);
        uint64_t r = 0;
        for (size_t i = a.d.size(); i-- > 0;) { r = (r << 32) | a.d[i]; q.d[i] = (uint32_t)(r / dv); r %= dv; }
        q.trim();
        if (r) rem.d.push_back((uint32_t)r);
        return;
    }
    const int s = __builtin_clz(b.d.back());
    const size_t n = b.d.size(), m = a.d.size() - n;
    std::vector<uint32_t> u(a.d.size() + 1), v(n);
    for (size_t i = n; i-- > 0;) v[i] = (b.d[i] << s) | ((s && i) ? (b.d[i - 1] >> (32 - s)) : 0);
    u[a.d.size()] = s ? (a.d.back() >> (32 - s)) : 0;
    for (size_t i = a.d.size(); i-- > 0;) u[i] = (a.d[i] << s) | ((s && i) ? (a.d[i - 1] >> (32 - s)) : 0);
    q.d.assign(m + 1, 0);
    for (size_t j = m + 1; j-- > 0;) {
        const uint64_t num = ((uint64_t)u[j + n] << 32) | u[j + n - 1];
        uint64_t qh = num / v[n - 1], rh = num % v[n - 1];
        while (qh >= (1ull << 32) || qh * v[n - 2] > ((rh << 32) | u[j + n - 2])) {
            --qh; rh += v[n - 1];
            if (rh >= (1ull << 32)) break;
        }
        int64_t borrow = 0;
        uint64_t carry = 0;
        for (size_t i = 0; i < n; ++i) {
            const uint64_t p = qh * v[i] + carry;
            carry = p >> 32;
            const int64_t t = (int64_t)u[i + j] - (int64_t)(uint32_t)p + borrow;
            u[i + j] = (uint32_t)t;
            borrow = t >> 32;
        }
        const int64_t t = (int64_t)u[j + n] - (int64_t)carry + borrow;
        u[j + n] = (uint32_t)t;
        if (t < 0) {  // add back
            --qh;
            uint64_t c = 0;
            for (size_t i = 0; i < n; ++i) { c += (uint64_t)u[i + j] + v[i]; u[i + j] = (uint32_t)c; c >>= 32; }
            u[j + n] += (uint32_t)c;
        }
        q.d[j] = (uint32_t)qh;
    }
    q.trim();
    rem.d.assign(n, 0);
    for (size_t i = 0; i < n; ++i) rem.d[i] = (u[i] >> s) | ((s && i + 1 <= n) ? (uint32_t)((uint64_t)u[i + 1] << (32 - s)) : 0);
    rem.trim();
}
// floor division and non-negative remainder for d > 0
void floordiv(const Big &a, const Big &d, Big &q, Big &r) {
    Big qm, rm;
    divmod_mag(a, d, qm, rm);
    if (!a.neg) { q = qm; r = rm; return; }
    if (rm.zero()) { q = -qm; r = Big(); return; }
    q = -(qm + Big(1));
    r = absb_(d) - rm;
}
Big fdiv(const Big &a, const Big &d) {  // floor(a / d), d > 0 (all divisions in the algorithm are exact or by d > 0)
    Big q, r;
    floordiv(a, d, q, r);
    return q;
}
Big absb(const Big &a) { Big r = a; r.neg = false; return r; }

// nearest integer to a/d (d > 0), ties towards zero
Big baldiv(const Big &a, const Big &d) {
    Big q, r;
    floordiv(a, d, q, r);
    const Big r2 = r + r;
    const int c = cmp(r2, d);
    if (c > 0 || (c == 0 && q.neg)) q = q + Big(1);
    return q;
}
// g = a*x + b*y >= 0 : classical Euclid on magnitudes, signs restored afterwards
void xgcd(const Big &a, const Big &b, Big &g, Big &x, Big &y) {
    Big u = absb(a), v = absb(b), u1(1), v1(0), u2(0), v2(1);
    while (!v.zero()) {
        Big q, r;
        floordiv(u, v, q, r);
        u = v; v = r;
        Big t = u1 - q * u2; u1 = u2; u2 = t;
        t = v1 - q * v2; v1 = v2; v2 = t;
    }
    if (a.neg) u1 = -u1;
    if (b.neg) v1 = -v1;
    g = u; x = u1; y = v1;
}

struct LLLState {
    int m, n, s = 0;
    std::vector<std::vector<Big>> B, U, lam;  // lam[k][j], 1-based
    std::vector<int> P;
    std::vector<Big> D;

    Big inner(int a, int b) const {
        Big r;
        for (int t = 0; t < n; ++t) r = r + B[a - 1][t] * B[b - 1][t];
        return r;
    }
    void incremental_gs(int k) {
        for (int j = 1; j < k; ++j) {
            const int posj = P[j];
            if (posj == 0) continue;
            Big u = inner(k, j);
            for (int i = 1; i < posj; ++i) u = fdiv(D[i] * u - lam[k][i] * lam[j][i], D[i - 1]);
            lam[k][posj] = u;
        }
        Big u = inner(k, k);
        for (int i = 1; i <= s; ++i) u = fdiv(D[i] * u - lam[k][i] * lam[k][i], D[i - 1]);
        if (u.zero()) P[k] = 0;
        else { ++s; P[k] = s; D[s] = u; }
    }
    void reduce(int k, int l) {
        const int pl = P[l];
        if (pl == 0) return;
        if (cmp(absb(lam[k][pl] + lam[k][pl]), D[pl]) <= 0) return;
        const Big r = baldiv(lam[k][pl], D[pl]);
        for (int t = 0; t < n; ++t) B[k - 1][t] = B[k - 1][t] - r * B[l - 1][t];
        for (int t = 0; t < m; ++t) U[k - 1][t] = U[k - 1][t] - r * U[l - 1][t];
        for (int j = 1; j < l; ++j) {
            const int pj = P[j];
            if (pj != 0) lam[k][pj] = lam[k][pj] - r * lam[l][pj];
        }
        lam[k][pl] = lam[k][pl] - r * D[pl];
    }
    int swap_rows(int k, int max_k) {
        if (P[k] != 0) {
            std::swap(B[k - 2], B[k - 1]);
            std::swap(U[k - 2], U[k - 1]);
            for (int j = 1; j <= k - 2; ++j) { const int pj = P[j]; if (pj) std::swap(lam[k - 1][pj], lam[k][pj]); }
            const int pk = P[k];
            for (int i = k + 1; i <= max_k; ++i) {
                const Big t1 = fdiv(lam[i][pk - 1] * lam[k][pk - 1] + lam[i][pk] * D[pk - 2], D[pk - 1]);
                const Big t2 = fdiv(lam[i][pk - 1] * D[pk] - lam[i][pk] * lam[k][pk - 1], D[pk - 1]);
                lam[i][pk - 1] = t1; lam[i][pk] = t2;
            }
            D[pk - 1] = fdiv(D[pk] * D[pk - 2] + lam[k][pk - 1] * lam[k][pk - 1], D[pk - 1]);
            return 0;
        }
        const int pk1 = P[k - 1];
        if (!lam[k][pk1].zero()) {
            Big e, x, y;
            xgcd(lam[k][pk1], D[pk1], e, x, y);
            const Big t1 = fdiv(lam[k][pk1], e);
            Big t2 = fdiv(D[pk1], e);
            const Big t3 = t2;
            t2 = -t2;
            auto rowtr = [&](std::vector<Big> &r1, std::vector<Big> &r2) {
                for (size_t t = 0; t < r1.size(); ++t) {
                    const Big c1 = r1[t], c2 = r2[t];
                    r1[t] = t1 * c1 + t2 * c2;
                    r2[t] = y * c1 + x * c2;
                }
            };
            rowtr(B[k - 2], B[k - 1]);
            rowtr(U[k - 2], U[k - 1]);
            for (int j = 1; j <= k - 2; ++j) {
                const int pj = P[j];
                if (!pj) continue;
                const Big c1 = lam[k - 1][pj], c2 = lam[k][pj];
                lam[k - 1][pj] = t1 * c1 + t2 * c2;
                lam[k][pj] = y * c1 + x * c2;
            }
            const Big t2s = t2 * t2;
            D[pk1] = fdiv(D[pk1], t2s);
            for (int i = k + 1; i <= max_k; ++i) {
                const int pi = P[i];
                if (!pi) continue;
                D[pi] = fdiv(D[pi], t2s);
                for (int j = i + 1; j <= max_k; ++j) lam[j][pi] = fdiv(lam[j][pi], t2s);
            }
            for (int i = k + 1; i <= max_k; ++i) lam[i][pk1] = fdiv(lam[i][pk1], t3);
            std::swap(P[k - 1], P[k]);
            return 1;
        }
        std::swap(B[k - 2], B[k - 1]);
        std::swap(U[k - 2], U[k - 1]);
        for (int j = 1; j <= k - 2; ++j) { const int pj = P[j]; if (pj) std::swap(lam[k - 1][pj], lam[k][pj]); }
        std::swap(P[k - 1], P[k]);
        return 0;
    }
    int run(int a, int b) {
        int k = 1, max_k = 0, force_reduce = 1;
        const Big A(a), Bq(b);
        while (k <= m) {
            if (k > max_k) { incremental_gs(k); max_k = k; }
            if (k == 1) { force_reduce = 1; ++k; continue; }
            if (force_reduce) for (int j = k - 1; j >= 1; --j) reduce(k, j);
            bool do_swap = false;
            if (P[k - 1] != 0) {
                if (P[k] == 0) do_swap = true;
                else {
                    const int pk = P[k];
                    const Big lhs = A * D[pk - 1] * D[pk - 1];
                    const Big rhs = Bq * (D[pk] * D[pk - 2] + lam[k][pk - 1] * lam[k][pk - 1]);
                    do_swap = cmp(lhs, rhs) > 0;
                }
            }
            if (do_swap) { force_reduce = swap_rows(k, max_k); --k; }
            else { force_reduce = 1; ++k; }
        }
        return s;
    }
};
}  // namespace

// LLL(B, U): libblt/ntl_wrapper.h:20, libblt/ntl_wrapper.cc:37-52.  B: r x c row-major, reduced in place; U: r x r out
// (may be NULL).  *rank receives the rank.  Returns BLT_STATUS_OK or BLT_STATUS_INPUT_ERROR.
// the arbitrary-precision path on its own (also a test seam: the fixed-capacity code is pinned against it)
extern "C" int bltcu_lll_arbitrary(int64_t *B, int64_t *U, int r, int c, int *rank);

// First the fixed-capacity arithmetic of lll_fixed.h with 512-bit, then with 1024-bit numbers (no allocation per
// operation: about twice as fast as the vectors below); it reports an overflow instead of truncating, and only then the
// arbitrary-precision integers take over.  All three perform the same operations, so the result does not depend on
// which one finished.
extern "C" int bltcu_lll(int64_t *B, int64_t *U, int r, int c, int *rank) {
    if (!B || r <= 0 || c <= 0 || !rank) return BLTCU_STATUS_INPUT_ERROR;
    // Fast pass (lll_fp.h): the same sequence of row operations with every decision taken in fp64 against a margin and
    // every row operation in exact int64; it either completes with the integral algorithm's output or defers, leaving
    // B untouched (independent rows only).  60 x 60 with entries up to 100: 5 ms against 90 ms for the passes below.
    if (r <= c && !getenv("BLTCU_LLL_EXACT_ONLY")) {
        std::vector<int64_t> b(B, B + (size_t)r * c), u((size_t)r * r, 0);
        std::vector<double> work(lllfp::scratch_doubles(r));
        if (lllfp::reduce_one(b.data(), u.data(), r, c, work.data(), lllfp::HostLanes()) == r) {
            memcpy(B, b.data(), b.size() * 8);
            if (U) memcpy(U, u.data(), u.size() * 8);
            *rank = r;
            return BLTCU_STATUS_OK;
        }
    }
    if ((long)r * c <= 128 * 128) {
        std::vector<int64_t> b((size_t)r * c), u((size_t)r * r);
        for (int pass = 0; pass < 2; ++pass) {
            const long bytes = pass == 0 ? lllfix::bytes_needed<16>(r, c) : lllfix::bytes_needed<32>(r, c);
            void *work = malloc((size_t)bytes);
            if (!work) break;
            std::copy(B, B + (size_t)r * c, b.begin());
            const int rk = pass == 0 ? lllfix::reduce_one<16, lllfix::OneLane>(b.data(), u.data(), r, c, work)
                                     : lllfix::reduce_one<32, lllfix::OneLane>(b.data(), u.data(), r, c, work);
            free(work);
            if (rk >= 0) {
                memcpy(B, b.data(), b.size() * 8);
                if (U) memcpy(U, u.data(), u.size() * 8);
                *rank = rk;
                return BLTCU_STATUS_OK;
            }
        }
    }
    return bltcu_lll_arbitrary(B, U, r, c, rank);
}

extern "C" int bltcu_lll_arbitrary(int64_t *B, int64_t *U, int r, int c, int *rank) {
    if (!B || r <= 0 || c <= 0 || !rank) return BLTCU_STATUS_INPUT_ERROR;
    LLLState S;
    S.m = r; S.n = c;
    S.B.assign(r, std::vector<Big>(c));
    S.U.assign(r, std::vector<Big>(r));
    for (int i = 0; i < r; ++i) {
        for (int j = 0; j < c; ++j) S.B[i][j] = Big(B[(size_t)i * c + j]);
        S.U[i][i] = Big(1);
    }
    S.P.assign(r + 1, 0);
    S.D.assign(r + 1, Big());
    S.D[0] = Big(1);
    S.lam.assign(r + 1, std::vector<Big>(r + 1));
    *rank = S.run(3, 4);
    for (int i = 0; i < r; ++i) {
        for (int j = 0; j < c; ++j) B[(size_t)i * c + j] = S.B[i][j].to_i64();
        if (U) for (int j = 0; j < r; ++j) U[(size_t)i * r + j] = S.U[i][j].to_i64();
    }
    return BLTCU_STATUS_OK;
}

// Test seam (not in include/bltcu.h): the fixed-capacity implementation that the batched device kernel runs, executed
// on the host, so that its arithmetic can be pinned against the arbitrary-precision one without a GPU.
// limbs: 8, 16 or 32 (x 32 bits).  Returns BLT_STATUS_ERROR when a number did not fit.
extern "C" int bltcu_lll_fixed_host(int64_t *B, int64_t *U, int r, int c, int *rank, int limbs) {
    if (!B || r <= 0 || c <= 0 || !rank) return BLTCU_STATUS_INPUT_ERROR;
    long bytes = 0;
    switch (limbs) {
    case 8: bytes = lllfix::bytes_needed<8>(r, c); break;
    case 16: bytes = lllfix::bytes_needed<16>(r, c); break;
    case 32: bytes = lllfix::bytes_needed<32>(r, c); break;
    default: return BLTCU_STATUS_INPUT_ERROR;
    }
    void *work = malloc((size_t)bytes);
    if (!work) return BLTCU_STATUS_ERROR;
    std::vector<int64_t> b(B, B + (size_t)r * c), u((size_t)r * r);
    int rk = -1;
    if (limbs == 8) rk = lllfix::reduce_one<8, lllfix::OneLane>(b.data(), u.data(), r, c, work);
    else if (limbs == 16) rk = lllfix::reduce_one<16, lllfix::OneLane>(b.data(), u.data(), r, c, work);
    else rk = lllfix::reduce_one<32, lllfix::OneLane>(b.data(), u.data(), r, c, work);
    free(work);
    if (rk < 0) return BLTCU_STATUS_ERROR;
    memcpy(B, b.data(), b.size() * 8);
    if (U) memcpy(U, u.data(), u.size() * 8);
    *rank = rk;
    return BLTCU_STATUS_OK;
}
