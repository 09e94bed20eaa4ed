// sound.cpp -- exact decision "does layer (u, k) meet the cube?", the question the reference puts to Yices in sound
// mode (yprune(), libblt/yices_wrapper.cc:47-157; consulted at libblt/closest.cc:258-267).
//
//     exists x in R^k :  | p_i - sum_{j>=k} u_j B_ji - sum_{j<k} x_j B_ji |  <=  l     for every coordinate i
//
// with p_i = (lo_i + hi_i) / 2 and l = (hi_0 - lo_0) / 2 (closest.cc:180-194).  Everything is doubled, which makes the
// system integral, and decided by Fourier-Motzkin elimination in exact 128-bit integer arithmetic with gcd
// normalisation.  Elimination is exact and complete but its size explodes with k: layers with more than
// BLTCU_YPRUNE_MAX_K free coefficients (or an intermediate system beyond the row limit, or a coefficient that does
// not fit 128 bits) are answered BLT_CHECK_UNKNOWN / LP_ERROR, as the reference answers a solver failure.  The GPU
// search does not need this routine: its sound mode keeps every layer whose bound is not clear of the radius
// (bltcu_opts.sound); this is the exact referee for small layers and the carrier of the reference's known answers
// (libblt/test_yices.cc:37-123).
#include <stdint.h>

#include <algorithm>
#include <vector>

#include "../../../include/bltcu.h"
#include "../../../include/codes.h"

namespace {
typedef __int128 i128;
const int MAX_ROWS = 20000;

i128 iabs(i128 x) { return x < 0 ? -x : x; }
i128 gcd128(i128 a, i128 b) {
    a = iabs(a); b = iabs(b);
    while (b) { const i128 t = a % b; a = b; b = t; }
    return a;
}
// row: c[0..k) . x <= c[k]
typedef std::vector<i128> Row;
bool mul_ok(i128 a, i128 b, i128 *out) { return !__builtin_mul_overflow(a, b, out); }
bool add_ok(i128 a, i128 b, i128 *out) { return !__builtin_add_overflow(a, b, out); }

void normalise(Row &r) {
    i128 g = 0;
    for (size_t i = 0; i + 1 < r.size(); ++i) g = gcd128(g, r[i]);
    if (g > 1) {
        for (size_t i = 0; i + 1 < r.size(); ++i) r[i] /= g;
        // floor division keeps the integer relaxation exact for the rational question as well: a.x <= b with integer a
        // after division by g is a.x/g <= b/g over the reals; keep b/g exact by scaling instead when it does not divide
        i128 &b = r.back();
        if (b % g == 0) b /= g;
        else { for (size_t i = 0; i + 1 < r.size(); ++i) r[i] *= g; }  // undo: keep the row as it was
    }
}
}  // namespace

extern "C" int bltcu_yprune(const int64_t *lat, int n, int m, const int64_t *u, int k, const int64_t *con, int *decision) {
    if (!lat || !u || !con || !decision || n <= 0 || m <= 0 || k < 0 || k > n) return LP_ERROR;
    *decision = BLT_CHECK_UNKNOWN;
    if (k > BLTCU_YPRUNE_MAX_K) return LP_ERROR;
    const i128 l2 = (i128)con[1] - (i128)con[0];  // 2 l
    std::vector<Row> sys;
    for (int i = 0; i < m; ++i) {
        i128 c2 = (i128)con[2 * i] + (i128)con[2 * i + 1];  // 2 p_i
        for (int j = k; j < n; ++j) c2 -= 2 * (i128)u[j] * (i128)lat[(size_t)j * m + i];
        // |c2 - 2 sum x_j B_ji| <= l2   <=>   -2 B.x <= l2 - c2   and   2 B.x <= l2 + c2
        Row a((size_t)k + 1), b((size_t)k + 1);
        for (int j = 0; j < k; ++j) { a[j] = -2 * (i128)lat[(size_t)j * m + i]; b[j] = 2 * (i128)lat[(size_t)j * m + i]; }
        a[k] = l2 - c2; b[k] = l2 + c2;
        normalise(a); normalise(b);
        sys.push_back(a); sys.push_back(b);
    }
    for (int v = k - 1; v >= 0; --v) {  // eliminate x_v
        std::vector<Row> pos, neg, rest;
        for (auto &r : sys) (r[v] > 0 ? pos : (r[v] < 0 ? neg : rest)).push_back(r);
        if ((long long)pos.size() * (long long)neg.size() + (long long)rest.size() > MAX_ROWS) return LP_ERROR;
        for (auto &p : pos)
            for (auto &q : neg) {
                Row r((size_t)k + 1);
                const i128 a = p[v], b = -q[v];  // b * p + a * q cancels x_v; a, b > 0
                bool ok = true;
                for (int c = 0; c <= k && ok; ++c) {
                    i128 t1, t2;
                    ok = mul_ok(b, p[c], &t1) && mul_ok(a, q[c], &t2) && add_ok(t1, t2, &r[c]);
                }
                if (!ok) return LP_ERROR;
                normalise(r);
                rest.push_back(r);
            }
        std::sort(rest.begin(), rest.end());
        rest.erase(std::unique(rest.begin(), rest.end()), rest.end());
        sys.swap(rest);
    }
    for (auto &r : sys)
        if (r[k] < 0) { *decision = BLT_CHECK_UNSAT; return LP_OK; }  // 0 <= negative
    *decision = BLT_CHECK_SAT;
    return LP_OK;
}
