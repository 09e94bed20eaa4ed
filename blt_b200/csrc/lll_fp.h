// lll_fp.h -- the exact integral LLL of bltcu_lll() (Cohen, Algorithm 2.6.7, delta = 3/4; what NTL::LLL(det2, B, U) runs,
// libblt/ntl_wrapper.cc:37-52) with its DECISIONS taken in double precision and its row operations in exact int64.
//
// The integral algorithm decides two things from its multiprecision state: whether (and by which integer) row k is
// reduced against row l -- a function of mu_kl = lam_kl / D_l -- and whether rows k-1, k are swapped -- the Lovasz test
// (3/4 - mu^2) |b*_{k-1}|^2 > |b*_k|^2.  Both are functions of the Gram-Schmidt data of the CURRENT integer basis.
// Here that data is recomputed in fp64 for the row being visited, every time it is visited, from exact integer inner
// products (Cholesky recurrence against the rows below, which do not change while the row is worked on), so no
// floating-point state is ever carried through a swap.  Every decision must clear a margin (1e-7 relative, several
// orders above the rounding error of the recurrence on the well-conditioned problems this path is for); a decision
// inside the margin, an inner product beyond 2^53, an entry beyond 2^62, a row that looks dependent or a step limit
// defer the WHOLE problem to the multiprecision routine (return value -1, B and U untouched by then are restored by
// the caller from its copy).  When it completes, the sequence of integer row operations is the integral algorithm's, so
// B, U and the rank are its output bit for bit.  Full row rank only: dependent generating sets (config 2) always
// defer.
//
// One source for host and device: `Lanes` deals loop iterations to the lanes of a warp (or to the one host "lane") and
// provides the sum over lanes.
#pragma once
#include <math.h>
#include <stdint.h>

#if defined(__CUDACC__)
#define LLLFP_HD __host__ __device__
#else
#define LLLFP_HD
#endif

namespace lllfp {

struct HostLanes {
    LLLFP_HD int lane() const { return 0; }
    LLLFP_HD int width() const { return 1; }
    LLLFP_HD void sync() const {}
    LLLFP_HD double sum(double x) const { return x; }
    LLLFP_HD int any(int x) const { return x; }
    LLLFP_HD double bcast(double x, int) const { return x; }
};
#if defined(__CUDACC__)
struct WarpLanes {
    __device__ int lane() const { return (int)(threadIdx.x & 31); }
    __device__ int width() const { return 32; }
    __device__ void sync() const { __syncwarp(); }
    __device__ double sum(double x) const {
        for (int off = 16; off >= 1; off >>= 1) x += __shfl_xor_sync(0xffffffffu, x, off);
        return x;
    }
    __device__ int any(int x) const { return __any_sync(0xffffffffu, x); }
    __device__ double bcast(double x, int src) const { return __shfl_sync(0xffffffffu, x, src); }
};
#endif

#define LLLFP_MARGIN 1.0e-7
#define LLLFP_MAX_ENTRY 4.0e18   // |entry| of B or U after an update must stay below this (int64 holds 9.2e18)

// doubles of scratch per problem: mu (r x r), rr (r), d (r), g (r)
LLLFP_HD inline size_t scratch_doubles(int r) { return (size_t)r * r + 3 * (size_t)r; }

// B: r x c, U: r x r (identity on entry is written here), work: scratch_doubles(r).  Returns r (the rank) or -1.
template <class Lanes>
LLLFP_HD int reduce_one(int64_t *B, int64_t *U, int r, int c, double *work, const Lanes &L) {
    const int lane = L.lane(), W = L.width();
    double *mu = work, *rr = mu + (size_t)r * r, *d = rr + r, *g = d + r;
    for (int e = lane; e < r * r; e += W) { U[e] = (e / r == e % r) ? 1 : 0; mu[e] = 0.0; }
    L.sync();
    int k = 1, max_k = 0, force_reduce = 1;   // 1-based like the integral routine
    long steps = 0;
    const long step_cap = 200000L + 2000L * r * r;
    while (k <= r) {
        if (++steps > step_cap) return -1;
        // ---- Gram-Schmidt data of row k against rows 1..k-1, from exact inner products ----
        const int64_t *bk = B + (size_t)(k - 1) * c;
        int bad = 0;
        for (int j = lane; j < k; j += W) {   // g[j] = <b_k, b_{j+1}>, j = 0..k-1 (the last one is <b_k, b_k>)
            const int64_t *bj = B + (size_t)j * c;
            double chk = 0.0;
            long long acc = 0;
            for (int t = 0; t < c; ++t) { acc += bk[t] * bj[t]; chk += fabs((double)bk[t] * (double)bj[t]); }
            if (chk > 9.0e15) bad = 1;      // 2^53: the sum would not be exact as a double (nor, further up, as an int64)
            g[j] = (double)acc;
        }
        if (L.any(bad)) return -1;
        L.sync();
        for (int j = 0; j < k - 1; ++j) {     // rr[j] = g[j] - sum_{i<j} mu[j][i] rr[i];  mu[k][j] = rr[j] / d[j]
            double part = 0.0;
            for (int i = lane; i < j; i += W) part = fma(mu[(size_t)j * r + i], rr[i], part);
            const double s = L.sum(part);
            if (lane == 0) { rr[j] = g[j] - s; mu[(size_t)(k - 1) * r + j] = rr[j] / d[j]; }
            L.sync();
        }
        {
            double part = 0.0;
            for (int j = lane; j < k - 1; j += W) part = fma(mu[(size_t)(k - 1) * r + j], rr[j], part);
            const double s = L.sum(part);
            const double dk = g[k - 1] - s;
            // a dependent row has |b*_k|^2 = 0: anything that small next to |b_k|^2 is left to the exact routine
            if (!(dk > 1.0e-6 * g[k - 1]) || !(g[k - 1] > 0.0)) return -1;
            if (lane == 0) d[k - 1] = dk;
            L.sync();
        }
        if (k > max_k) max_k = k;
        if (k == 1) { force_reduce = 1; k = 2; continue; }
        // ---- size reduction of row k against rows k-1 .. 1 ----
        if (force_reduce) {
            for (int l = k - 1; l >= 1; --l) {
                const double x = mu[(size_t)(k - 1) * r + (l - 1)];
                const double ax = fabs(x);
                if (!(ax < 2.0e9)) return -1;
                const double m = LLLFP_MARGIN * (1.0 + ax);
                if (ax <= 0.5 - m) continue;                  // 2 |lam| <= D: nothing to do
                if (ax < 0.5 + m) return -1;                  // too close to the tie to call
                const double q = rint(x);                     // nearest integer; a tie is excluded by the margin below
                if (fabs(x - q) > 0.5 - m) return -1;
                const long long qi = (long long)q;
                int64_t *rowk = B + (size_t)(k - 1) * c;
                const int64_t *rowl = B + (size_t)(l - 1) * c;
                int over = 0;
                for (int t = lane; t < c; t += W) {
                    if (fabs((double)rowk[t] - q * (double)rowl[t]) > LLLFP_MAX_ENTRY) over = 1;
                    rowk[t] -= qi * rowl[t];
                }
                int64_t *uk = U + (size_t)(k - 1) * r;
                const int64_t *ul = U + (size_t)(l - 1) * r;
                for (int t = lane; t < r; t += W) {
                    if (fabs((double)uk[t] - q * (double)ul[t]) > LLLFP_MAX_ENTRY) over = 1;
                    uk[t] -= qi * ul[t];
                }
                if (L.any(over)) return -1;
                for (int j = lane; j < l - 1; j += W) mu[(size_t)(k - 1) * r + j] -= q * mu[(size_t)(l - 1) * r + j];
                if (lane == 0) mu[(size_t)(k - 1) * r + (l - 1)] = x - q;
                L.sync();
            }
            // |b*_k|^2 does not change under size reduction; mu[k][k-1] did: the Lovasz test below uses the new one
        }
        // ---- Lovasz test, 3 D_{k-1}^2 > 4 (D_k D_{k-2} + lam^2)  <=>  (3/4 - mu^2) d_{k-1} > d_k ----
        const double m1 = mu[(size_t)(k - 1) * r + (k - 2)];
        const double lhs = (0.75 - m1 * m1) * d[k - 2], rhs = d[k - 1];
        if (fabs(lhs - rhs) <= LLLFP_MARGIN * (fabs(lhs) + fabs(rhs))) return -1;
        if (lhs > rhs) {
            int64_t *r1 = B + (size_t)(k - 2) * c, *r2 = B + (size_t)(k - 1) * c;
            for (int t = lane; t < c; t += W) { const int64_t x = r1[t]; r1[t] = r2[t]; r2[t] = x; }
            int64_t *u1 = U + (size_t)(k - 2) * r, *u2 = U + (size_t)(k - 1) * r;
            for (int t = lane; t < r; t += W) { const int64_t x = u1[t]; u1[t] = u2[t]; u2[t] = x; }
            L.sync();
            force_reduce = 0;   // the integral routine does not re-reduce the row it has just moved down
            k -= 1;
        } else {
            force_reduce = 1;
            k += 1;
        }
    }
    return r;
}

}  // namespace lllfp
