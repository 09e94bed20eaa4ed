// search.cu -- sm_100a implementation of BLT's bounded lattice-point search (closest(), libblt/closest.cc:166-327)
// and of its per-node layer-distance solve (infdist(), libblt/glpk_wrapper.cc:33-232).
//
// Execution model (DESIGN.md sections 3-5):
//   * ONE persistent kernel per search, one CTA per SM, W warps per CTA.  A warp owns one depth-first search of
//     one subtree at a time; its 32 lanes cooperate on that subtree's current node.
//   * A node is a small dense LP  min_x |G x - c|_inf  (G = first kk basis vectors as columns, m rows).  It is
//     solved by an exchange method on reference sets: K = kk+1 active rows with signs, the K x K inverse H of
//     [sigma | G_S] in shared memory, dual feasibility kept as the invariant, so t = lambda . c_S is a valid lower
//     bound at every step and a node is pruned as soon as the bound crosses the radius.
//   * Warm starts ("side chains"): a node starts from the final basis of the previous sibling on the same side of
//     the level's real centre, or, for the first node of a side, from its parent's optimal basis with the level's
//     coefficient fixed.  Both live in a per-warp level store in global memory (L2 resident).
//   * Load balance: a global MPMC queue of subtree records.  A warp that explores a child while other warps are
//     idle writes the child as a record (its optimal basis, right-hand side, path) instead of descending.  Every
//     node is still solved exactly once, from the same starting basis as in the sequential order, so niter and
//     nprune do not depend on scheduling.
//
// Floating point: every operation below has a twin in oracle/blt_oracle.c with the same operand order; compile
// with -fmad=false so that only the explicit fma() calls fuse.  FP64 FMA pipe is the roofline (no tensor cores:
// nothing here is a dense contraction worth a tcgen05 tile).
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <chrono>
#include <mutex>
#include <vector>
#include <algorithm>

#include "../../include/bltcu.h"

#define FULLMASK 0xffffffffu
#define DIST_EPS 1.0e-4  // libblt/glpk_wrapper.h:13
#define EPS_OPT 1.0e-9
#define PIV_TOL 1.0e-11
#define MAXD BLTCU_MAX_DIM

enum { NODE_EXPLORE = 0, NODE_PRUNE = 1, NODE_FAIL = -1 };
enum { REC_ROOT = 1 };

// ------------------------------------------------------------------------------------------------------------
// Global control block (one per search, zeroed by the host before launch)
// ------------------------------------------------------------------------------------------------------------
struct QCtl {
    unsigned int head, tail;    // work ring: claimed / reserved tickets
    unsigned int fhead, ftail;  // free-slot ring
    int pending;                // records queued or being processed
    int idle;                   // warps waiting for work
    int abort;                  // stop everything (solution found with early exit, or failure)
    int status;                 // BLT status code of the search
    int found;                  // a solution has been written to sol
    unsigned int out_count;     // static split: number of collected subtrees
    int timed_out;              // watchdog fired: 1 node loop, 2 idle loop, 3 pop spin, 4 push spin
    int pad1;
    unsigned long long deadline_ns;  // %globaltimer value after which every loop gives up (0 = none)
    unsigned long long niter, nprune, nodes, pivots, quick, refactors, colds, donated;
    unsigned long long hist[MAXD + 1];
};

struct Params {
    int n, m, LD, LDB, KMAX;
    const double *B;        // n x m (double)lat
    const double *p;        // m centre
    double radius;
    const int64_t *lat;     // n x m
    const int64_t *con;     // m x 2
    QCtl *q;
    int *ring;              // work ring (record indices, -1 = empty)
    int *fring;             // free-slot ring
    unsigned int ring_mask; // capacity - 1
    unsigned char *pool;    // subtree records
    size_t rec_stride;
    int pool_cap;
    double *scrD;           // per-warp level store (doubles)
    size_t scrD_stride;
    int *scrI;              // per-warp level store (ints)
    size_t scrI_stride;
    int64_t *sol;           // n, written by the finder
    int64_t *best_u;        // per warp: n (last recorded candidate)
    unsigned char *best_key;// per warp: n
    int *best_valid;        // per warp
    int *out_list;          // static split output (record indices)
    int static_depth;       // records shallower than this emit every child; children at this depth are collected
    int dynamic;            // 1 = donate to idle warps
    int donate_min_level;
    int lowwater;
    int no_early_exit;
    int warps_per_cta;
    size_t warp_smem_doubles;
};

// record field offsets (bytes), computed identically on host and device
struct RecLayout {
    size_t off_u, off_ord, off_cpar, off_row, off_sg, off_H, stride;
};
__host__ __device__ inline RecLayout rec_layout(int n, int m) {
    RecLayout L;
    size_t o = 32;  // header: int k, int flags, double xc, pad
    L.off_u = o; o += (size_t)8 * n;
    L.off_ord = o; o += ((size_t)n + 7) / 8 * 8;
    L.off_cpar = o; o += (size_t)8 * m;
    L.off_row = o; o += ((size_t)4 * (n + 1) + 7) / 8 * 8;
    L.off_sg = o; o += ((size_t)4 * (n + 1) + 7) / 8 * 8;
    L.off_H = o; o += (size_t)8 * (n + 1) * (n + 1);
    L.stride = (o + 255) / 256 * 256;
    return L;
}
// level store: chains for k = 2..n (K = k), two sides; packed
__host__ __device__ inline size_t chain_off(int k) {  // doubles before level k: 2 * sum_{j=2}^{k-1} j^2
    const size_t kk = (size_t)k;
    return 2 * ((kk - 1) * kk * (2 * kk - 1) / 6 - 1);
}
__host__ __device__ inline size_t chain_ioff(int k) {  // ints before level k: 2 sides * (row,sg) * sum j
    const size_t kk = (size_t)k;
    return 4 * ((kk - 1) * kk / 2 - 1);
}

// ------------------------------------------------------------------------------------------------------------
// Per-warp state (pointers into shared memory) and primitives
// ------------------------------------------------------------------------------------------------------------
struct WS {
    double *W, *z, *g, *w, *a, *cS, *c, *cpar, *xc;
    int *row, *sg, *tmpi, *tmps, *step;
    long long *u;
    unsigned char *ord, *flags;  // flags[k]: bit0/1 = chain initialised side 0/1, bit2 = pruned, bit3 = curside
    const double *Bs;
    int LD, LDB, n, m, K, lane;
    unsigned long long pivots, quick, refactors, colds;
};

__device__ __forceinline__ bool approxLE(double a, double b) {  // libblt/util.cc:17-31
    const double mx = (fabs(a) < fabs(b) ? fabs(b) : fabs(a)) * DIST_EPS;
    return ((b - a) > mx) || (fabs(a - b) <= mx);
}
__device__ __forceinline__ int sgn_d(double z) { return z > 0 ? 1 : -1; }  // closest.cc:343-346

template <class FA, class FB>
__device__ __forceinline__ double dot4(int len, FA fa, FB fb) {
    double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;
    int e = 0;
    for (; e + 3 < len; e += 4) {
        a0 = fma(fa(e), fb(e), a0);
        a1 = fma(fa(e + 1), fb(e + 1), a1);
        a2 = fma(fa(e + 2), fb(e + 2), a2);
        a3 = fma(fa(e + 3), fb(e + 3), a3);
    }
    if (e < len) { a0 = fma(fa(e), fb(e), a0); ++e; }
    if (e < len) { a1 = fma(fa(e), fb(e), a1); ++e; }
    if (e < len) { a2 = fma(fa(e), fb(e), a2); }
    return (a0 + a1) + (a2 + a3);
}

__device__ __forceinline__ double warp_max(double v) {
    for (int off = 16; off >= 1; off >>= 1) { const double o = __shfl_xor_sync(FULLMASK, v, off); v = o > v ? o : v; }
    return v;
}
// (largest value, smallest index); idx < 0 never wins
__device__ __forceinline__ void warp_argmax(double &v, int &idx, double &aux) {
    for (int off = 16; off >= 1; off >>= 1) {
        const double ov = __shfl_xor_sync(FULLMASK, v, off);
        const int oi = __shfl_xor_sync(FULLMASK, idx, off);
        const double oa = __shfl_xor_sync(FULLMASK, aux, off);
        const bool take = (oi >= 0) && (idx < 0 || ov > v || (ov == v && oi < idx));
        if (take) { v = ov; idx = oi; aux = oa; }
    }
}
// (smallest value, smallest index); idx < 0 never wins
__device__ __forceinline__ void warp_argmin(double &v, int &idx) {
    for (int off = 16; off >= 1; off >>= 1) {
        const double ov = __shfl_xor_sync(FULLMASK, v, off);
        const int oi = __shfl_xor_sync(FULLMASK, idx, off);
        const bool take = (oi >= 0) && (idx < 0 || ov < v || (ov == v && oi < idx));
        if (take) { v = ov; idx = oi; }
    }
}

__device__ __forceinline__ void lp_gather(WS &s) {
    for (int e = s.lane; e < s.K; e += 32) s.cS[e] = s.row[e] >= 0 ? s.c[s.row[e]] : 0.0;
    __syncwarp();
}
__device__ __forceinline__ double lp_treedot(WS &s) {  // lambda . cS with the oracle's treedot order
    double acc = 0.0;
    for (int e = s.lane; e < s.K; e += 32) acc = fma(s.W[e], s.cS[e], acc);
    for (int off = 16; off >= 1; off >>= 1) acc = acc + __shfl_xor_sync(FULLMASK, acc, off);
    return acc;
}
__device__ __forceinline__ void lp_point(WS &s) {
    const int K = s.K;
    for (int v = s.lane; v < K; v += 32) {
        const double *Hv = s.W + (size_t)v * s.LD;
        const double *cS = s.cS;
        s.z[v] = dot4(K, [&](int e) { return Hv[e]; }, [&](int e) { return cS[e]; });
    }
    __syncwarp();
}
__device__ __forceinline__ void lp_row(WS &s, int q, int sq) {
    if (s.lane == 0) s.a[0] = (double)sq;
    for (int i = s.lane; i + 1 < s.K; i += 32) s.a[i + 1] = s.Bs[(size_t)i * s.LDB + q];
    __syncwarp();
}
__device__ __forceinline__ void lp_w(WS &s) {
    const int K = s.K, LD = s.LD;
    for (int c = s.lane; c < K; c += 32) {
        const double *a = s.a;
        const double *Hc = s.W + c;
        s.w[c] = dot4(K, [&](int e) { return a[e]; }, [&](int e) { return Hc[(size_t)e * LD]; });
    }
    __syncwarp();
}
// H <- H - g (w - e_l)^T with g = H[:,l] / w_l, for rows v < nrows
__device__ __forceinline__ void lp_update(WS &s, int l, int nrows, bool set_col) {
    const int K = s.K, LD = s.LD;
    const double wl = s.w[l];
    for (int v = s.lane; v < nrows; v += 32) s.g[v] = s.W[(size_t)v * LD + l] / wl;
    __syncwarp();
    for (int c = s.lane; c < K; c += 32) {
        const double wc = s.w[c];
        double *Hc = s.W + c;
        if (c == l) {
            if (set_col) for (int v = 0; v < nrows; ++v) Hc[(size_t)v * LD] = s.g[v];
        } else {
#pragma unroll 4
            for (int v = 0; v < nrows; ++v) Hc[(size_t)v * LD] = fma(-s.g[v], wc, Hc[(size_t)v * LD]);
        }
    }
    __syncwarp();
    s.pivots++;
}
__device__ __forceinline__ void lp_pivot(WS &s, int l) { lp_update(s, l, s.K, true); }

__device__ __forceinline__ int lp_ratio(WS &s, int sq) {
    const int K = s.K;
    double wm = 0.0;
    for (int e = s.lane; e < K; e += 32) { const double a = fabs(s.w[e]); wm = a > wm ? a : wm; }
    wm = warp_max(wm);
    const double thr = PIV_TOL * wm;
    int best = -1;
    double br = 0.0;
    for (int e = s.lane; e < K; e += 32) {
        const double den = (double)(sq * s.sg[e]) * s.w[e];
        if (!(den > thr)) continue;
        double num = (double)s.sg[e] * s.W[e];
        if (num < 0.0) num = 0.0;
        const double r = num / den;
        if (best < 0 || r < br) { best = e; br = r; }
    }
    warp_argmin(br, best);
    return best;
}

__device__ __forceinline__ void lp_identity(WS &s, int K) {
    const int LD = s.LD;
    for (int e = s.lane; e < K * K; e += 32) { const int v = e / K, c = e - v * K; s.W[(size_t)v * LD + c] = (v == c) ? 1.0 : 0.0; }
    __syncwarp();
}

__device__ int lp_refactor(WS &s) {
    const int K = s.K;
    for (int e = s.lane; e < K; e += 32) { s.tmpi[e] = s.row[e]; s.tmps[e] = s.sg[e]; s.row[e] = -1; s.sg[e] = 1; }
    __syncwarp();
    lp_identity(s, K);
    for (int i = 0; i < K; ++i) {
        lp_row(s, s.tmpi[i], s.tmps[i]);
        lp_w(s);
        int l = -1; double bw = 0.0, aux = 0.0;
        for (int e = s.lane; e < K; e += 32) if (s.row[e] < 0 && fabs(s.w[e]) > bw) { bw = fabs(s.w[e]); l = e; }
        warp_argmax(bw, l, aux);
        if (l < 0) return -1;
        lp_pivot(s, l);
        if (s.lane == 0) { s.row[l] = s.tmpi[i]; s.sg[l] = s.tmps[i]; }
        __syncwarp();
    }
    s.refactors++;
    return 0;
}

__device__ __forceinline__ double residual(const WS &s, int j) {
    const double *Bj = s.Bs + j;
    const double *z1 = s.z + 1;
    const int LDB = s.LDB;
    return s.c[j] - dot4(s.K - 1, [&](int e) { return Bj[(size_t)e * LDB]; }, [&](int e) { return z1[e]; });
}

// cold start for LP(K-1); returns 0 ok, 1 square system (stage A only), -1 failure.  `used` aliases tmpi.
__device__ int lp_cold(WS &s, int K) {
    const int LD = s.LD, m = s.m, kk = K - 1, LDB = s.LDB;
    s.K = K; s.colds++;
    for (int e = s.lane; e < K; e += 32) { s.row[e] = -1; s.sg[e] = 1; }
    unsigned long long used0 = 0, used1 = 0;  // bitmask over m <= 128 rows, replicated in every lane
    lp_identity(s, K);
    for (int v = 1; v <= kk; ++v) {
        int q = -1; double bp = 0.0, aux = 0.0;
        const double *Hv = s.W + v;
        for (int j = s.lane; j < m; j += 32) {
            const bool u = j < 64 ? ((used0 >> j) & 1ull) : ((used1 >> (j - 64)) & 1ull);
            if (u) continue;
            const double *Bj = s.Bs + j;
            const double pv = fabs(dot4(K, [&](int e) { return e == 0 ? 1.0 : Bj[(size_t)(e - 1) * LDB]; },
                                        [&](int e) { return Hv[(size_t)e * LD]; }));
            if (pv > bp) { bp = pv; q = j; }
        }
        warp_argmax(bp, q, aux);
        if (q < 0) return -1;
        lp_row(s, q, 1); lp_w(s); lp_pivot(s, v);
        if (s.lane == 0) { s.row[v] = q; s.sg[v] = 1; }
        if (q < 64) used0 |= 1ull << q; else used1 |= 1ull << (q - 64);
        __syncwarp();
    }
    lp_gather(s); lp_point(s);
    if (kk >= m) return 1;
    {
        int q = -1; double br = -1.0, aux = 0.0;
        for (int j = s.lane; j < m; j += 32) {
            const bool u = j < 64 ? ((used0 >> j) & 1ull) : ((used1 >> (j - 64)) & 1ull);
            if (u) continue;
            const double r = fabs(residual(s, j));
            if (r > br) { br = r; q = j; }
        }
        warp_argmax(br, q, aux);
        if (q < 0) return -1;
        const double *Bq = s.Bs + q;
        const double *H0 = s.W + LD;
        const double rho = dot4(kk, [&](int e) { return Bq[(size_t)e * LDB]; }, [&](int e) { return H0[(size_t)e * LD]; });
        const int sq = (rho >= 0.0) ? 1 : -1;
        lp_row(s, q, sq); lp_w(s);
        if (s.w[0] == 0.0) return -1;
        lp_pivot(s, 0);
        if (s.lane == 0) { s.row[0] = q; s.sg[0] = sq; }
        __syncwarp();
    }
    {   // sign fix (Sherman-Morrison on the sigma column); lam -> a, Delta -> w, sigma' -> cS
        bool any = false;
        for (int e = s.lane; e < K; e += 32) {
            const double l = s.W[e];
            s.a[e] = l;
            const int ns = l > 0.0 ? 1 : (l < 0.0 ? -1 : s.sg[e]);
            s.w[e] = (double)(ns - s.sg[e]);
            s.cS[e] = (double)ns;
            s.tmps[e] = ns;
            if (ns != s.sg[e]) any = true;
        }
        __syncwarp();
        if (__any_sync(FULLMASK, any)) {
            const double *lam = s.a, *sp = s.cS, *dl = s.w;
            const double D = dot4(K, [&](int e) { return lam[e]; }, [&](int e) { return sp[e]; });
            if (!(D > 0.0)) return -1;
            for (int v = s.lane; v < K; v += 32) {
                const double *Hv = s.W + (size_t)v * LD;
                s.g[v] = dot4(K, [&](int e) { return Hv[e]; }, [&](int e) { return dl[e]; }) / D;
            }
            __syncwarp();
            for (int c = s.lane; c < K; c += 32) {
                const double lc = lam[c];
                double *Hc = s.W + c;
                for (int v = 0; v < K; ++v) Hc[(size_t)v * LD] = fma(-s.g[v], lc, Hc[(size_t)v * LD]);
            }
            for (int e = s.lane; e < K; e += 32) s.sg[e] = s.tmps[e];
            __syncwarp();
        }
    }
    return 0;
}

__device__ __forceinline__ void lp_flip(WS &s) {
    for (int e = s.lane; e < s.K; e += 32) { s.W[e] = -s.W[e]; s.sg[e] = -s.sg[e]; }
    if (s.lane == 0) s.z[0] = -s.z[0];
    __syncwarp();
}

// node solve from the current dual-feasible basis; radius = +inf disables pruning
__device__ int lp_solve(WS &s, double radius, double &d_out, double &xlast_out) {
    const int K = s.K, m = s.m;
    lp_gather(s);
    {
        double tq = lp_treedot(s);
        if (tq < 0.0) tq = -tq;
        if (!approxLE(tq, radius)) { s.quick++; return NODE_PRUNE; }
    }
    lp_point(s);
    if (s.z[0] < 0.0) lp_flip(s);
    double cmax = 0.0;
    for (int j = s.lane; j < m; j += 32) { const double a = fabs(s.c[j]); cmax = a > cmax ? a : cmax; }
    cmax = warp_max(cmax);
    const int cap = 100 * K + 1000;
    bool refactored = false;
    for (int iter = 0;; ++iter) {
        const double t = s.z[0];
        if (!approxLE(t, radius)) return NODE_PRUNE;
        int q = -1; double rmax = -1.0, rq = 0.0;
        for (int j = s.lane; j < m; j += 32) {
            const double r = residual(s, j);
            if (fabs(r) > rmax) { rmax = fabs(r); q = j; rq = r; }
        }
        warp_argmax(rmax, q, rq);
        const double tol = EPS_OPT * (cmax + t);
        if (rmax <= t + tol) { d_out = t; xlast_out = s.z[K - 1]; return NODE_EXPLORE; }
        if (iter >= cap) return NODE_FAIL;
        bool dup = false;
        for (int e = s.lane; e < K; e += 32) if (s.row[e] == q) dup = true;
        if (__any_sync(FULLMASK, dup)) {
            if (refactored) return NODE_FAIL;
            refactored = true;
            if (lp_refactor(s)) return NODE_FAIL;
            lp_gather(s); lp_point(s);
            if (s.z[0] < 0.0) lp_flip(s);
            continue;
        }
        const int sq = rq > 0.0 ? 1 : -1;
        lp_row(s, q, sq); lp_w(s);
        const int l = lp_ratio(s, sq);
        if (l < 0) return NODE_FAIL;
        lp_pivot(s, l);
        if (s.lane == 0) { s.row[l] = q; s.sg[l] = sq; }
        const double delta = (double)sq * (rmax - t);
        for (int v = s.lane; v < K; v += 32) s.z[v] = fma(s.g[v], delta, s.z[v]);
        __syncwarp();
    }
}

// child transform: last variable becomes fixed; dir = sgn(x_last - u).  K -> K-1.
__device__ int lp_child(WS &s, int dir) {
    const int K = s.K, LD = s.LD;
    for (int e = s.lane; e < K; e += 32) s.w[e] = s.W[(size_t)(K - 1) * LD + e];
    __syncwarp();
    const int l = lp_ratio(s, -dir);
    if (l < 0) return -1;
    lp_update(s, l, K - 1, false);
    if (l != K - 1) {
        for (int v = s.lane; v < K - 1; v += 32) s.W[(size_t)v * LD + l] = s.W[(size_t)v * LD + K - 1];
        if (s.lane == 0) { s.row[l] = s.row[K - 1]; s.sg[l] = s.sg[K - 1]; }
    }
    __syncwarp();
    s.K = K - 1;
    return 0;
}

// ------------------------------------------------------------------------------------------------------------
// Level store and records
// ------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ void chain_store(const WS &s, double *dH, int *dI) {  // W (K x K) -> packed
    const int K = s.K, LD = s.LD;
    for (int e = s.lane; e < K * K; e += 32) { const int v = e / K, c = e - v * K; dH[e] = s.W[(size_t)v * LD + c]; }
    for (int e = s.lane; e < K; e += 32) { dI[e] = s.row[e]; dI[K + e] = s.sg[e]; }
}
__device__ __forceinline__ void chain_load(WS &s, const double *dH, const int *dI, int K) {
    const int LD = s.LD;
    s.K = K;
    for (int e = s.lane; e < K * K; e += 32) { const int v = e / K, c = e - v * K; s.W[(size_t)v * LD + c] = dH[e]; }
    for (int e = s.lane; e < K; e += 32) { s.row[e] = dI[e]; s.sg[e] = dI[K + e]; }
    __syncwarp();
}

__device__ __forceinline__ unsigned long long gtimer() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
// watchdog: a search that overruns its deadline stops with BLT_STATUS_ERROR instead of hanging the device
__device__ __forceinline__ bool overdue(QCtl *q, int where) {
    const unsigned long long dl = *(volatile unsigned long long *)&q->deadline_ns;
    if (dl == 0 || gtimer() < dl) return false;
    if (atomicCAS(&q->timed_out, 0, where) == 0) { q->status = BLTCU_STATUS_ERROR; }
    q->abort = 1;
    __threadfence();
    return true;
}

// MPMC rings -------------------------------------------------------------------------------------------------
__device__ __forceinline__ int ring_pop(QCtl *q, volatile unsigned int *head, volatile unsigned int *tail, int *ring, unsigned int mask) {
    for (;;) {
        const unsigned int h = *head, t = *tail;
        if ((int)(t - h) <= 0) return -1;
        if (atomicCAS((unsigned int *)head, h, h + 1) == h) {
            volatile int *slot = ring + (h & mask);
            int v;
            unsigned int spins = 0;
            while ((v = *slot) < 0) { __nanosleep(20); if ((++spins & 1023u) == 0 && overdue(q, 3)) return -1; }
            *slot = -1;
            __threadfence();
            return v;
        }
    }
}
__device__ __forceinline__ void ring_push(QCtl *q, unsigned int *tail, int *ring, unsigned int mask, int val) {
    const unsigned int t = atomicAdd(tail, 1u);
    volatile int *slot = ring + (t & mask);
    unsigned int spins = 0;
    while (*slot >= 0) { __nanosleep(20); if ((++spins & 1023u) == 0 && overdue(q, 4)) return; }
    __threadfence();
    *slot = val;
}

// exact check, closest.cc:330-340 (int64 products wrap like the reference's)
__device__ bool check_soln_dev(const Params &P, const long long *u, int lane) {
    bool ok = true;
    for (int j = lane; j < P.m; j += 32) {
        unsigned long long z = 0;
        for (int i = 0; i < P.n; ++i) z += (unsigned long long)u[i] * (unsigned long long)P.lat[(size_t)i * P.m + j];
        const long long zz = (long long)z;
        if (P.con[2 * j] > zz || zz > P.con[2 * j + 1]) ok = false;
    }
    return __all_sync(FULLMASK, ok);
}

// ------------------------------------------------------------------------------------------------------------
// One subtree
// ------------------------------------------------------------------------------------------------------------
struct Counters { unsigned long long niter, nprune, nodes, donated; };

__device__ void run_task(const Params &P, WS &s, int rec, int gwarp, Counters &cnt, unsigned int *hist) {
    const int n = P.n, m = P.m, lane = s.lane;
    const RecLayout RL = rec_layout(n, m);
    unsigned char *R = P.pool + (size_t)rec * RL.stride;
    const int k0 = ((int *)R)[0];
    const int rflags = ((int *)R)[1];
    double xc0 = ((double *)R)[1];
    const bool is_root = (rflags & REC_ROOT) != 0;
    double *scrD = P.scrD + (size_t)gwarp * P.scrD_stride;
    int *scrI = P.scrI + (size_t)gwarp * P.scrI_stride;
    double *cparL = scrD + chain_off(n + 1);  // (n+2) x m after the chains
    auto chainH = [&](int k, int side) { return scrD + chain_off(k) + (size_t)side * k * k; };
    auto chainI = [&](int k, int side) { return scrI + chain_ioff(k) + (size_t)side * 2 * k; };

    // ---- load the record ----
    for (int e = lane; e < n; e += 32) { s.u[e] = ((long long *)(R + RL.off_u))[e]; s.ord[e] = (R + RL.off_ord)[e]; s.flags[e + 1] = 0; }
    for (int j = lane; j < m; j += 32) { const double v = ((double *)(R + RL.off_cpar))[j]; s.cpar[j] = v; cparL[(size_t)k0 * m + j] = v; }
    int wk = -1, wside = -1;  // which chain the working matrix W currently equals
    if (!is_root) {
        chain_load(s, (const double *)(R + RL.off_H), (const int *)(R + RL.off_row), k0 + 1);
        // row and sg are adjacent arrays of (n+1) ints in the record, but chain_load expects sg at dI[K+e]
        for (int e = lane; e < k0 + 1; e += 32) s.sg[e] = ((const int *)(R + RL.off_sg))[e];
        __syncwarp();
        if (k0 + 1 <= n) { chain_store(s, chainH(k0 + 1, 0), chainI(k0 + 1, 0)); wk = k0 + 1; wside = 0; }
    }
    __syncwarp();
    __threadfence();
    if (lane == 0) ring_push(P.q, &P.q->ftail, P.fring, P.ring_mask, rec);  // the record slot is free again

    double d = 0.0, xlast = 0.0;
    if (is_root) {  // initial solve LP(n), closest.cc:206
        for (int j = lane; j < m; j += 32) s.c[j] = s.cpar[j];
        __syncwarp();
        const int st = lp_cold(s, n + 1);
        cnt.nodes++; if (lane == 0) hist[n]++;
        int bad = 0;
        if (st < 0) bad = 1;
        else if (st == 1) xlast = s.z[n];
        else if (lp_solve(s, INFINITY, d, xlast) != NODE_EXPLORE) bad = 1;
        if (bad) { if (lane == 0) { P.q->status = BLTCU_STATUS_LP_FAIL; P.q->abort = 1; } return; }
        xc0 = xlast;
    }
    int k = k0;
    if (lane == 0) {
        const long long un = (long long)round(xc0);
        s.u[k - 1] = un; s.step[k - 1] = sgn_d(xc0 - (double)un); s.xc[k] = xc0; s.ord[k - 1] = 0; s.flags[k] = 0;
    }
    __syncwarp();
    const int depth0 = n - k0;
    const bool forced = depth0 < P.static_depth;  // static split: emit every explored child of this record

    unsigned int wd = 0;
    for (;;) {
        {   // one lane reads the stop flag so that the whole warp takes the same branch
            int ab = 0;
            if (lane == 0) {
                ab = *(volatile int *)&P.q->abort;
                if (!ab && (++wd & 255u) == 0) ab = overdue(P.q, 1) ? 1 : 0;
            }
            if (__shfl_sync(FULLMASK, ab, 0)) return;
        }
        // c = cpar - u[k-1] * B[k-1]
        {
            const double mu = -(double)s.u[k - 1];
            const double *Bk = s.Bs + (size_t)(k - 1) * s.LDB;
            for (int j = lane; j < m; j += 32) s.c[j] = fma(mu, Bk[j], s.cpar[j]);
            __syncwarp();
        }
        cnt.nodes++; if (lane == 0) hist[k - 1]++;
        int status, side = 0;
        const bool second_prune_possible = (s.flags[k] & 4) != 0;
        if (k == 1) {
            double dm = 0.0;
            for (int j = lane; j < m; j += 32) { const double a = fabs(s.c[j]); dm = a > dm ? a : dm; }
            d = warp_max(dm);
            status = approxLE(d, P.radius) ? NODE_EXPLORE : NODE_PRUNE;
        } else {
            const int dir = sgn_d(s.xc[k] - (double)s.u[k - 1]);
            side = dir > 0 ? 0 : 1;
            const bool inited = (s.flags[k] >> side) & 1;
            if (!inited) {
                if (is_root && k == n) {
                    if (lp_cold(s, k) != 0) { if (lane == 0) { P.q->status = BLTCU_STATUS_LP_FAIL; P.q->abort = 1; } return; }
                } else {
                    const int ps = (s.flags[k + 1] >> 3) & 1;
                    if (!(wk == k + 1 && wside == ps)) chain_load(s, chainH(k + 1, ps), chainI(k + 1, ps), k + 1);
                    if (lp_child(s, dir)) { if (lane == 0) { P.q->status = BLTCU_STATUS_LP_FAIL; P.q->abort = 1; } return; }
                }
                if (lane == 0) s.flags[k] |= (unsigned char)(1 << side);
                __syncwarp();
            } else if (!(wk == k && wside == side)) chain_load(s, chainH(k, side), chainI(k, side), k);
            status = lp_solve(s, P.radius, d, xlast);
            if (status == NODE_FAIL) { if (lane == 0) { P.q->status = BLTCU_STATUS_LP_FAIL; P.q->abort = 1; } return; }
            wk = k; wside = side;
            // keep the chain head unless this is the prune that leaves the level
            if (!(status == NODE_PRUNE && (second_prune_possible || (is_root && k == k0)))) {
                chain_store(s, chainH(k, side), chainI(k, side));
                __syncwarp();
            }
        }
        if (status == NODE_PRUNE) {
            if (is_root && k == k0) return;  // closest.cc:280-286: first prune at the top level ends the search
            cnt.nprune++;
            if (s.flags[k] & 4) {  // second prune at this level: leave it
                if (k == k0) return;
                k++;
                for (int j = lane; j < m; j += 32) s.cpar[j] = cparL[(size_t)k * m + j];
            } else if (lane == 0) s.flags[k] |= 4;
            __syncwarp();
            if (lane == 0) {
                const int st = s.step[k - 1];
                s.u[k - 1] += st; s.step[k - 1] = -st - (st > 0 ? 1 : -1);
                if (s.ord[k - 1] < 255) s.ord[k - 1]++;
            }
            __syncwarp();
            continue;
        }
        // explore
        if (k > 1) {
            // hand the child to the queue instead of descending?
            bool emit = false;
            int slot = -1;
            if (forced && k == k0) emit = true;
            else if (P.dynamic && (k - 1) >= P.donate_min_level && s.ord[k - 1] > 0) {  // keep the first (nearest) child
                int want = 0;
                if (lane == 0) {
                    const int idle = *(volatile int *)&P.q->idle;
                    const int qlen = (int)(*(volatile unsigned int *)&P.q->tail - *(volatile unsigned int *)&P.q->head);
                    want = (idle > 0 && qlen < idle + P.lowwater) ? 1 : 0;
                }
                emit = __shfl_sync(FULLMASK, want, 0) != 0;
            }
            if (emit) {
                if (lane == 0) slot = ring_pop(P.q, &P.q->fhead, &P.q->ftail, P.fring, P.ring_mask);
                slot = __shfl_sync(FULLMASK, slot, 0);
                if (slot < 0) {
                    if (forced && k == k0) { if (lane == 0) { P.q->status = BLTCU_STATUS_ERROR; P.q->abort = 1; } return; }
                    emit = false;
                }
            }
            if (emit) {
                unsigned char *C = P.pool + (size_t)slot * RL.stride;
                if (lane == 0) { ((int *)C)[0] = k - 1; ((int *)C)[1] = 0; ((double *)C)[1] = xlast; }
                for (int e = lane; e < n; e += 32) { ((long long *)(C + RL.off_u))[e] = s.u[e]; (C + RL.off_ord)[e] = s.ord[e]; }
                for (int j = lane; j < m; j += 32) ((double *)(C + RL.off_cpar))[j] = s.c[j];
                {
                    const int K = s.K, LD = s.LD;
                    double *dH = (double *)(C + RL.off_H);
                    for (int e = lane; e < K * K; e += 32) { const int v = e / K, c2 = e - v * K; dH[e] = s.W[(size_t)v * LD + c2]; }
                    for (int e = lane; e < K; e += 32) { ((int *)(C + RL.off_row))[e] = s.row[e]; ((int *)(C + RL.off_sg))[e] = s.sg[e]; }
                }
                __syncwarp();
                __threadfence();
                if (lane == 0) {
                    if (forced && (n - (k - 1)) >= P.static_depth) {
                        const unsigned int o = atomicAdd(&P.q->out_count, 1u);
                        P.out_list[o] = slot;
                    } else {
                        atomicAdd(&P.q->pending, 1);
                        ring_push(P.q, &P.q->tail, P.ring, P.ring_mask, slot);
                    }
                    // continue with the next sibling as if the subtree had returned
                    const int st = s.step[k - 1];
                    s.u[k - 1] += st; s.step[k - 1] = -st - (st > 0 ? 1 : -1);
                    if (s.ord[k - 1] < 255) s.ord[k - 1]++;
                }
                cnt.donated++;
                cnt.niter++;
                __syncwarp();
                continue;
            }
            // descend, closest.cc:88-95
            for (int j = lane; j < m; j += 32) { const double v = s.c[j]; cparL[(size_t)(k - 1) * m + j] = v; s.cpar[j] = v; }
            if (lane == 0) {
                s.flags[k] = (unsigned char)((s.flags[k] & ~8) | (side << 3));
                s.xc[k - 1] = xlast;
                const long long un = (long long)round(xlast);
                s.u[k - 2] = un; s.step[k - 2] = sgn_d(xlast - (double)un);
                s.flags[k - 1] = 0; s.ord[k - 2] = 0;
            }
            k--;
            __syncwarp();
        } else {
            // leaf: record, closest.cc:300-312
            {
                unsigned char *bk = P.best_key + (size_t)gwarp * n;
                int64_t *bu = P.best_u + (size_t)gwarp * n;
                bool newer = true;
                int bvld = 0;
                if (lane == 0) bvld = P.best_valid[gwarp];
                bvld = __shfl_sync(FULLMASK, bvld, 0);
                if (bvld) {  // lexicographic compare of ordinal paths from the top level down
                    int cmp = 0;
                    for (int e = n - 1; e >= 0 && cmp == 0; --e) cmp = (int)s.ord[e] - (int)bk[e];
                    newer = cmp >= 0;
                }
                if (newer) {
                    for (int e = lane; e < n; e += 32) { bk[e] = s.ord[e]; bu[e] = s.u[e]; }
                    if (lane == 0) P.best_valid[gwarp] = 1;
                }
                __syncwarp();
            }
            if (check_soln_dev(P, s.u, lane)) {
                int won = 0;
                if (lane == 0) won = (atomicCAS(&P.q->found, 0, 1) == 0);
                won = __shfl_sync(FULLMASK, won, 0);
                if (won) {
                    for (int e = lane; e < n; e += 32) P.sol[e] = s.u[e];
                    __threadfence();
                }
                if (!P.no_early_exit) { if (lane == 0) P.q->abort = 1; return; }
            }
            if (n == 1) return;  // D4 of oracle/blt_oracle.c
            // explore(), closest.cc:99-104
            k = 2;
            for (int j = lane; j < m; j += 32) s.cpar[j] = cparL[(size_t)2 * m + j];
            if (lane == 0) {
                const int st = s.step[1];
                s.u[1] += st; s.step[1] = -st - (st > 0 ? 1 : -1);
                if (s.ord[1] < 255) s.ord[1]++;
            }
            __syncwarp();
        }
        cnt.niter++;
    }
}

// ------------------------------------------------------------------------------------------------------------
// The persistent kernel
// ------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256, 1) bltcu_search_kernel(const Params P) {
    extern __shared__ __align__(16) double smem[];
    const int n = P.n, m = P.m;
    double *Bs = smem;
    for (int e = threadIdx.x; e < n * P.LDB; e += blockDim.x) {
        const int i = e / P.LDB, j = e - i * P.LDB;
        Bs[e] = j < m ? P.B[(size_t)i * m + j] : 0.0;
    }
    __syncthreads();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int gwarp = blockIdx.x * P.warps_per_cta + warp;
    double *base = smem + (size_t)n * P.LDB + (size_t)warp * P.warp_smem_doubles;
    WS s;
    const int KM = P.KMAX;
    s.W = base; base += (size_t)KM * P.LD;
    s.z = base; base += KM; s.g = base; base += KM; s.w = base; base += KM; s.a = base; base += KM; s.cS = base; base += KM;
    s.c = base; base += m; s.cpar = base; base += m; s.xc = base; base += n + 2;
    s.u = (long long *)base; base += n;
    int *ib = (int *)base;
    s.row = ib; ib += KM; s.sg = ib; ib += KM; s.tmpi = ib; ib += KM; s.tmps = ib; ib += KM; s.step = ib; ib += n + 1;
    unsigned int *hist = (unsigned int *)ib; ib += n + 1;
    s.ord = (unsigned char *)ib; s.flags = s.ord + ((n + 8) / 8 * 8);
    s.Bs = Bs; s.LD = P.LD; s.LDB = P.LDB; s.n = n; s.m = m; s.K = 0; s.lane = lane;
    s.pivots = s.quick = s.refactors = s.colds = 0;
    for (int e = lane; e <= n; e += 32) hist[e] = 0;
    __syncwarp();
    Counters cnt = {0, 0, 0, 0};
    QCtl *q = P.q;
    bool am_idle = false;
    for (;;) {
        int rec = -1, stop = 0;
        if (lane == 0) {
            if (*(volatile int *)&q->abort) stop = 1;
            else {
                rec = ring_pop(q, &q->head, &q->tail, P.ring, P.ring_mask);
                if (rec < 0) {
                    if (*(volatile int *)&q->pending == 0) stop = 1;
                    else {
                        if (!am_idle) { atomicAdd(&q->idle, 1); am_idle = true; }
                        __nanosleep(200);
                        if (overdue(q, 2)) stop = 1;
                    }
                } else if (am_idle) { atomicSub(&q->idle, 1); am_idle = false; }
            }
        }
        stop = __shfl_sync(FULLMASK, stop, 0);
        rec = __shfl_sync(FULLMASK, rec, 0);
        if (stop) break;
        if (rec < 0) continue;
        run_task(P, s, rec, gwarp, cnt, hist);
        __syncwarp();
        if (lane == 0) { __threadfence(); atomicSub(&q->pending, 1); }
    }
    if (lane == 0) {
        if (am_idle) atomicSub(&q->idle, 1);
        atomicAdd(&q->niter, cnt.niter); atomicAdd(&q->nprune, cnt.nprune); atomicAdd(&q->nodes, cnt.nodes);
        atomicAdd(&q->donated, cnt.donated);
        atomicAdd(&q->pivots, s.pivots); atomicAdd(&q->quick, s.quick); atomicAdd(&q->refactors, s.refactors);
        atomicAdd(&q->colds, s.colds);
    }
    __syncwarp();
    for (int e = lane; e <= n; e += 32) if (hist[e]) atomicAdd(&q->hist[e], (unsigned long long)hist[e]);
}

// small helper kernels ---------------------------------------------------------------------------------------
__global__ void bltcu_setup_kernel(const int64_t *lat, const int64_t *con, int n, int m, double *B, double *p,
                                   unsigned long long *now) {
    const int tid = blockIdx.x * blockDim.x + threadIdx.x;
    if (tid == 0) *now = gtimer();
    if (tid < n * m) B[tid] = (double)lat[tid];                                   // closest.cc:203
    if (tid < m) p[tid] = (double)(con[2 * tid + 1] + con[2 * tid]) / 2.0;        // closest.cc:180-185
}
__global__ void bltcu_check_kernel(const int64_t *lat, const int64_t *con, int n, int m, const int64_t *u, int *ok) {
    Params P; P.n = n; P.m = m; P.lat = lat; P.con = con;
    const bool r = check_soln_dev(P, (const long long *)u, threadIdx.x & 31);
    if (threadIdx.x == 0) *ok = r ? 1 : 0;
}
// one cold layer-distance solve (test seam): c computed as glpk_wrapper.cc:156-166 does
__global__ void bltcu_infdist_kernel(const double *B, int n, int m, int LD, int LDB, const int64_t *u, int k,
                                     const double *p, double *x, double *dist, int *rc) {
    extern __shared__ __align__(16) double smem[];
    double *Bs = smem;
    const int lane = threadIdx.x;
    for (int e = lane; e < n * LDB; e += 32) { const int i = e / LDB, j = e - i * LDB; Bs[e] = j < m ? B[(size_t)i * m + j] : 0.0; }
    __syncwarp();
    const int KM = n + 1;
    double *base = smem + (size_t)n * LDB;
    WS s;
    s.W = base; base += (size_t)KM * LD;
    s.z = base; base += KM; s.g = base; base += KM; s.w = base; base += KM; s.a = base; base += KM; s.cS = base; base += KM;
    s.c = base; base += m; s.cpar = base; base += m; s.xc = base; base += n + 2;
    s.u = (long long *)base; base += n;
    int *ib = (int *)base;
    s.row = ib; ib += KM; s.sg = ib; ib += KM; s.tmpi = ib; ib += KM; s.tmps = ib; ib += KM; s.step = ib;
    s.Bs = Bs; s.LD = LD; s.LDB = LDB; s.n = n; s.m = m; s.K = 0; s.lane = lane;
    s.pivots = s.quick = s.refactors = s.colds = 0;
    for (int i = lane; i < m; i += 32) {
        double rhs1 = 0.0;
        for (int j = k; j < n; ++j) rhs1 += (double)(-u[j]) * B[(size_t)j * m + i];
        rhs1 += p[i];
        s.c[i] = rhs1;
    }
    __syncwarp();
    int r = 0;
    double d = 0.0, xl = 0.0;
    if (k == 0) {
        double dm = 0.0;
        for (int j = lane; j < m; j += 32) { const double a = fabs(s.c[j]); dm = a > dm ? a : dm; }
        d = warp_max(dm);
    } else {
        const int st = lp_cold(s, k + 1);
        if (st < 0) r = -1;
        else if (st == 1) {
            double dm = 0.0;
            for (int j = lane; j < m; j += 32) { const double a = fabs(residual(s, j)); dm = a > dm ? a : dm; }
            d = warp_max(dm);
        } else if (lp_solve(s, INFINITY, d, xl) != NODE_EXPLORE) r = -1;
        if (r == 0) for (int i = lane; i < k; i += 32) x[i] = s.z[i + 1];
    }
    if (lane == 0) { *dist = d; *rc = r; }
}

// ------------------------------------------------------------------------------------------------------------
// Host side
// ------------------------------------------------------------------------------------------------------------
#define CK(call)                                                                                       \
    do {                                                                                               \
        cudaError_t e_ = (call);                                                                       \
        if (e_ != cudaSuccess) {                                                                       \
            fprintf(stderr, "bltcu: CUDA error %s at %s:%d: %s\n", cudaGetErrorName(e_), __FILE__, __LINE__, \
                    cudaGetErrorString(e_));                                                           \
            return BLTCU_STATUS_ERROR;                                                                 \
        }                                                                                              \
    } while (0)

namespace {
struct DevWs {  // cached per device, grow-only
    int device = -1;
    int sms = 0;
    size_t smem_optin = 0;
    void *buf = nullptr;
    size_t cap = 0;
    int64_t *d_in = nullptr;  // staging for host lat/con
    size_t in_cap = 0;
    cudaStream_t stream = nullptr;
    cudaEvent_t e0 = nullptr, e1 = nullptr;
};
std::mutex g_mu;
std::vector<DevWs *> g_ws;

int get_ws(int device, DevWs **out) {
    for (auto *w : g_ws) if (w->device == device) { CK(cudaSetDevice(device)); *out = w; return 0; }
    CK(cudaSetDevice(device));
    DevWs *w = new DevWs();
    w->device = device;
    int v = 0;
    CK(cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, device)); w->sms = v;
    CK(cudaDeviceGetAttribute(&v, cudaDevAttrMaxSharedMemoryPerBlockOptin, device)); w->smem_optin = (size_t)v;
    CK(cudaStreamCreateWithFlags(&w->stream, cudaStreamNonBlocking));
    CK(cudaEventCreate(&w->e0)); CK(cudaEventCreate(&w->e1));
    g_ws.push_back(w);
    *out = w;
    return 0;
}
int ws_reserve(DevWs *w, size_t bytes) {
    if (bytes <= w->cap) return 0;
    if (w->buf) { CK(cudaFree(w->buf)); w->buf = nullptr; w->cap = 0; }
    CK(cudaMalloc(&w->buf, bytes));
    w->cap = bytes;
    return 0;
}
size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }
}  // namespace

static int search_impl(DevWs *w, const int64_t *d_lat, const int64_t *d_con, int n, int m, const bltcu_opts *opts,
                       int64_t *sol, bltcu_stats *stats) {
    const auto t_start = std::chrono::steady_clock::now();
    bltcu_opts o;
    memset(&o, 0, sizeof o);
    if (opts) o = *opts;
    const int world = o.world > 1 ? o.world : 1;
    const int rank = world > 1 ? o.rank : 0;
    if (world > 1 && o.split_depth <= 0) o.split_depth = std::min(6, std::max(1, n - 3));
    if (o.split_depth > n - 2) o.split_depth = std::max(0, n - 2);

    // ---- launch shape from the shared-memory budget ----
    const int KMAX = n + 1;
    const int LD = KMAX | 1, LDB = m | 1;
    size_t wd = (size_t)KMAX * LD + 5 * (size_t)KMAX + 2 * (size_t)m + (n + 2) + n;           // doubles
    size_t wi = 4 * (size_t)KMAX + (n + 1) + (n + 1);                                         // ints
    size_t wb = 2 * (size_t)((n + 8) / 8 * 8) + 16;                                            // bytes
    size_t warp_bytes = align_up(wd * 8 + wi * 4 + wb, 16);
    const size_t b_bytes = (size_t)n * LDB * 8;
    const size_t budget = w->smem_optin;
    int wpc = (int)((budget > b_bytes ? budget - b_bytes : 0) / warp_bytes);
    if (wpc > 8) wpc = 8;
    if (o.warps_per_cta > 0 && o.warps_per_cta < wpc) wpc = o.warps_per_cta;
    if (wpc < 1) {
        fprintf(stderr, "bltcu: problem %d x %d does not fit the shared-memory budget (%zu bytes)\n", n, m, budget);
        return BLTCU_STATUS_ERROR;
    }
    int ctas = o.ctas > 0 ? o.ctas : w->sms;
    const int nwarps = ctas * wpc;
    const size_t smem_bytes = b_bytes + (size_t)wpc * warp_bytes;

    // ---- carve the global workspace ----
    const RecLayout RL = rec_layout(n, m);
    int pool_cap = 4096;
    while ((size_t)pool_cap * RL.stride > ((size_t)3 << 30) && pool_cap > 256) pool_cap >>= 1;
    const unsigned int ring_cap = (unsigned int)pool_cap;  // power of two
    const size_t scrD_stride = align_up(chain_off(n + 1) + (size_t)(n + 2) * m, 32);
    const size_t scrI_stride = align_up(chain_ioff(n + 1) + 8, 32);
    size_t off = 0;
    auto carve = [&](size_t bytes) { size_t r = off; off = align_up(off + bytes, 256); return r; };
    const size_t o_q = carve(sizeof(QCtl));
    const size_t o_B = carve((size_t)n * m * 8);
    const size_t o_p = carve((size_t)m * 8);
    const size_t o_ring = carve((size_t)ring_cap * 4);
    const size_t o_fring = carve((size_t)ring_cap * 4);
    const size_t o_out = carve((size_t)pool_cap * 4);
    const size_t o_sol = carve((size_t)n * 8);
    const size_t o_bu = carve((size_t)nwarps * n * 8);
    const size_t o_bk = carve((size_t)nwarps * n);
    const size_t o_bv = carve((size_t)nwarps * 4);
    const size_t o_pool = carve((size_t)pool_cap * RL.stride);
    const size_t o_scrD = carve((size_t)nwarps * scrD_stride * 8);
    const size_t o_scrI = carve((size_t)nwarps * scrI_stride * 4);
    if (ws_reserve(w, off)) return BLTCU_STATUS_ERROR;
    unsigned char *base = (unsigned char *)w->buf;
    cudaStream_t st = w->stream;

    Params P;
    memset(&P, 0, sizeof P);
    P.n = n; P.m = m; P.LD = LD; P.LDB = LDB; P.KMAX = KMAX;
    P.B = (double *)(base + o_B); P.p = (double *)(base + o_p);
    P.lat = d_lat; P.con = d_con;
    P.q = (QCtl *)(base + o_q);
    P.ring = (int *)(base + o_ring); P.fring = (int *)(base + o_fring); P.ring_mask = ring_cap - 1;
    P.pool = base + o_pool; P.rec_stride = RL.stride; P.pool_cap = pool_cap;
    P.scrD = (double *)(base + o_scrD); P.scrD_stride = scrD_stride;
    P.scrI = (int *)(base + o_scrI); P.scrI_stride = scrI_stride;
    P.sol = (int64_t *)(base + o_sol);
    P.best_u = (int64_t *)(base + o_bu); P.best_key = base + o_bk; P.best_valid = (int *)(base + o_bv);
    P.out_list = (int *)(base + o_out);
    P.donate_min_level = o.donate_min_level > 0 ? std::max(2, o.donate_min_level) : std::min(6, std::max(2, n / 3));
    P.lowwater = std::min(nwarps, pool_cap / 4);
    P.no_early_exit = o.no_early_exit;
    P.warps_per_cta = wpc;
    P.warp_smem_doubles = warp_bytes / 8;

    CK(cudaFuncSetAttribute(bltcu_search_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes));
    int launches = 0;
    bltcu_setup_kernel<<<(n * m + 255) / 256, 256, 0, st>>>(d_lat, d_con, n, m, (double *)P.B, (double *)P.p,
                                                          (unsigned long long *)P.sol);
    CK(cudaGetLastError());
    launches++;
    // radius: closest.cc:193-194 uses row 0 only; device clock for the watchdog deadline
    int64_t con0[2];
    unsigned long long now_ns = 0, deadline_ns = 0;
    CK(cudaMemcpyAsync(con0, d_con, 16, cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(&now_ns, P.sol, 8, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    P.radius = (double)(con0[1] - con0[0]) / 2.0;
    {
        long long tmo = o.timeout_ms;
        if (tmo <= 0) { const char *e = getenv("BLTCU_TIMEOUT_MS"); if (e) tmo = atoll(e); }
        if (tmo > 0) deadline_ns = now_ns + (unsigned long long)tmo * 1000000ull;
    }

    // host images of the control block and rings
    std::vector<int> ring(ring_cap, -1), fring(ring_cap);
    QCtl hq;
    auto reset_ctl = [&](const std::vector<int> &initial, int first_free) -> int {
        memset(&hq, 0, sizeof hq);
        std::fill(ring.begin(), ring.end(), -1);
        for (size_t i = 0; i < initial.size(); ++i) ring[i] = initial[i];
        hq.tail = (unsigned int)initial.size();
        hq.pending = (int)initial.size();
        // free ring: slots not in use
        std::fill(fring.begin(), fring.end(), -1);
        unsigned int nf = 0;
        std::vector<char> used(pool_cap, 0);
        for (int r : initial) used[r] = 1;
        for (int i = 0; i < pool_cap; ++i) if (!used[i] && i >= first_free) fring[nf++] = i;
        for (int i = 0; i < std::min(first_free, pool_cap); ++i) if (!used[i]) fring[nf++] = i;
        hq.ftail = nf;
        hq.deadline_ns = deadline_ns;
        CK(cudaMemcpyAsync(P.q, &hq, sizeof hq, cudaMemcpyHostToDevice, st));
        CK(cudaMemcpyAsync(P.ring, ring.data(), ring_cap * 4, cudaMemcpyHostToDevice, st));
        CK(cudaMemcpyAsync(P.fring, fring.data(), ring_cap * 4, cudaMemcpyHostToDevice, st));
        return 0;
    };
    CK(cudaMemsetAsync(P.best_valid, 0, (size_t)nwarps * 4, st));
    CK(cudaMemsetAsync(P.sol, 0, (size_t)n * 8, st));

    // root record in slot 0: k = n, flags ROOT, cpar = p
    {
        std::vector<unsigned char> rec(RL.stride, 0);
        ((int *)rec.data())[0] = n;
        ((int *)rec.data())[1] = REC_ROOT;
        CK(cudaMemcpyAsync(P.pool, rec.data(), RL.stride, cudaMemcpyHostToDevice, st));
        CK(cudaStreamSynchronize(st));
        // cpar = p lives on the device: copy it there
        CK(cudaMemcpyAsync(P.pool + RL.off_cpar, P.p, (size_t)m * 8, cudaMemcpyDeviceToDevice, st));
    }

    float kernel_ms = 0.f;
    QCtl acc;  // totals
    memset(&acc, 0, sizeof acc);
    auto add_counts = [&](const QCtl &x) {
        acc.niter += x.niter; acc.nprune += x.nprune; acc.nodes += x.nodes; acc.pivots += x.pivots; acc.quick += x.quick;
        acc.refactors += x.refactors; acc.colds += x.colds; acc.donated += x.donated;
        for (int i = 0; i <= n; ++i) acc.hist[i] += x.hist[i];
    };
    auto launch = [&]() -> int {
        CK(cudaEventRecord(w->e0, st));
        bltcu_search_kernel<<<ctas, wpc * 32, smem_bytes, st>>>(P);
        CK(cudaGetLastError());
        CK(cudaEventRecord(w->e1, st));
        CK(cudaMemcpyAsync(&hq, P.q, sizeof hq, cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));
        float ms = 0.f;
        CK(cudaEventElapsedTime(&ms, w->e0, w->e1));
        kernel_ms += ms;
        launches++;
        if (hq.timed_out)
            fprintf(stderr, "bltcu: watchdog fired (where=%d) pending=%d idle=%d head=%u tail=%u fhead=%u ftail=%u found=%d\n",
                    hq.timed_out, hq.pending, hq.idle, hq.head, hq.tail, hq.fhead, hq.ftail, hq.found);
        return 0;
    };

    int status = BLTCU_STATUS_OK, found = 0;
    std::vector<int> initial;
    if (o.split_depth > 0) {
        // phase A: deterministic static split of the top levels (every rank computes the same list)
        initial.assign(1, 0);
        if (reset_ctl(initial, 1)) return BLTCU_STATUS_ERROR;
        P.static_depth = o.split_depth; P.dynamic = 0;
        if (launch()) return BLTCU_STATUS_ERROR;
        status = hq.status; found = hq.found;
        if (rank == 0) add_counts(hq);
        std::vector<int> out(hq.out_count);
        if (hq.out_count) CK(cudaMemcpy(out.data(), P.out_list, (size_t)hq.out_count * 4, cudaMemcpyDeviceToHost));
        // order the collected subtrees by their path (DFS order) so that every rank deals the same hands
        std::vector<std::vector<unsigned char>> keys(out.size(), std::vector<unsigned char>(n));
        for (size_t i = 0; i < out.size(); ++i) {
            CK(cudaMemcpy(keys[i].data(), P.pool + (size_t)out[i] * RL.stride + RL.off_ord, n, cudaMemcpyDeviceToHost));
            std::reverse(keys[i].begin(), keys[i].end());  // top level first
        }
        std::vector<int> idx(out.size());
        for (size_t i = 0; i < idx.size(); ++i) idx[i] = (int)i;
        std::sort(idx.begin(), idx.end(), [&](int a, int b) { return keys[a] < keys[b]; });
        initial.clear();
        for (size_t i = 0; i < idx.size(); ++i) if ((int)(i % world) == rank) initial.push_back(out[idx[i]]);
        const bool stop = (status != BLTCU_STATUS_OK) || (found && !o.no_early_exit && (hq.abort != 0));
        if (!stop && !initial.empty()) {
            // phase B keeps every record of phase A alive only if it is in this rank's hand; the rest is free
            if (reset_ctl(initial, 0)) return BLTCU_STATUS_ERROR;
            if (found) { hq.found = 1; CK(cudaMemcpyAsync(P.q, &hq, sizeof hq, cudaMemcpyHostToDevice, st)); }
            P.static_depth = 0; P.dynamic = 1;
            if (launch()) return BLTCU_STATUS_ERROR;
            if (hq.status != BLTCU_STATUS_OK) status = hq.status;
            found = found || hq.found;
            add_counts(hq);
        }
    } else {
        initial.assign(1, 0);
        if (reset_ctl(initial, 1)) return BLTCU_STATUS_ERROR;
        P.static_depth = 0; P.dynamic = 1;
        if (launch()) return BLTCU_STATUS_ERROR;
        status = hq.status; found = hq.found;
        add_counts(hq);
    }

    // ---- result: the point found, else the last recorded candidate in DFS order (uhat) ----
    std::vector<int64_t> hsol(n, 0);
    if (found) {
        CK(cudaMemcpy(hsol.data(), P.sol, (size_t)n * 8, cudaMemcpyDeviceToHost));
    } else {
        std::vector<int> bv(nwarps);
        CK(cudaMemcpy(bv.data(), P.best_valid, (size_t)nwarps * 4, cudaMemcpyDeviceToHost));
        int bestw = -1;
        std::vector<unsigned char> bkey(n), key(n);
        for (int i = 0; i < nwarps; ++i) {
            if (!bv[i]) continue;
            CK(cudaMemcpy(key.data(), P.best_key + (size_t)i * n, n, cudaMemcpyDeviceToHost));
            std::reverse(key.begin(), key.end());
            if (bestw < 0 || key >= bkey) { bestw = i; bkey = key; }
        }
        if (bestw >= 0) CK(cudaMemcpy(hsol.data(), P.best_u + (size_t)bestw * n, (size_t)n * 8, cudaMemcpyDeviceToHost));
    }
    if (sol) memcpy(sol, hsol.data(), (size_t)n * 8);
    if (stats) {
        memset(stats, 0, sizeof *stats);
        stats->niter = acc.niter; stats->nprune = acc.nprune; stats->nodes = acc.nodes; stats->pivots = acc.pivots;
        stats->quick_prunes = acc.quick; stats->refactors = acc.refactors; stats->cold_starts = acc.colds;
        stats->donated = acc.donated;
        for (int i = 0; i <= n; ++i) stats->hist[i] = acc.hist[i];
        stats->kernel_ms = kernel_ms;
        stats->found = found;
        stats->launches = launches;
        stats->warps = nwarps; stats->ctas = ctas;
        stats->scratch_bytes = off;
        stats->total_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_start).count();
    }
    return status;
}

static int validate(int n, int m) {
    if (n <= 0 || m <= 0 || m < n) return BLTCU_STATUS_INPUT_ERROR;  // closest.cc:173-177
    if (n > MAXD || m > MAXD) {
        fprintf(stderr, "bltcu: dimension %d x %d exceeds BLTCU_MAX_DIM=%d\n", n, m, MAXD);
        return BLTCU_STATUS_ERROR;
    }
    return BLTCU_STATUS_OK;
}

extern "C" int bltcu_closest_dev(const int64_t *d_lat, const int64_t *d_con, int n, int m, const bltcu_opts *opts,
                                 int64_t *sol, bltcu_stats *stats) {
    const int v = validate(n, m);
    if (v != BLTCU_STATUS_OK) return v;
    std::lock_guard<std::mutex> lk(g_mu);
    DevWs *w = nullptr;
    if (get_ws(opts ? opts->device : 0, &w)) return BLTCU_STATUS_ERROR;
    return search_impl(w, d_lat, d_con, n, m, opts, sol, stats);
}

extern "C" int bltcu_closest_ex(const int64_t *lat, const int64_t *con, int n, int m, const bltcu_opts *opts,
                                int64_t *sol, bltcu_stats *stats) {
    const int v = validate(n, m);
    if (v != BLTCU_STATUS_OK) return v;
    if (!lat || !con) return BLTCU_STATUS_INPUT_ERROR;
    const auto t0 = std::chrono::steady_clock::now();
    std::lock_guard<std::mutex> lk(g_mu);
    DevWs *w = nullptr;
    if (get_ws(opts ? opts->device : 0, &w)) return BLTCU_STATUS_ERROR;
    const size_t need = ((size_t)n * m + 2 * (size_t)m) * 8;
    if (need > w->in_cap) {
        if (w->d_in) CK(cudaFree(w->d_in));
        w->d_in = nullptr; w->in_cap = 0;
        CK(cudaMalloc((void **)&w->d_in, need));
        w->in_cap = need;
    }
    CK(cudaMemcpyAsync(w->d_in, lat, (size_t)n * m * 8, cudaMemcpyHostToDevice, w->stream));
    CK(cudaMemcpyAsync(w->d_in + (size_t)n * m, con, (size_t)2 * m * 8, cudaMemcpyHostToDevice, w->stream));
    const int rc = search_impl(w, w->d_in, w->d_in + (size_t)n * m, n, m, opts, sol, stats);
    if (stats) stats->total_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    return rc;
}

extern "C" int bltcu_closest(const int64_t *lat, const int64_t *con, int n, int m, int64_t *sol, uint64_t *niter,
                             uint64_t *nprune) {
    bltcu_stats st;
    const int rc = bltcu_closest_ex(lat, con, n, m, nullptr, sol, &st);
    if (rc == BLTCU_STATUS_OK || rc == BLTCU_STATUS_LP_FAIL) {
        if (niter) *niter = st.niter;
        if (nprune) *nprune = st.nprune;
    }
    return rc;
}

extern "C" int bltcu_check_soln(const int64_t *lat, const int64_t *con, int n, int m, const int64_t *u, int *ok) {
    if (n <= 0 || m <= 0 || !lat || !con || !u || !ok) return BLTCU_STATUS_INPUT_ERROR;
    std::lock_guard<std::mutex> lk(g_mu);
    DevWs *w = nullptr;
    if (get_ws(0, &w)) return BLTCU_STATUS_ERROR;
    const size_t bytes = ((size_t)n * m + 2 * (size_t)m + n + 1) * 8;
    int64_t *d = nullptr;
    CK(cudaMalloc((void **)&d, bytes));
    CK(cudaMemcpy(d, lat, (size_t)n * m * 8, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(d + (size_t)n * m, con, (size_t)2 * m * 8, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(d + (size_t)n * m + 2 * m, u, (size_t)n * 8, cudaMemcpyHostToDevice));
    int *dok = (int *)(d + (size_t)n * m + 2 * m + n);
    bltcu_check_kernel<<<1, 32, 0, w->stream>>>(d, d + (size_t)n * m, n, m, d + (size_t)n * m + 2 * m, dok);
    CK(cudaGetLastError());
    CK(cudaStreamSynchronize(w->stream));
    CK(cudaMemcpy(ok, dok, 4, cudaMemcpyDeviceToHost));
    CK(cudaFree(d));
    return BLTCU_STATUS_OK;
}

extern "C" int bltcu_infdist(const double *B, int n, int m, const int64_t *u, int k, const double *p, double *x,
                             double *dist) {
    if (n <= 0 || m <= 0 || k < 0 || k > n || k > m || n > MAXD || m > MAXD) return BLTCU_STATUS_INPUT_ERROR;
    std::lock_guard<std::mutex> lk(g_mu);
    DevWs *w = nullptr;
    if (get_ws(0, &w)) return BLTCU_STATUS_ERROR;
    const int KMAX = n + 1, LD = KMAX | 1, LDB = m | 1;
    const size_t smem = ((size_t)n * LDB + (size_t)KMAX * LD + 5 * (size_t)KMAX + 2 * (size_t)m + (n + 2) + n) * 8 +
                        (5 * (size_t)KMAX + 8) * 4;
    if (smem > w->smem_optin) return BLTCU_STATUS_ERROR;
    const size_t nb = (size_t)n * m * 8, nu = (size_t)n * 8, np = (size_t)m * 8;
    unsigned char *d = nullptr;
    CK(cudaMalloc((void **)&d, nb + nu + np + nu + 16));
    double *dB = (double *)d; int64_t *du = (int64_t *)(d + nb); double *dp = (double *)(d + nb + nu);
    double *dx = (double *)(d + nb + nu + np); double *dd = (double *)(d + nb + nu + np + nu); int *drc = (int *)(dd + 1);
    CK(cudaMemcpy(dB, B, nb, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(du, u, nu, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(dp, p, np, cudaMemcpyHostToDevice));
    CK(cudaFuncSetAttribute(bltcu_infdist_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    bltcu_infdist_kernel<<<1, 32, smem, w->stream>>>(dB, n, m, LD, LDB, du, k, dp, dx, dd, drc);
    CK(cudaGetLastError());
    CK(cudaStreamSynchronize(w->stream));
    int rc = 0;
    CK(cudaMemcpy(&rc, drc, 4, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(dist, dd, 8, cudaMemcpyDeviceToHost));
    if (k > 0 && rc == 0) CK(cudaMemcpy(x, dx, (size_t)k * 8, cudaMemcpyDeviceToHost));
    CK(cudaFree(d));
    return rc == 0 ? BLTCU_STATUS_OK : BLTCU_STATUS_LP_FAIL;
}

extern "C" void bltcu_release(void) {
    std::lock_guard<std::mutex> lk(g_mu);
    for (auto *w : g_ws) {
        cudaSetDevice(w->device);
        if (w->buf) cudaFree(w->buf);
        if (w->d_in) cudaFree(w->d_in);
        if (w->e0) cudaEventDestroy(w->e0);
        if (w->e1) cudaEventDestroy(w->e1);
        if (w->stream) cudaStreamDestroy(w->stream);
        delete w;
    }
    g_ws.clear();
}

extern "C" const char *bltcu_version(void) { return "blt-b200 0.1 (sm_100a)"; }

// ------------------------------------------------------------------------------------------------------------
// Allocation-free variant of the search
// ------------------------------------------------------------------------------------------------------------
// Who needs this: code that runs the search from a place where device memory may not be requested or returned --
// a plugin hosted inside another application whose memory manager owns the GPU heap, a solver callback in a
// real-time loop where a cudaMalloc (which may synchronise the whole device) is not acceptable, or a process that
// has given all free HBM to an arena and would get cudaErrorMemoryAllocation from any further request.  The normal
// entry points grow a cached workspace with cudaMalloc; this one never does.  All global scratch of the search
// kernel (control block, subtree queue and pool, per-warp level store, staging for the inputs) is carved out of
// the fixed array below, which is part of the library image: every GPU that loads libbltcu.so holds these 20 MiB
// from the moment the library is loaded until the process exits, whether or not this entry point is ever called.
#define BLTCU_NOALLOC_BYTES (20u << 20)
#define BLTCU_NOALLOC_WPC 2       // warps per CTA of this variant (fixed so that the level store has a fixed size)
#define BLTCU_NOALLOC_CTAS 148    // at most this many CTAs
#define BLTCU_NOALLOC_POOL 512    // subtree records (power of two)
#define BLTCU_NOALLOC_MAXDEV 64

static __device__ __align__(256) unsigned char g_noalloc_scratch[BLTCU_NOALLOC_BYTES];

// In host code `g_noalloc_scratch` is only a handle for the runtime's symbol table, not an address a kernel can
// dereference; the device address is looked up once per device and cached here.
static unsigned char *g_noalloc_dev[BLTCU_NOALLOC_MAXDEV];
static int g_noalloc_sms[BLTCU_NOALLOC_MAXDEV];
static size_t g_noalloc_smem[BLTCU_NOALLOC_MAXDEV];

// Puts the shared array into the state a fresh search expects.  The array still holds whatever the previous call
// left behind, so everything the kernel reads before writing is rewritten here: the control block, both rings, the
// root record, the result slot and the per-warp "recorded a candidate" flags.
__global__ void bltcu_noalloc_init_kernel(Params P, unsigned long long deadline_ns, int nwarps) {
    const int tid = blockIdx.x * blockDim.x + threadIdx.x, nth = gridDim.x * blockDim.x;
    const RecLayout RL = rec_layout(P.n, P.m);
    unsigned char *q = (unsigned char *)P.q;
    for (size_t i = tid; i < sizeof(QCtl); i += nth) q[i] = 0;
    for (int i = tid; i <= (int)P.ring_mask; i += nth) {
        P.ring[i] = (i == 0) ? 0 : -1;                      // the root record sits in slot 0
        P.fring[i] = (i + 1 <= (int)P.ring_mask) ? i + 1 : -1;  // slots 1..cap-1 are free
    }
    for (size_t i = tid; i < RL.stride; i += nth) P.pool[i] = 0;
    for (int i = tid; i < P.n; i += nth) P.sol[i] = 0;
    for (int i = tid; i < nwarps; i += nth) P.best_valid[i] = 0;
    __threadfence();
    __syncthreads();
    if (blockIdx.x == 0) {
        if (threadIdx.x == 0) {
            P.q->tail = 1; P.q->pending = 1; P.q->ftail = P.ring_mask; P.q->deadline_ns = deadline_ns;
            ((int *)P.pool)[0] = P.n;
            ((int *)P.pool)[1] = REC_ROOT;
        }
        for (int j = threadIdx.x; j < P.m; j += blockDim.x) ((double *)(P.pool + RL.off_cpar))[j] = P.p[j];
    }
}

// bltcu_closest_noalloc: closest() (libblt/closest.h:50-54) without any device or host allocation.
//
//   lat, con, sol, niter, nprune, found   host pointers, same meaning as bltcu_closest(); n, m <= BLTCU_NOALLOC_MAX_DIM
//
// * The first call on a device looks up the device address of the scratch array (cudaGetSymbolAddress) and the
//   device's SM count; make that first call before starting any stream capture.  Later calls reuse the cached values.
// * Work is issued on the legacy default stream of `device`; the call blocks until the result is on the host.
// * All calls share the one scratch array.  Two calls whose kernels overlap on the same GPU (two host threads, or
//   two processes' worth of streams through this library in one process) corrupt each other's search: serialise them.
// * Returns BLT_STATUS_INPUT_ERROR with a message on stderr when n or m exceed the limit, otherwise the same
//   status codes as bltcu_closest().
extern "C" int bltcu_closest_noalloc(const int64_t *lat, const int64_t *con, int n, int m, int device, int64_t *sol,
                                     uint64_t *niter, uint64_t *nprune, int *found) {
    if (n <= 0 || m <= 0 || m < n || !lat || !con) return BLTCU_STATUS_INPUT_ERROR;  // closest.cc:173-177
    if (n > BLTCU_NOALLOC_MAX_DIM || m > BLTCU_NOALLOC_MAX_DIM) {
        fprintf(stderr, "bltcu_closest_noalloc: %d x %d exceeds the allocation-free limit of %d x %d "
                        "(the scratch array is fixed at %u MiB); use bltcu_closest() for larger problems\n",
                n, m, BLTCU_NOALLOC_MAX_DIM, BLTCU_NOALLOC_MAX_DIM, BLTCU_NOALLOC_BYTES >> 20);
        return BLTCU_STATUS_INPUT_ERROR;
    }
    if (device < 0 || device >= BLTCU_NOALLOC_MAXDEV) return BLTCU_STATUS_INPUT_ERROR;
    std::lock_guard<std::mutex> lk(g_mu);
    CK(cudaSetDevice(device));
    if (!g_noalloc_dev[device]) {  // once per device, with that device current
        unsigned char *ptr = nullptr;
        CK(cudaGetSymbolAddress((void **)&ptr, g_noalloc_scratch));
        int v = 0;
        CK(cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, device)); g_noalloc_sms[device] = v;
        CK(cudaDeviceGetAttribute(&v, cudaDevAttrMaxSharedMemoryPerBlockOptin, device)); g_noalloc_smem[device] = (size_t)v;
        g_noalloc_dev[device] = ptr;
    }
    unsigned char *base = g_noalloc_dev[device];
    const int wpc = BLTCU_NOALLOC_WPC;
    const int ctas = std::min(g_noalloc_sms[device], BLTCU_NOALLOC_CTAS);
    const int nwarps = ctas * wpc;

    // same shared-memory and workspace carving as search_impl(), but inside the fixed array
    const int KMAX = n + 1, LD = KMAX | 1, LDB = m | 1;
    const size_t wd = (size_t)KMAX * LD + 5 * (size_t)KMAX + 2 * (size_t)m + (n + 2) + n;
    const size_t wi = 4 * (size_t)KMAX + (n + 1) + (n + 1);
    const size_t wb = 2 * (size_t)((n + 8) / 8 * 8) + 16;
    const size_t warp_bytes = align_up(wd * 8 + wi * 4 + wb, 16);
    const size_t smem_bytes = (size_t)n * LDB * 8 + (size_t)wpc * warp_bytes;
    if (smem_bytes > g_noalloc_smem[device]) return BLTCU_STATUS_ERROR;
    const RecLayout RL = rec_layout(n, m);
    const size_t scrD_stride = align_up(chain_off(n + 1) + (size_t)(n + 2) * m, 32);
    const size_t scrI_stride = align_up(chain_ioff(n + 1) + 8, 32);
    size_t off = 0;
    auto carve = [&](size_t bytes) { size_t r = off; off = align_up(off + bytes, 256); return r; };
    const size_t o_q = carve(sizeof(QCtl));
    const size_t o_in = carve(((size_t)n * m + 2 * (size_t)m) * 8);
    const size_t o_B = carve((size_t)n * m * 8);
    const size_t o_p = carve((size_t)m * 8);
    const size_t o_ring = carve((size_t)BLTCU_NOALLOC_POOL * 4);
    const size_t o_fring = carve((size_t)BLTCU_NOALLOC_POOL * 4);
    const size_t o_out = carve((size_t)BLTCU_NOALLOC_POOL * 4);
    const size_t o_sol = carve((size_t)n * 8);
    const size_t o_bu = carve((size_t)nwarps * n * 8);
    const size_t o_bk = carve((size_t)nwarps * n);
    const size_t o_bv = carve((size_t)nwarps * 4);
    const size_t o_pool = carve((size_t)BLTCU_NOALLOC_POOL * RL.stride);
    const size_t o_scrD = carve((size_t)nwarps * scrD_stride * 8);
    const size_t o_scrI = carve((size_t)nwarps * scrI_stride * 4);
    if (off > (size_t)BLTCU_NOALLOC_BYTES) {
        fprintf(stderr, "bltcu_closest_noalloc: internal layout (%zu bytes) exceeds the scratch array\n", off);
        return BLTCU_STATUS_ERROR;
    }
    int64_t *d_lat = (int64_t *)(base + o_in), *d_con = d_lat + (size_t)n * m;
    Params P;
    memset(&P, 0, sizeof P);
    P.n = n; P.m = m; P.LD = LD; P.LDB = LDB; P.KMAX = KMAX;
    P.B = (double *)(base + o_B); P.p = (double *)(base + o_p);
    P.lat = d_lat; P.con = d_con;
    P.q = (QCtl *)(base + o_q);
    P.ring = (int *)(base + o_ring); P.fring = (int *)(base + o_fring); P.ring_mask = BLTCU_NOALLOC_POOL - 1;
    P.pool = base + o_pool; P.rec_stride = RL.stride; P.pool_cap = BLTCU_NOALLOC_POOL;
    P.scrD = (double *)(base + o_scrD); P.scrD_stride = scrD_stride;
    P.scrI = (int *)(base + o_scrI); P.scrI_stride = scrI_stride;
    P.sol = (int64_t *)(base + o_sol);
    P.best_u = (int64_t *)(base + o_bu); P.best_key = base + o_bk; P.best_valid = (int *)(base + o_bv);
    P.out_list = (int *)(base + o_out);
    P.static_depth = 0; P.dynamic = 1;
    P.donate_min_level = std::min(6, std::max(2, n / 3));
    P.lowwater = std::min(nwarps, BLTCU_NOALLOC_POOL / 4);
    P.no_early_exit = 0;
    P.warps_per_cta = wpc;
    P.warp_smem_doubles = warp_bytes / 8;
    P.radius = (double)(con[1] - con[0]) / 2.0;  // closest.cc:193-194: row 0 only

    cudaStream_t st = 0;  // legacy default stream: creating a stream would be an allocation
    CK(cudaMemcpyAsync(d_lat, lat, (size_t)n * m * 8, cudaMemcpyHostToDevice, st));
    CK(cudaMemcpyAsync(d_con, con, (size_t)2 * m * 8, cudaMemcpyHostToDevice, st));
    bltcu_setup_kernel<<<(n * m + 255) / 256, 256, 0, st>>>(d_lat, d_con, n, m, (double *)P.B, (double *)P.p,
                                                          (unsigned long long *)P.sol);
    CK(cudaGetLastError());
    unsigned long long deadline_ns = 0;
    if (const char *e = getenv("BLTCU_TIMEOUT_MS")) {
        const long long tmo = atoll(e);
        if (tmo > 0) {
            unsigned long long now_ns = 0;
            CK(cudaMemcpyAsync(&now_ns, P.sol, 8, cudaMemcpyDeviceToHost, st));
            CK(cudaStreamSynchronize(st));
            deadline_ns = now_ns + (unsigned long long)tmo * 1000000ull;
        }
    }
    bltcu_noalloc_init_kernel<<<1, 256, 0, st>>>(P, deadline_ns, nwarps);  // one block: its barrier orders the clears
    CK(cudaGetLastError());
    CK(cudaFuncSetAttribute(bltcu_search_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes));
    bltcu_search_kernel<<<ctas, wpc * 32, smem_bytes, st>>>(P);
    CK(cudaGetLastError());
    static QCtl hq;  // host copies live in static storage: no heap use on the host either
    static int64_t h_bu[BLTCU_NOALLOC_CTAS * BLTCU_NOALLOC_WPC * BLTCU_NOALLOC_MAX_DIM];
    static unsigned char h_bk[BLTCU_NOALLOC_CTAS * BLTCU_NOALLOC_WPC * BLTCU_NOALLOC_MAX_DIM];
    static int h_bv[BLTCU_NOALLOC_CTAS * BLTCU_NOALLOC_WPC];
    CK(cudaMemcpyAsync(&hq, P.q, sizeof hq, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    if (hq.timed_out) fprintf(stderr, "bltcu_closest_noalloc: watchdog fired (where=%d)\n", hq.timed_out);
    int64_t hsol[BLTCU_NOALLOC_MAX_DIM] = {0};
    if (hq.found) {
        CK(cudaMemcpy(hsol, P.sol, (size_t)n * 8, cudaMemcpyDeviceToHost));
    } else {  // last recorded candidate in depth-first order, as the reference returns on UNSAT
        CK(cudaMemcpy(h_bv, P.best_valid, (size_t)nwarps * 4, cudaMemcpyDeviceToHost));
        CK(cudaMemcpy(h_bk, P.best_key, (size_t)nwarps * n, cudaMemcpyDeviceToHost));
        CK(cudaMemcpy(h_bu, P.best_u, (size_t)nwarps * n * 8, cudaMemcpyDeviceToHost));
        int bestw = -1;
        for (int i = 0; i < nwarps; ++i) {
            if (!h_bv[i]) continue;
            bool ge = true;
            if (bestw >= 0) {
                for (int e = n - 1; e >= 0; --e) {  // ordinal paths, top level first
                    const int a = h_bk[(size_t)i * n + e], b = h_bk[(size_t)bestw * n + e];
                    if (a != b) { ge = a > b; break; }
                }
            }
            if (ge) bestw = i;
        }
        if (bestw >= 0) for (int e = 0; e < n; ++e) hsol[e] = h_bu[(size_t)bestw * n + e];
    }
    if (sol) for (int e = 0; e < n; ++e) sol[e] = hsol[e];
    if (niter) *niter = hq.niter;
    if (nprune) *nprune = hq.nprune;
    if (found) *found = hq.found;
    return hq.status;
}
