/*
 * blt_oracle.c -- CPU restatement of BLT's lattice-search hot path.  TEST INFRASTRUCTURE ONLY.
 *
 * This file is the parity oracle for the CUDA search kernels in blt_b200/csrc/.  Only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load it; the product
 * (libbltcu.so, libblt.so, run_blt) never links or calls anything in oracle/.
 *
 * What it restates (all file:line relative to /root/reference):
 *   - closest()            libblt/closest.cc:166-327   (DFS loop, zig-zag order, one-prune-per-side rule)
 *   - SearchState          libblt/closest.cc:40-129    (set_index/descend/explore/ascend/record)
 *   - check_soln()         libblt/closest.cc:330-340   (exact int64 u*lat within bounds)
 *   - sgn()                libblt/closest.cc:343-346   (sgn(0) = -1)
 *   - infdist()/glpk_infdist()  libblt/glpk_wrapper.cc:33-70, :120-232
 *                          (the LP  min t  s.t. |sum_{i<k} x_i B_ij + sum_{i>=k} u_i B_ij - p_j| <= t)
 *   - approximatelyLE()    libblt/util.cc:17-31        (Knuth relative comparison; the prune predicate)
 *
 * PARITY STATUS -- read this.  The LP itself is solved in the reference by GLPK 4.55 (glp_simplex, primal,
 * cold slack basis per node), a third-party dependency that is not under /root/reference and cannot be built in
 * this image (no GLPK/NTL/Boost headers).  This oracle therefore restates the *published problem* (the LP above,
 * set up exactly as glpk_wrapper.cc:141-186 does) and solves it with its own deterministic exchange ("dual
 * simplex on reference sets") method, specified in DESIGN.md section 3.  Pinned against the reference's own known
 * answers: every infdist value / minimiser / rounding in libblt/test_glpk.cc:35-169, the prune sets of
 * libblt/test_yices.cc:37-123, closest() results of libblt/test_closest.cc and the SAT decisions of
 * libblt/test_regression.cc:84-106 (tests/test_oracle_kats.py).  NOT pinned by anything in the reference, hence
 * "parity unpinned" for these: niter/nprune, which optimal vertex x an LP returns when it is not unique, and
 * behaviour after an LP failure.  For those, "reference" means this oracle.
 *
 * Deliberate, documented divergences from the reference (they only matter where the reference is itself
 * ill-defined):
 *   D1. infdist's process-global rescale-and-retry state (glpk_wrapper.cc:39-62) is replaced by "scale 2^0 only;
 *       failure => BLT_STATUS_LP_FAIL".
 *   D2. The right-hand side c = p - sum_{i>=k} u_i B_i is carried incrementally per level with one fma per
 *       element instead of being re-summed from scratch (glpk_wrapper.cc:156-166).  Both are exact, hence
 *       identical, whenever every partial sum is an integer (or half-integer) below 2^53.
 *   D3. A node is abandoned ("pruned") as soon as a dual-feasible basis proves t > radius*(1+eps); the reference
 *       solves to optimality and then applies the same predicate.  Same decision, fewer pivots.
 *   D4. n == 1 and the leaf passes the LP test but fails the exact test: the reference indexes u(1) out of
 *       bounds (closest.cc:99-104); we return BLT_STATUS_OK with uhat recorded.
 *
 * Floating-point contract (so that the CUDA kernels can match bit for bit): compile with -ffp-contract=off;
 * every multiply-add that the spec fuses is an explicit fma(); reductions use dot4()/treedot() below, whose
 * summation order mirrors what a 32-lane warp does.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define BLT_STATUS_OK 0
#define BLT_STATUS_ERROR 1
#define BLT_STATUS_LP_FAIL 2
#define BLT_STATUS_INPUT_ERROR 3
#define ORACLE_TRUNCATED 100 /* node limit hit (bench sampling only) */

#define DIST_EPS 1.0e-4   /* glpk_wrapper.h:13 */
#define EPS_OPT 1.0e-9    /* optimality tolerance, relative to (|c|_inf + t) */
#define PIV_TOL 1.0e-11   /* ratio-test pivot threshold, relative to |w|_inf */

enum { NODE_EXPLORE = 0, NODE_PRUNE = 1, NODE_FAIL = -1 };

/* ---- util.cc:17-31, literal ---- */
static int approximatelyEqual(double a, double b, double eps) {
    return fabs(a - b) <= ((fabs(a) < fabs(b) ? fabs(b) : fabs(a)) * eps);
}
static int definitelyLess(double a, double b, double eps) {
    return (b - a) > ((fabs(a) < fabs(b) ? fabs(b) : fabs(a)) * eps);
}
static int approximatelyLE(double a, double b, double eps) {
    return definitelyLess(a, b, eps) || approximatelyEqual(a, b, eps);
}
/* closest.cc:343-346 */
static int sgn(double z) { return z > 0 ? 1 : -1; }

/* ---- reduction primitives (order is part of the spec) ---- */
/* element e goes to accumulator e%4, ascending e, fused multiply-add; result (a0+a1)+(a2+a3) */
static double dot4(const double *a, int sa, const double *b, int sb, int len) {
    double acc[4] = {0.0, 0.0, 0.0, 0.0};
    for (int e = 0; e < len; ++e) acc[e & 3] = fma(a[(size_t)e * sa], b[(size_t)e * sb], acc[e & 3]);
    return (acc[0] + acc[1]) + (acc[2] + acc[3]);
}
/* 32 lane-strided partial sums (lane l takes e = l, l+32, ...), then xor-butterfly 16,8,4,2,1 */
static double treedot(const double *a, const double *b, int len) {
    double p[32], q[32];
    for (int l = 0; l < 32; ++l) {
        double acc = 0.0;
        for (int e = l; e < len; e += 32) acc = fma(a[e], b[e], acc);
        p[l] = acc;
    }
    for (int off = 16; off >= 1; off >>= 1) {
        for (int l = 0; l < 32; ++l) q[l] = p[l] + p[l ^ off];
        memcpy(p, q, sizeof p);
    }
    return p[0];
}

/* ---- LP engine: basis inverse H (K x K), reference rows/signs, current point z = (t, x_0..x_{K-2}) ---- */
typedef struct {
    int n, m, ld;
    const double *B; /* n x m row-major, (double)lat  (closest.cc:203) */
    int K;
    double *H;       /* ld x ld; H[v*ld+s], v = variable (0 = t, i+1 = x_i), s = reference slot */
    int *row, *sg;   /* per slot: lattice coordinate index j (or -1 = artificial), sign sigma */
    double *z, *w, *g, *a, *cS, *lam;
    int *tmpi, *tmps;
    unsigned char *used;
    uint64_t pivots, refactors, quick, colds;
} LP;

static void lp_gather(LP *lp, const double *c) {
    for (int s = 0; s < lp->K; ++s) lp->cS[s] = lp->row[s] >= 0 ? c[lp->row[s]] : 0.0;
}
static void lp_point(LP *lp) { /* z = H cS, one dot4 per variable */
    for (int v = 0; v < lp->K; ++v) lp->z[v] = dot4(lp->H + (size_t)v * lp->ld, 1, lp->cS, 1, lp->K);
}
/* entering row for coordinate q with sign sq: a = (sq, B[0][q], ..., B[K-2][q]) */
static void lp_row(LP *lp, int q, int sq) {
    lp->a[0] = (double)sq;
    for (int i = 0; i + 1 < lp->K; ++i) lp->a[i + 1] = lp->B[(size_t)i * lp->m + q];
}
static void lp_w(LP *lp) { /* w_s = sum_v a_v H[v][s] */
    for (int s = 0; s < lp->K; ++s) lp->w[s] = dot4(lp->a, 1, lp->H + s, lp->ld, lp->K);
}
/* replace slot l by the row whose coefficients w.r.t. the current rows are w: H <- H - g (w - e_l)^T, g = H[:,l]/w_l */
static void lp_pivot(LP *lp, int l) {
    const int K = lp->K, ld = lp->ld;
    const double wl = lp->w[l];
    for (int v = 0; v < K; ++v) lp->g[v] = lp->H[(size_t)v * ld + l] / wl;
    for (int v = 0; v < K; ++v) {
        double *Hv = lp->H + (size_t)v * ld;
        const double gv = lp->g[v];
        for (int s = 0; s < K; ++s) Hv[s] = (s == l) ? gv : fma(-gv, lp->w[s], Hv[s]);
    }
    lp->pivots++;
}
/* ratio test: among slots with sq*sg_s*w_s > PIV_TOL*|w|_inf pick min max(0,sg_s*lambda_s)/(sq*sg_s*w_s); ties: lowest s */
static int lp_ratio(LP *lp, int sq) {
    const int K = lp->K;
    double wmax = 0.0;
    for (int s = 0; s < K; ++s) if (fabs(lp->w[s]) > wmax) wmax = fabs(lp->w[s]);
    const double thr = PIV_TOL * wmax;
    int best = -1;
    double br = 0.0;
    for (int s = 0; s < K; ++s) {
        const double den = (double)(sq * lp->sg[s]) * lp->w[s];
        if (!(den > thr)) continue;
        double num = (double)lp->sg[s] * lp->H[s]; /* row 0 of H = lambda */
        if (num < 0.0) num = 0.0;
        const double r = num / den;
        if (best < 0 || r < br) { best = s; br = r; }
    }
    return best;
}

/* rebuild H from the current reference rows/signs (Gauss-Jordan as K pivots from the identity) */
static int lp_refactor(LP *lp) {
    const int K = lp->K, ld = lp->ld;
    for (int s = 0; s < K; ++s) { lp->tmpi[s] = lp->row[s]; lp->tmps[s] = lp->sg[s]; lp->row[s] = -1; lp->sg[s] = 1; }
    for (int v = 0; v < K; ++v) for (int s = 0; s < K; ++s) lp->H[(size_t)v * ld + s] = (v == s) ? 1.0 : 0.0;
    for (int i = 0; i < K; ++i) {
        lp_row(lp, lp->tmpi[i], lp->tmps[i]);
        lp_w(lp);
        int l = -1; double bw = 0.0;
        for (int s = 0; s < K; ++s) if (lp->row[s] < 0 && fabs(lp->w[s]) > bw) { bw = fabs(lp->w[s]); l = s; }
        if (l < 0) return -1;
        lp_pivot(lp, l);
        lp->row[l] = lp->tmpi[i]; lp->sg[l] = lp->tmps[i];
    }
    lp->refactors++;
    return 0;
}

static double residual(const LP *lp, const double *c, int j) {
    return c[j] - dot4(lp->B + j, lp->m, lp->z + 1, 1, lp->K - 1);
}

/* cold start for LP(K-1) with right-hand side c: build a dual-feasible basis of K real rows.
 * If K-1 == m (square system, only reachable for the root solve) there is no K-th row: stop after stage A with
 * the interpolating point in z and return 1. */
static int lp_cold(LP *lp, int K, const double *c) {
    const int ld = lp->ld, m = lp->m, kk = K - 1;
    lp->K = K; lp->colds++;
    for (int s = 0; s < K; ++s) { lp->row[s] = -1; lp->sg[s] = 1; }
    for (int v = 0; v < K; ++v) for (int s = 0; s < K; ++s) lp->H[(size_t)v * ld + s] = (v == s) ? 1.0 : 0.0;
    memset(lp->used, 0, (size_t)m);
    /* stage A: one real row per x-variable slot, largest pivot first */
    for (int v = 1; v <= kk; ++v) {
        int q = -1; double bp = 0.0;
        for (int j = 0; j < m; ++j) {
            if (lp->used[j]) continue;
            lp_row(lp, j, 1);
            const double pv = fabs(dot4(lp->a, 1, lp->H + v, ld, K));
            if (pv > bp) { bp = pv; q = j; }
        }
        if (q < 0) return -1;
        lp_row(lp, q, 1); lp_w(lp); lp_pivot(lp, v);
        lp->row[v] = q; lp->sg[v] = 1; lp->used[q] = 1;
    }
    lp_gather(lp, c); lp_point(lp);
    if (kk >= m) return 1;
    /* stage B: the unused row with the largest interpolation residual takes slot 0 (the t slot) */
    {
        int q = -1; double br = -1.0;
        for (int j = 0; j < m; ++j) {
            if (lp->used[j]) continue;
            const double r = fabs(residual(lp, c, j));
            if (r > br) { br = r; q = j; }
        }
        if (q < 0) return -1;
        const double rho = dot4(lp->B + q, m, lp->H + ld, ld, kk); /* sum_i B[i][q] * H[i+1][0] */
        const int sq = (rho >= 0.0) ? 1 : -1;
        lp_row(lp, q, sq); lp_w(lp);
        if (lp->w[0] == 0.0) return -1;
        lp_pivot(lp, 0);
        lp->row[0] = q; lp->sg[0] = sq; lp->used[q] = 1;
    }
    /* sign fix: sigma'_s = sign(lambda_s); Sherman-Morrison on the t column of M */
    {
        int any = 0;
        for (int s = 0; s < K; ++s) {
            const double l = lp->H[s];
            lp->lam[s] = l;
            const int ns = l > 0.0 ? 1 : (l < 0.0 ? -1 : lp->sg[s]);
            lp->w[s] = (double)(ns - lp->sg[s]); /* Delta */
            lp->cS[s] = (double)ns;
            if (ns != lp->sg[s]) any = 1;
            lp->tmps[s] = ns;
        }
        if (any) {
            const double D = dot4(lp->lam, 1, lp->cS, 1, K);
            if (!(D > 0.0)) return -1;
            for (int v = 0; v < K; ++v) lp->g[v] = dot4(lp->H + (size_t)v * ld, 1, lp->w, 1, K) / D;
            for (int v = 0; v < K; ++v) {
                double *Hv = lp->H + (size_t)v * ld;
                for (int s = 0; s < K; ++s) Hv[s] = fma(-lp->g[v], lp->lam[s], Hv[s]);
            }
            for (int s = 0; s < K; ++s) lp->sg[s] = lp->tmps[s];
        }
    }
    return 0;
}

static void lp_flip(LP *lp) { /* flip every sign: negate row 0 of H */
    for (int s = 0; s < lp->K; ++s) { lp->H[s] = -lp->H[s]; lp->sg[s] = -lp->sg[s]; }
    lp->z[0] = -lp->z[0];
}

/* Solve the node LP(K-1) from the current (dual-feasible) basis.  radius = +inf disables pruning. */
static int lp_solve(LP *lp, const double *c, double radius, double *d_out, double *xlast_out) {
    const int K = lp->K, m = lp->m;
    lp_gather(lp, c);
    /* quick bound: |lambda . cS| is a lower bound on the optimum */
    {
        double tq = treedot(lp->H, lp->cS, K);
        if (tq < 0.0) tq = -tq;
        if (!approximatelyLE(tq, radius, DIST_EPS)) { lp->quick++; return NODE_PRUNE; }
    }
    lp_point(lp);
    if (lp->z[0] < 0.0) lp_flip(lp);
    double cmax = 0.0;
    for (int j = 0; j < m; ++j) if (fabs(c[j]) > cmax) cmax = fabs(c[j]);
    const int cap = 100 * K + 1000;
    int refactored = 0;
    for (int iter = 0;; ++iter) {
        const double t = lp->z[0];
        if (!approximatelyLE(t, radius, DIST_EPS)) return NODE_PRUNE;
        int q = -1; double rmax = -1.0, rq = 0.0;
        for (int j = 0; j < m; ++j) {
            const double r = residual(lp, c, j);
            if (fabs(r) > rmax) { rmax = fabs(r); q = j; rq = r; }
        }
        const double tol = EPS_OPT * (cmax + t);
        if (rmax <= t + tol) { *d_out = t; *xlast_out = lp->z[K - 1]; return NODE_EXPLORE; }
        if (iter >= cap) return NODE_FAIL;
        int dup = 0;
        for (int s = 0; s < K; ++s) if (lp->row[s] == q) dup = 1;
        if (dup) { /* a reference row is violated: H has drifted; rebuild once */
            if (refactored) return NODE_FAIL;
            refactored = 1;
            if (lp_refactor(lp)) return NODE_FAIL;
            lp_gather(lp, c); lp_point(lp);
            if (lp->z[0] < 0.0) lp_flip(lp);
            continue;
        }
        const int sq = rq > 0.0 ? 1 : -1;
        lp_row(lp, q, sq); lp_w(lp);
        const int l = lp_ratio(lp, sq);
        if (l < 0) return NODE_FAIL;
        lp_pivot(lp, l);
        lp->row[l] = q; lp->sg[l] = sq;
        const double delta = (double)sq * (rmax - t);
        for (int v = 0; v < K; ++v) lp->z[v] = fma(lp->g[v], delta, lp->z[v]);
    }
}

/* child transform: variable x_{K-2} (last row of H) becomes fixed; dir = sgn(x_last - u).  K -> K-1. */
static int lp_child(LP *lp, int dir) {
    const int K = lp->K, ld = lp->ld;
    for (int s = 0; s < K; ++s) lp->w[s] = lp->H[(size_t)(K - 1) * ld + s];
    const int l = lp_ratio(lp, -dir);
    if (l < 0) return -1;
    const double wl = lp->w[l];
    for (int v = 0; v < K - 1; ++v) lp->g[v] = lp->H[(size_t)v * ld + l] / wl;
    for (int v = 0; v < K - 1; ++v) {
        double *Hv = lp->H + (size_t)v * ld;
        const double gv = lp->g[v];
        for (int s = 0; s < K; ++s) if (s != l) Hv[s] = fma(-gv, lp->w[s], Hv[s]);
        if (l != K - 1) Hv[l] = Hv[K - 1];
    }
    if (l != K - 1) { lp->row[l] = lp->row[K - 1]; lp->sg[l] = lp->sg[K - 1]; }
    lp->K = K - 1;
    lp->pivots++;
    return 0;
}

static LP *lp_new(int n, int m, const double *B) {
    LP *lp = (LP *)calloc(1, sizeof(LP));
    const int ld = n + 1;
    lp->n = n; lp->m = m; lp->ld = ld; lp->B = B;
    lp->H = (double *)calloc((size_t)ld * ld, sizeof(double));
    lp->row = (int *)calloc(ld, sizeof(int)); lp->sg = (int *)calloc(ld, sizeof(int));
    lp->tmpi = (int *)calloc(ld, sizeof(int)); lp->tmps = (int *)calloc(ld, sizeof(int));
    lp->z = (double *)calloc(ld, sizeof(double)); lp->w = (double *)calloc(ld, sizeof(double));
    lp->g = (double *)calloc(ld, sizeof(double)); lp->a = (double *)calloc(ld, sizeof(double));
    lp->cS = (double *)calloc(ld, sizeof(double)); lp->lam = (double *)calloc(ld, sizeof(double));
    lp->used = (unsigned char *)calloc(m > 0 ? m : 1, 1);
    return lp;
}
static void lp_free(LP *lp) {
    free(lp->H); free(lp->row); free(lp->sg); free(lp->tmpi); free(lp->tmps); free(lp->z); free(lp->w);
    free(lp->g); free(lp->a); free(lp->cS); free(lp->lam); free(lp->used); free(lp);
}

/* ---- closest.cc:330-340 ---- */
int oracle_check_soln(const int64_t *lat, const int64_t *con, int n, int m, const int64_t *u) {
    for (int j = 0; j < m; ++j) {
        uint64_t z = 0; /* wraps like the reference's int64 prod() */
        for (int i = 0; i < n; ++i) z += (uint64_t)u[i] * (uint64_t)lat[(size_t)i * m + j];
        const int64_t zz = (int64_t)z;
        if (con[2 * j] > zz || zz > con[2 * j + 1]) return 0;
    }
    return 1;
}

/* ---- infdist (glpk_wrapper.cc:120-232) as a standalone cold solve: used for the reference's LP known answers.
 * B is n x m doubles, u has n entries, k free coefficients.  x receives k values.  Returns 0 on success. ---- */
int oracle_infdist(const double *B, int n, int m, const int64_t *u, int k, const double *p, double *x, double *dist) {
    if (k < 0 || k > n || k > m) return -1;
    double *c = (double *)calloc((size_t)m, sizeof(double));
    for (int i = 0; i < m; ++i) { /* glpk_wrapper.cc:156-166, literal order */
        double rhs1 = 0.0;
        for (int j = k; j < n; ++j) rhs1 += (double)(-u[j]) * B[(size_t)j * m + i];
        rhs1 += p[i];
        c[i] = rhs1;
    }
    int rc = 0;
    if (k == 0) {
        double d = 0.0;
        for (int j = 0; j < m; ++j) if (fabs(c[j]) > d) d = fabs(c[j]);
        *dist = d;
    } else {
        LP *lp = lp_new(n, m, B);
        const int st = lp_cold(lp, k + 1, c);
        if (st < 0) rc = -1;
        else if (st == 1) {
            double d = 0.0;
            for (int j = 0; j < m; ++j) { const double r = fabs(residual(lp, c, j)); if (r > d) d = r; }
            *dist = d;
            for (int i = 0; i < k; ++i) x[i] = lp->z[i + 1];
        } else {
            double d, xl;
            if (lp_solve(lp, c, INFINITY, &d, &xl) != NODE_EXPLORE) rc = -1;
            else { *dist = d; for (int i = 0; i < k; ++i) x[i] = lp->z[i + 1]; }
        }
        lp_free(lp);
    }
    free(c);
    return rc;
}

/* ---- closest() ----
 * lat: n x m int64 row-major (rows = basis vectors), con: m x 2.  sol: n (uhat).  hist: n+1 counters, hist[kk] =
 * number of layer-distance solves with kk free coefficients.  stats[0..3] = pivots, refactors, quick prunes, cold
 * starts.  max_nodes = 0 means unlimited, otherwise stop with ORACLE_TRUNCATED after that many solves.
 *
 * Warm-start policy ("side chains"): at every level the siblings below the real centre xc and
 * the siblings above it form two chains.  The first node of a chain starts from the parent's optimal basis with the
 * level's coefficient fixed (lp_child); every later node of the chain starts from the final basis of the previous
 * node on the same side.  The basis a node starts from is therefore a function of its path alone. */
typedef struct { double *H; int *row, *sg; int init; } Chain;

static void chain_store(Chain *ch, const LP *lp) {
    const int K = lp->K, ld = lp->ld;
    memcpy(ch->H, lp->H, (size_t)K * ld * sizeof(double));
    memcpy(ch->row, lp->row, K * sizeof(int)); memcpy(ch->sg, lp->sg, K * sizeof(int));
}
static void chain_load(LP *lp, const Chain *ch, int K) {
    const int ld = lp->ld;
    lp->K = K;
    memcpy(lp->H, ch->H, (size_t)K * ld * sizeof(double));
    memcpy(lp->row, ch->row, K * sizeof(int)); memcpy(lp->sg, ch->sg, K * sizeof(int));
}

int oracle_closest(const int64_t *lat, const int64_t *con, int n, int m, int64_t *sol, uint64_t *niter_out,
                   uint64_t *nprune_out, uint64_t *hist, uint64_t *stats, uint64_t max_nodes) {
    if (n <= 0 || m <= 0 || m < n) return BLT_STATUS_INPUT_ERROR; /* closest.cc:173-177 */
    double *B = (double *)malloc((size_t)n * m * sizeof(double));
    for (size_t i = 0; i < (size_t)n * m; ++i) B[i] = (double)lat[i];
    double *p = (double *)malloc((size_t)m * sizeof(double));
    for (int j = 0; j < m; ++j) p[j] = (double)(con[2 * j + 1] + con[2 * j]) / 2.0; /* closest.cc:180-185 */
    const double radius = (double)(con[1] - con[0]) / 2.0;                          /* closest.cc:193-194: row 0 only */

    LP *lp = lp_new(n, m, B);
    const int ld = lp->ld;
    int64_t *u = (int64_t *)calloc(n, sizeof(int64_t)), *uhat = (int64_t *)calloc(n, sizeof(int64_t));
    int64_t *step = (int64_t *)calloc(n, sizeof(int64_t));
    unsigned char *pruned = (unsigned char *)calloc(n, 1);
    double *c = (double *)calloc(m, sizeof(double));
    /* per level k (state k solves LP(k-1)): two chains, the parent's right-hand side, the real centre, the side of
     * the node we descended through */
    Chain (*ch)[2] = (Chain (*)[2])calloc(n + 2, sizeof(Chain[2]));
    double **cparL = (double **)calloc(n + 2, sizeof(double *));
    double *xc = (double *)calloc(n + 2, sizeof(double));
    int *curside = (int *)calloc(n + 2, sizeof(int));
    for (int k = 0; k <= n + 1; ++k) {
        cparL[k] = (double *)calloc(m, sizeof(double));
        for (int s = 0; s < 2; ++s) {
            ch[k][s].H = (double *)calloc((size_t)ld * ld, sizeof(double));
            ch[k][s].row = (int *)calloc(ld, sizeof(int)); ch[k][s].sg = (int *)calloc(ld, sizeof(int));
        }
    }
    uint64_t counter = 0, nprune = 0, nodes = 0;
    int rc = BLT_STATUS_OK;
    int k = n;
    double xlast = 0.0, d = 0.0;
    if (hist) memset(hist, 0, (size_t)(n + 1) * sizeof(uint64_t));

    /* initial solve LP(n): closest.cc:206; only x(n-1) is used */
    {
        memcpy(c, p, (size_t)m * sizeof(double));
        const int st = lp_cold(lp, n + 1, c);
        if (hist) hist[n]++;
        nodes++;
        if (st < 0) { rc = BLT_STATUS_LP_FAIL; goto done; }
        if (st == 1) xlast = lp->z[n];
        else if (lp_solve(lp, c, INFINITY, &d, &xlast) != NODE_EXPLORE) { rc = BLT_STATUS_LP_FAIL; goto done; }
    }
    u[n - 1] = (int64_t)round(xlast);          /* set_index, closest.cc:74-82 */
    step[n - 1] = sgn(xlast - (double)u[n - 1]);
    xc[n] = xlast;
    memcpy(cparL[n], p, (size_t)m * sizeof(double));

    for (;;) {
        if (max_nodes && nodes >= max_nodes) { rc = ORACLE_TRUNCATED; goto done; }
        /* right-hand side of this node: c = cpar - u[k-1] * B[k-1]  (D2) */
        {
            const double mu = -(double)u[k - 1];
            const double *Bk = B + (size_t)(k - 1) * m, *cp = cparL[k];
            for (int j = 0; j < m; ++j) c[j] = fma(mu, Bk[j], cp[j]);
        }
        int status, side = 0;
        nodes++;
        if (hist) hist[k - 1]++;
        if (k == 1) { /* LP(0): plain max-abs */
            d = 0.0;
            for (int j = 0; j < m; ++j) if (fabs(c[j]) > d) d = fabs(c[j]);
            status = approximatelyLE(d, radius, DIST_EPS) ? NODE_EXPLORE : NODE_PRUNE;
        } else {
            const int dir = sgn(xc[k] - (double)u[k - 1]);
            side = dir > 0 ? 0 : 1;
            if (!ch[k][side].init) {
                if (k == n) {
                    if (lp_cold(lp, k, c) != 0) { rc = BLT_STATUS_LP_FAIL; goto done; }
                } else {
                    chain_load(lp, &ch[k + 1][curside[k + 1]], k + 1);
                    if (lp_child(lp, dir)) { rc = BLT_STATUS_LP_FAIL; goto done; }
                }
                ch[k][side].init = 1;
            } else chain_load(lp, &ch[k][side], k);
            status = lp_solve(lp, c, radius, &d, &xlast);
            if (status == NODE_FAIL) { rc = BLT_STATUS_LP_FAIL; goto done; }
            chain_store(&ch[k][side], lp);
        }
        if (status == NODE_PRUNE) {
            if (k < n) { /* ascend, closest.cc:110-122 */
                if (pruned[k - 1]) k++;
                else pruned[k - 1] = 1;
                u[k - 1] += step[k - 1];
                step[k - 1] = -step[k - 1] - sgn((double)step[k - 1]);
                nprune++;
                continue;
            }
            rc = BLT_STATUS_OK; /* closest.cc:280-286: first prune at the top level ends the search */
            goto done;
        }
        /* explore */
        if (k > 1) { /* descend, closest.cc:88-95 */
            curside[k] = side;
            memcpy(cparL[k - 1], c, (size_t)m * sizeof(double));
            xc[k - 1] = xlast;
            k--;
            u[k - 1] = (int64_t)round(xlast);
            step[k - 1] = sgn(xlast - (double)u[k - 1]);
            pruned[k - 1] = 0;
            ch[k][0].init = ch[k][1].init = 0;
        } else {
            memcpy(uhat, u, n * sizeof(int64_t)); /* record */
            if (oracle_check_soln(lat, con, n, m, u)) { rc = BLT_STATUS_OK; goto done; }
            if (n == 1) { rc = BLT_STATUS_OK; goto done; } /* D4 */
            /* explore(), closest.cc:99-104 */
            k = 2;
            u[1] += step[1];
            step[1] = -step[1] - sgn((double)step[1]);
        }
        counter++;
    }
done:
    memcpy(sol, uhat, n * sizeof(int64_t));
    if (niter_out) *niter_out = counter;
    if (nprune_out) *nprune_out = nprune;
    if (stats) { stats[0] = lp->pivots; stats[1] = lp->refactors; stats[2] = lp->quick; stats[3] = lp->colds; }
    for (int kk = 0; kk <= n + 1; ++kk) {
        free(cparL[kk]);
        for (int s = 0; s < 2; ++s) { free(ch[kk][s].H); free(ch[kk][s].row); free(ch[kk][s].sg); }
    }
    free(ch); free(cparL); free(xc); free(curside);
    free(u); free(uhat); free(step); free(pruned); free(c); free(B); free(p);
    lp_free(lp);
    return rc;
}
