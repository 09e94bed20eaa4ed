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
 * set up exactly as glpk_wrapper.cc:141-186 does) and solves it with its own deterministic dual simplex in the
 * coordinates of the orthogonal complement of the free basis vectors, specified in DESIGN.md section 3.  Pinned against the reference's own known
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
#define HARRIS_DELTA 1.0e-10 /* dual infeasibility the ratio test tolerates */
#define STALL_MAX 40      /* degenerate steps before the anti-cycling rule takes over */
#define PIV_REL 1.0e-9    /* preferred pivots are at least this large next to the leaving row's scale */

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

static double butterfly32(double *p) {
    double q[32];
    for (int off = 16; off >= 1; off >>= 1) {
        for (int l = 0; l < 32; ++l) q[l] = p[l] + p[l ^ off];
        memcpy(p, q, 32 * sizeof(double));
    }
    return p[0];
}

/* ---- geometry shared by every node of one problem: B^T = Q R (orthonormal Q by Gram-Schmidt applied twice) ----
 * q_0..q_{n-1} span the basis rows in order, q_n..q_{m-1} complete R^m.  Qrev[j][i] = q_{m-1-i}[j], so that the
 * first `qd` entries of row j of Qrev are row j of N = [q_{m-1} .. q_{m-qd}], the complement of the first m-qd
 * directions.  Rm[j][i] = q_j . b_i (upper triangular), QTp[j] = q_j . p. */
typedef struct {
    int n, m;
    double *Q;    /* m x m, row t = vector q_t */
    double *Qrev; /* m x m */
    double *Rm;   /* n x n */
    double *QTp;  /* m */
} Geo;

static void geo_project(const Geo *g, int t, double *v, double *coef) { /* v -= (q_j . v) q_j for j < t, twice */
    const int m = g->m;
    for (int pass = 0; pass < 2; ++pass)
        for (int j = 0; j < t; ++j) {
            const double *qj = g->Q + (size_t)j * m;
            const double h = treedot(qj, v, m);
            for (int e = 0; e < m; ++e) v[e] = fma(-h, qj[e], v[e]);
            if (coef) coef[j] += h;
        }
}
static int geo_build(Geo *g, const double *B, const double *p, int n, int m) {
    g->n = n; g->m = m;
    g->Q = (double *)calloc((size_t)m * m, sizeof(double)); g->Qrev = (double *)calloc((size_t)m * m, sizeof(double));
    g->Rm = (double *)calloc((size_t)n * n, sizeof(double)); g->QTp = (double *)calloc(m, sizeof(double));
    double *v = (double *)calloc(m, sizeof(double)), *coef = (double *)calloc(m, sizeof(double));
    int ok = 1;
    for (int t = 0; t < m && ok; ++t) {
        if (t < n) memcpy(v, B + (size_t)t * m, m * sizeof(double));
        else { /* completion: the unit vector with the largest part outside the current span (ties: lowest j) */
            int best = -1; double bv = -1.0;
            for (int j = 0; j < m; ++j) {
                double sres = 1.0;
                for (int i = 0; i < t; ++i) { const double x = g->Q[(size_t)i * m + j]; sres = fma(-x, x, sres); }
                if (sres > bv) { bv = sres; best = j; }
            }
            memset(v, 0, m * sizeof(double)); v[best] = 1.0;
        }
        memset(coef, 0, m * sizeof(double));
        geo_project(g, t, v, coef);
        const double nrm = sqrt(treedot(v, v, m));
        if (!(nrm > 0.0)) { ok = 0; break; }
        for (int e = 0; e < m; ++e) g->Q[(size_t)t * m + e] = v[e] / nrm;
        if (t < n) { for (int j = 0; j < t; ++j) g->Rm[(size_t)j * n + t] = coef[j]; g->Rm[(size_t)t * n + t] = nrm; }
    }
    for (int j = 0; j < m; ++j) for (int i = 0; i < m; ++i) g->Qrev[(size_t)j * m + i] = g->Q[(size_t)(m - 1 - i) * m + j];
    for (int t = 0; t < m; ++t) g->QTp[t] = treedot(g->Q + (size_t)t * m, p, m);
    free(v); free(coef);
    return ok ? 0 : -1;
}
static void geo_free(Geo *g) { free(g->Q); free(g->Qrev); free(g->Rm); free(g->QTp); }

/* ---- LP engine in complement coordinates -------------------------------------------------------------------
 * Node LP(kk):  min |r|_inf  s.t.  N^T r = chat   (N = first qd = m-kk columns of Qrev, chat = N^T c).
 * Vertex: rows in Z (qd-1 of them, "basic") are strictly inside, all other rows sit at r_j = sg_j t.
 * A = [ s | n_z1 .. n_z(qd-1) ] with s = sum_{j active} sg_j n_j;  H = A^-1;  (t, r_Z) = H chat;  y = row 0 of H is
 * the dual vector, invariant sg_j (n_j . y) >= 0 for active j, hence t = y . chat is a lower bound of the optimum. */
typedef struct {
    const Geo *g;
    int m, ld, qd;
    double *H;
    int *zrow, *pos;     /* zrow[p] = row held by basic position p >= 1 (-1 = artificial); pos[j] = position or 0 */
    signed char *sg;     /* sign of each active row */
    double *chat, *val, *alpha, *beta, *w, *a, *r, *v;
    uint64_t pivots, refactors, quick, colds, repairs;
} LP;

static LP *lp_new(const Geo *g) {
    LP *lp = (LP *)calloc(1, sizeof(LP));
    const int m = g->m, ld = m + 1;
    lp->g = g; lp->m = m; lp->ld = ld;
    lp->H = (double *)calloc((size_t)ld * ld, sizeof(double));
    lp->zrow = (int *)calloc(ld, sizeof(int)); lp->pos = (int *)calloc(m, sizeof(int));
    lp->sg = (signed char *)calloc(m, 1);
    lp->chat = (double *)calloc(ld, sizeof(double)); lp->val = (double *)calloc(ld, sizeof(double));
    lp->alpha = (double *)calloc(m, sizeof(double)); lp->beta = (double *)calloc(m, sizeof(double));
    lp->w = (double *)calloc(ld, sizeof(double)); lp->a = (double *)calloc(ld, sizeof(double));
    lp->r = (double *)calloc(m, sizeof(double)); lp->v = (double *)calloc(ld, sizeof(double));
    return lp;
}
static void lp_free(LP *lp) {
    free(lp->H); free(lp->zrow); free(lp->pos); free(lp->sg); free(lp->chat); free(lp->val); free(lp->alpha);
    free(lp->beta); free(lp->w); free(lp->a); free(lp->r); free(lp->v); free(lp);
}
static void lp_values(LP *lp) {
    for (int p = 0; p < lp->qd; ++p) lp->val[p] = dot4(lp->H + (size_t)p * lp->ld, 1, lp->chat, 1, lp->qd);
}
static void lp_flip(LP *lp) {
    for (int c = 0; c < lp->qd; ++c) lp->H[c] = -lp->H[c];
    for (int j = 0; j < lp->m; ++j) lp->sg[j] = (signed char)-lp->sg[j];
    lp->val[0] = -lp->val[0];
}
/* alpha_j = n_j . y, beta_j = n_j . H[pe] for the active rows; returns |N y|_1 */
static double lp_alpha_beta(LP *lp, int pe) {
    const int m = lp->m, qd = lp->qd;
    double part[32];
    for (int l = 0; l < 32; ++l) part[l] = 0.0;
    for (int j = 0; j < m; ++j) {
        if (lp->pos[j]) { lp->alpha[j] = 0.0; lp->beta[j] = 0.0; continue; }
        const double *nj = lp->g->Qrev + (size_t)j * m;
        lp->alpha[j] = dot4(nj, 1, lp->H, 1, qd);
        lp->beta[j] = dot4(nj, 1, lp->H + (size_t)pe * lp->ld, 1, qd);
        part[j & 31] += fabs(lp->alpha[j]);
    }
    return butterfly32(part);
}
/* Sign repair.  The basis Z fixes the dual vector y up to scale; the signs are supposed to be sg_j = sign(n_j . y).  When
 * round-off has pushed some sg_j*alpha_j below -HARRIS_DELTA those rows are flipped to the other bound: column 0 of A
 * changes by Delta = sum_flipped -2 sg_j n_j, a rank-one correction of H.  Needs alpha from lp_alpha_beta.  Returns 1
 * if anything was flipped. */
static int lp_repair(LP *lp) {
    const int m = lp->m, qd = lp->qd, ld = lp->ld;
    int any = 0;
    for (int j = 0; j < m; ++j) {
        lp->r[j] = 0.0;
        if (!lp->pos[j] && (double)lp->sg[j] * lp->alpha[j] < -HARRIS_DELTA) { lp->r[j] = -2.0 * (double)lp->sg[j]; any = 1; }
    }
    if (!any) return 0;
    for (int c = 0; c < qd; ++c) lp->a[c] = dot4(lp->r, 1, lp->g->Qrev + c, m, m);   /* Delta */
    for (int i = 0; i < qd; ++i) lp->w[i] = dot4(lp->H + (size_t)i * ld, 1, lp->a, 1, qd);
    const double den = 1.0 + lp->w[0];
    for (int c = 0; c < qd; ++c) lp->v[c] = lp->H[c];
    for (int i = 0; i < qd; ++i) {
        const double f = lp->w[i] / den;
        double *Hi = lp->H + (size_t)i * ld;
        for (int c = 0; c < qd; ++c) Hi[c] = fma(-f, lp->v[c], Hi[c]);
    }
    for (int j = 0; j < m; ++j) if (lp->r[j] != 0.0) lp->sg[j] = (signed char)-lp->sg[j];
    lp->repairs++;
    return 1;
}
/* Entering row (Harris two-pass ratio test).  With den_j = -se*sg_j*beta_j and num_j = max(0, sg_j*alpha_j):
 * pass 1: theta_max = min over den_j > dmin of max(0, sg_j*alpha_j + HARRIS_DELTA) / den_j;
 * pass 2: among den_j > dmin with num_j/den_j <= theta_max the largest den_j wins (ties: lowest j).
 * dmin = 0; when an artificial leaves its pivot is den itself, and a winner below 1e-3*PIV_TOL*scale is refused.
 * If no row qualifies and a row of Z leaves, the dual vector moves all the way onto the leaving row's own ray, where
 * exactly the rows with beta_j = 0 vanish: take the one with the largest alpha among |beta_j| <= PIV_TOL*scale. */
static int lp_ratio(LP *lp, int se, int pe, int bland) {
    const int m = lp->m;
    const int artificial = lp->zrow[pe] < 0;
    double scale = 0.0;
    for (int c = 0; c < lp->qd; ++c) if (fabs(lp->H[(size_t)pe * lp->ld + c]) > scale) scale = fabs(lp->H[(size_t)pe * lp->ld + c]);
    for (int j = 0; j < m; ++j) if (!lp->pos[j] && fabs(lp->beta[j]) > scale) scale = fabs(lp->beta[j]);
    const double tiny = PIV_TOL * scale;
    int best = -1; double bd = 0.0;
    /* tier 0 only accepts pivots that are not tiny next to the row's scale (rows below the bar may then drift a little
     * below zero, which the bound tolerates); tier 1 accepts any positive pivot */
    for (int tier = 0; tier < 2 && best < 0; ++tier) {
        const double dmin = tier == 0 ? PIV_REL * scale : 0.0;
        double tmax = INFINITY;
        for (int j = 0; j < m; ++j) {
            if (lp->pos[j]) continue;
            const double den = -(double)(se * lp->sg[j]) * lp->beta[j];
            if (!(den > dmin)) continue;
            double num = (double)lp->sg[j] * lp->alpha[j] + HARRIS_DELTA; /* rows already below zero get less room, not more */
            if (num < 0.0) num = 0.0;
            const double rt = num / den;
            if (rt < tmax) tmax = rt;
        }
        bd = 0.0;
        for (int j = 0; j < m; ++j) {
            if (lp->pos[j]) continue;
            const double den = -(double)(se * lp->sg[j]) * lp->beta[j];
            if (!(den > dmin)) continue;
            double num = (double)lp->sg[j] * lp->alpha[j];
            if (num < 0.0) num = 0.0;
            if (num / den <= tmax && (bland ? best < 0 : den > bd)) { bd = den; best = j; }
        }
    }
    if (artificial) return (best >= 0 && bd > 1.0e-3 * tiny) ? best : -1;  /* its pivot is den itself: refuse a vanishing one */
    if (best >= 0) return best;
    double ba = -1.0;
    for (int j = 0; j < m; ++j) {
        if (lp->pos[j] || fabs(lp->beta[j]) > tiny) continue;
        if (fabs(lp->alpha[j]) > ba) { ba = fabs(lp->alpha[j]); best = j; }
    }
    return best;
}
/* basic position pe (row zrow[pe], or artificial) leaves with sign se, active row l takes the position */
static int lp_replace(LP *lp, int pe, int l, int se) {
    const int qd = lp->qd, ld = lp->ld, m = lp->m;
    double *H = lp->H, *Hp = H + (size_t)pe * ld;
    const int e = lp->zrow[pe];
    if (e >= 0) for (int c = 0; c < qd; ++c) Hp[c] = Hp[c] - (double)se * H[c];       /* s += se * n_e */
    const double *nl = lp->g->Qrev + (size_t)l * m;
    for (int i = 0; i < qd; ++i) lp->w[i] = dot4(H + (size_t)i * ld, 1, nl, 1, qd);
    const double wp = lp->w[pe];
    if (wp == 0.0) return -1;
    for (int c = 0; c < qd; ++c) lp->v[c] = Hp[c] / wp;
    for (int i = 0; i < qd; ++i) {
        if (i == pe) continue;
        double *Hi = H + (size_t)i * ld;
        const double wi = lp->w[i];
        for (int c = 0; c < qd; ++c) Hi[c] = fma(-wi, lp->v[c], Hi[c]);
    }
    const double sl = (double)lp->sg[l];
    for (int c = 0; c < qd; ++c) Hp[c] = lp->v[c] + sl * H[c];                        /* s -= sg_l * n_l */
    if (e >= 0) { lp->sg[e] = (signed char)se; lp->pos[e] = 0; }
    lp->zrow[pe] = l; lp->pos[l] = pe;
    lp->pivots++;
    return 0;
}
/* first coordinate: qd = 1, no basic rows */
static int lp_init1(LP *lp) {
    const int m = lp->m;
    double part[32];
    for (int l = 0; l < 32; ++l) part[l] = 0.0;
    for (int j = 0; j < m; ++j) {
        const double x = lp->g->Qrev[(size_t)j * m];
        lp->sg[j] = x >= 0.0 ? 1 : -1; lp->pos[j] = 0;
        part[j & 31] += fabs(x);
    }
    const double s = butterfly32(part);
    if (!(s > 0.0)) return -1;
    lp->H[0] = 1.0 / s; lp->qd = 1; lp->zrow[0] = -1;
    lp->colds++;
    return 0;
}
/* one more coordinate (index qd): border H with an artificial basic column at position qd */
static void lp_extend(LP *lp) {
    const int m = lp->m, qd = lp->qd, ld = lp->ld;
    const double *Qr = lp->g->Qrev;
    double part[32];
    for (int l = 0; l < 32; ++l) part[l] = 0.0;
    for (int j = 0; j < m; ++j) if (!lp->pos[j]) part[j & 31] = fma((double)lp->sg[j], Qr[(size_t)j * m + qd], part[j & 31]);
    lp->a[0] = butterfly32(part);
    for (int p = 1; p < qd; ++p) lp->a[p] = Qr[(size_t)lp->zrow[p] * m + qd];
    for (int c = 0; c < qd; ++c) lp->H[(size_t)qd * ld + c] = -dot4(lp->a, 1, lp->H + c, ld, qd);
    for (int p = 0; p < qd; ++p) lp->H[(size_t)p * ld + qd] = 0.0;
    lp->H[(size_t)qd * ld + qd] = 1.0;
    lp->zrow[qd] = -1;
    lp->qd = qd + 1;
}
/* drive the artificial at the last position out; sf = sign of its value (= side of the fixed coefficient) */
static int lp_drop_artificial(LP *lp, int sf) {
    const int pe = lp->qd - 1;
    lp_alpha_beta(lp, pe);
    if (lp_repair(lp)) lp_alpha_beta(lp, pe);
    const int l = lp_ratio(lp, sf, pe, 0);
    if (l < 0) return -1;
    return lp_replace(lp, pe, l, sf);
}
/* rebuild H = A^-1 from Z and the signs: plain Gauss-Jordan with row pivoting (rare path) */
static int lp_refactor(LP *lp) {
    const int qd = lp->qd, ld = lp->ld, m = lp->m;
    const double *Qr = lp->g->Qrev;
    double *A = (double *)calloc((size_t)qd * qd, sizeof(double));
    for (int c = 0; c < qd; ++c) { /* A[c][0] = s_c, A[c][p] = n_{z_p}[c] */
        double s = 0.0;
        for (int j = 0; j < m; ++j) if (!lp->pos[j]) s = fma((double)lp->sg[j], Qr[(size_t)j * m + c], s);
        A[(size_t)c * qd] = s;
        for (int p = 1; p < qd; ++p) A[(size_t)c * qd + p] = lp->zrow[p] >= 0 ? Qr[(size_t)lp->zrow[p] * m + c] : (c == p ? 1.0 : 0.0);
    }
    for (int i = 0; i < qd; ++i) for (int c = 0; c < qd; ++c) lp->H[(size_t)i * ld + c] = (i == c) ? 1.0 : 0.0;
    int rc = 0;
    for (int col = 0; col < qd; ++col) {
        int pr = -1; double bv = 0.0;
        for (int i = col; i < qd; ++i) if (fabs(A[(size_t)i * qd + col]) > bv) { bv = fabs(A[(size_t)i * qd + col]); pr = i; }
        if (pr < 0) { rc = -1; break; }
        if (pr != col) {
            for (int c = 0; c < qd; ++c) {
                double t = A[(size_t)pr * qd + c]; A[(size_t)pr * qd + c] = A[(size_t)col * qd + c]; A[(size_t)col * qd + c] = t;
                t = lp->H[(size_t)pr * ld + c]; lp->H[(size_t)pr * ld + c] = lp->H[(size_t)col * ld + c]; lp->H[(size_t)col * ld + c] = t;
            }
        }
        const double pv = A[(size_t)col * qd + col];
        for (int c = 0; c < qd; ++c) { A[(size_t)col * qd + c] /= pv; lp->H[(size_t)col * ld + c] /= pv; }
        for (int i = 0; i < qd; ++i) {
            if (i == col) continue;
            const double f = A[(size_t)i * qd + col];
            if (f == 0.0) continue;
            for (int c = 0; c < qd; ++c) {
                A[(size_t)i * qd + c] = fma(-f, A[(size_t)col * qd + c], A[(size_t)i * qd + c]);
                lp->H[(size_t)i * ld + c] = fma(-f, lp->H[(size_t)col * ld + c], lp->H[(size_t)i * ld + c]);
            }
        }
    }
    free(A);
    lp->refactors++;
    return rc;
}

/* full residual vector and its image: v[i] = sum_j Qrev[j][i] r_j for i < nv */
static void lp_image(LP *lp, int nv) {
    const int m = lp->m;
    for (int j = 0; j < m; ++j) lp->r[j] = lp->pos[j] ? lp->val[lp->pos[j]] : (double)lp->sg[j] * lp->val[0];
    for (int i = 0; i < nv; ++i) lp->v[i] = dot4(lp->g->Qrev + i, m, lp->r, 1, m);
}

/* Solve from the current dual-feasible vertex.  radius = +inf disables pruning.  On NODE_EXPLORE *d_out is the
 * optimum and lp->v[qd] (if qd < m) the next coordinate of the residual's image (used for x_last). */
static int lp_solve(LP *lp, double radius, double cmax, double *d_out) {
    const int qd = lp->qd, m = lp->m;
    lp_values(lp);
    if (lp->val[0] < 0.0) lp_flip(lp);
    const int cap = 100 * m + 1000;
    int refactored = 0, stall = 0, bland = 0;
    double tbest = -1.0;
    for (int iter = 0;; ++iter) {
        const double t = lp->val[0];
        if (!approximatelyLE(t, radius, DIST_EPS)) { if (iter == 0) lp->quick++; return NODE_PRUNE; }
        /* degenerate stalling: after STALL_MAX steps without progress switch to lowest-index (Bland) choices */
        if (t > tbest * (1.0 + 1.0e-12)) { tbest = t; stall = 0; } else if (++stall >= 2 * STALL_MAX) bland = 1;
        const int dantzig = stall >= STALL_MAX; /* first resort against stalling: leave by largest violation */
        const double tol = EPS_OPT * (cmax + t);
        /* Pricing: among the basic rows outside the box (|r_z| > t + tol) the one with the steepest dual edge leaves,
         * score = (|r_z| - t)^2 / |H[p] - se H[0]|^2.  The columns of N are orthonormal, so that norm IS the norm of the
         * dual ray sum_j (n_j . rho)^2 -- exact dual steepest-edge weights for q flops per row, no recurrences.
         * Ties: lowest position.  In Bland mode: the violated row with the lowest row index. */
        int pe = -1; double best = -1.0; int br = -1;
        for (int p = 1; p < qd; ++p) {
            const double a = fabs(lp->val[p]);
            if (!(a > t + tol)) continue;
            if (bland) { if (br < 0 || lp->zrow[p] < br) { br = lp->zrow[p]; pe = p; } continue; }
            const double se_p = lp->val[p] > 0.0 ? 1.0 : -1.0;
            const double *Hp = lp->H + (size_t)p * lp->ld, *H0 = lp->H;
            double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0; /* dot4 order */
            int e = 0;
            for (; e + 3 < qd; e += 4) {
                const double x0 = Hp[e] - se_p * H0[e], x1 = Hp[e + 1] - se_p * H0[e + 1];
                const double x2 = Hp[e + 2] - se_p * H0[e + 2], x3 = Hp[e + 3] - se_p * H0[e + 3];
                a0 = fma(x0, x0, a0); a1 = fma(x1, x1, a1); a2 = fma(x2, x2, a2); a3 = fma(x3, x3, a3);
            }
            if (e < qd) { const double x = Hp[e] - se_p * H0[e]; a0 = fma(x, x, a0); ++e; }
            if (e < qd) { const double x = Hp[e] - se_p * H0[e]; a1 = fma(x, x, a1); ++e; }
            if (e < qd) { const double x = Hp[e] - se_p * H0[e]; a2 = fma(x, x, a2); }
            const double nr = (a0 + a1) + (a2 + a3);
            const double viol = a - t;
            const double score = dantzig ? viol : viol * viol / nr;
            if (score > best) { best = score; pe = p; }
        }
        if (pe < 0) {
            /* optimal by the books: verify the vertex against the data before trusting it */
            const int nv = qd < m ? qd + 1 : qd;
            lp_image(lp, nv);
            double err = 0.0;
            for (int i = 0; i < qd; ++i) { const double e = fabs(lp->v[i] - lp->chat[i]); if (e > err) err = e; }
            if (err <= 1.0e-8 * ((double)m * cmax + t)) { *d_out = t; return NODE_EXPLORE; }
            if (refactored) return NODE_FAIL;
            refactored = 1;
            if (lp_refactor(lp)) return NODE_FAIL;
            lp_values(lp);
            if (lp->val[0] < 0.0) lp_flip(lp);
            continue;
        }
        if (iter >= cap) return NODE_FAIL;
        const int se = lp->val[pe] > 0.0 ? 1 : -1;
        const double nrm1 = lp_alpha_beta(lp, pe);
        if (lp_repair(lp)) { /* the vertex moved a little: price again */
            lp_values(lp);
            if (lp->val[0] < 0.0) lp_flip(lp);
            continue;
        }
        if (fabs(nrm1 - 1.0) > 1.0e-5) { /* the dual vector has drifted */
            if (refactored) return NODE_FAIL;
            refactored = 1;
            if (lp_refactor(lp)) return NODE_FAIL;
            lp_values(lp);
            if (lp->val[0] < 0.0) lp_flip(lp);
            continue;
        }
        const int l = lp_ratio(lp, se, pe, bland);
        if (l < 0) return NODE_FAIL;
        if (lp_replace(lp, pe, l, se)) return NODE_FAIL;
        lp_values(lp);
    }
}

/* lp_solve preceded by the removal of a fresh coordinate's artificial column (art = its sign, 0 = none) */
static int lp_solve_art(LP *lp, double radius, double cmax, double *d_out, int art) {
    if (art && lp_drop_artificial(lp, art)) return NODE_FAIL;
    return lp_solve(lp, radius, cmax, d_out);
}
/* Last resort after a failed warm start: rebuild the vertex of LP(m - qd) coordinate by coordinate from nothing, each
 * intermediate problem solved to optimality.  lp->chat must hold all qd entries. */
static int lp_cold_solve(LP *lp, int qd, double radius, double cmax, double *d_out) {
    if (lp_init1(lp)) return NODE_FAIL;
    int status = NODE_FAIL;
    for (;;) {
        int art = 0; double rad = radius;
        if (lp->qd < qd) {
            lp_extend(lp);
            lp_values(lp);
            art = lp->val[lp->qd - 1] > 0.0 ? 1 : -1;
            if (lp->qd < qd) rad = INFINITY;
        }
        status = lp_solve_art(lp, rad, cmax, d_out, art);
        if (status == NODE_FAIL) return NODE_FAIL;
        if (lp->qd == qd) return status;
        if (status != NODE_EXPLORE) return NODE_FAIL;
    }
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
    double cmax = 0.0;
    for (int j = 0; j < m; ++j) if (fabs(c[j]) > cmax) cmax = fabs(c[j]);
    if (k == 0) { *dist = cmax; free(c); return 0; }
    Geo g;
    if (geo_build(&g, B, p, n, m)) { free(c); geo_free(&g); return -1; }
    LP *lp = lp_new(&g);
    const int qd = m - k;
    double d = 0.0;
    for (int j = 0; j < m; ++j) lp->r[j] = 0.0;
    if (qd > 0) {
        for (int i = 0; i < qd; ++i) lp->chat[i] = dot4(g.Qrev + i, m, c, 1, m);
        if (lp_init1(lp)) rc = -1;
        for (int i = 1; i < qd && rc == 0; ++i) {
            lp_extend(lp);
            lp_values(lp);
            if (lp_drop_artificial(lp, lp->val[i] > 0.0 ? 1 : -1)) rc = -1;
        }
        if (rc == 0 && lp_solve(lp, INFINITY, cmax, &d) != NODE_EXPLORE) rc = -1;
        if (rc == 0) for (int j = 0; j < m; ++j) lp->r[j] = lp->pos[j] ? lp->val[lp->pos[j]] : (double)lp->sg[j] * lp->val[0];
    }
    if (rc == 0) { /* x = R_k^-1 Q_k^T (c - r) */
        *dist = d;
        double *y = (double *)calloc(k, sizeof(double));
        for (int i = 0; i < k; ++i) {
            double s = 0.0;
            for (int j = 0; j < m; ++j) s = fma(g.Q[(size_t)i * m + j], c[j] - lp->r[j], s);
            y[i] = s;
        }
        for (int i = k - 1; i >= 0; --i) {
            double s = y[i];
            for (int t = i + 1; t < k; ++t) s = fma(-g.Rm[(size_t)i * n + t], x[t], s);
            x[i] = s / g.Rm[(size_t)i * n + i];
        }
        free(y);
    }
    lp_free(lp); geo_free(&g); free(c);
    return rc;
}

/* ---- closest() ----
 * lat: n x m int64 row-major (rows = basis vectors), con: m x 2.  sol: n (uhat).  hist: n+1 counters, hist[kk] =
 * number of layer-distance solves with kk free coefficients.  stats[0..3] = pivots, refactors, quick prunes, cold
 * starts.  max_nodes = 0 means unlimited, otherwise stop with ORACLE_TRUNCATED after that many solves.
 *
 * Warm-start policy ("side chains"): at every level the siblings below the real centre xc and
 * the siblings above it form two chains.  The first node of a chain starts from the parent's optimal vertex extended
 * by the level's coordinate; every later node of the chain starts from the final vertex of the previous node on the
 * same side.  The vertex a node starts from is therefore a function of its path alone. */
typedef struct { double *H; int *zrow; signed char *sg; int init; } Chain;

static void chain_store(Chain *ch, const LP *lp) {
    memcpy(ch->H, lp->H, (size_t)lp->qd * lp->ld * sizeof(double));
    memcpy(ch->zrow, lp->zrow, lp->qd * sizeof(int));
    memcpy(ch->sg, lp->sg, lp->m);
}
static void chain_load(LP *lp, const Chain *ch, int qd) {
    lp->qd = qd;
    memcpy(lp->H, ch->H, (size_t)qd * lp->ld * sizeof(double));
    memcpy(lp->zrow, ch->zrow, qd * sizeof(int));
    memcpy(lp->sg, ch->sg, lp->m);
    for (int j = 0; j < lp->m; ++j) lp->pos[j] = 0;
    for (int p = 1; p < qd; ++p) if (lp->zrow[p] >= 0) lp->pos[lp->zrow[p]] = p;
}

/* Sound mode (bltcu_opts.sound): the node LPs prune against radius * (1 + margin); 0 = the reference's plain rule.
 * The reference's own sound mode asks Yices about every LP prune (closest.cc:258-267); Yices is not available here. */
static double g_sound_margin = 0.0;
static uint64_t g_inband = 0;
void oracle_set_sound(double margin) { g_sound_margin = margin; }
uint64_t oracle_inband(void) { return g_inband; }

#ifdef ORACLE_CENSUS
/* Study aid (profiles/census_study.py compiles its own copy with -DORACLE_CENSUS; the checker is built without it):
 * sizes of the subtrees below the explored nodes of one level, and how many of their nodes lie within `near`
 * levels of the root. */
static int g_census_level = 0, g_census_near = 2, g_census_cap = 0, g_census_count = 0, g_census_cur = -1;
static uint64_t *g_census_size = 0, *g_census_probe = 0;
void oracle_set_census(int level, int near, uint64_t *size, uint64_t *probe, int cap) {
    g_census_level = level; g_census_near = near; g_census_size = size; g_census_probe = probe; g_census_cap = cap;
    g_census_count = 0; g_census_cur = -1;
}
int oracle_census_count(void) { return g_census_count; }
#endif

int oracle_closest(const int64_t *lat, const int64_t *con, int n, int m, int64_t *sol, uint64_t *niter_out,
                   uint64_t *nprune_out, uint64_t *hist, uint64_t *stats, uint64_t max_nodes) {
    if (n <= 0 || m <= 0 || m < n) return BLT_STATUS_INPUT_ERROR; /* closest.cc:173-177 */
    double *B = (double *)malloc((size_t)n * m * sizeof(double));
    for (size_t i = 0; i < (size_t)n * m; ++i) B[i] = (double)lat[i];
    double *p = (double *)malloc((size_t)m * sizeof(double));
    for (int j = 0; j < m; ++j) p[j] = (double)(con[2 * j + 1] + con[2 * j]) / 2.0; /* closest.cc:180-185 */
    const double radius = (double)(con[1] - con[0]) / 2.0;                          /* closest.cc:193-194: row 0 only */
    const double radius_lp = g_sound_margin > 0.0 ? radius * (1.0 + g_sound_margin) : radius;
    g_inband = 0;

    Geo g;
    uint64_t counter = 0, nprune = 0, nodes = 0;
    int rc = BLT_STATUS_OK;
    if (hist) memset(hist, 0, (size_t)(n + 1) * sizeof(uint64_t));
    if (geo_build(&g, B, p, n, m)) { geo_free(&g); free(B); free(p); return BLT_STATUS_LP_FAIL; }
    LP *lp = lp_new(&g);
    const int ld = lp->ld;
    int64_t *u = (int64_t *)calloc(n, sizeof(int64_t)), *uhat = (int64_t *)calloc(n, sizeof(int64_t));
    int64_t *step = (int64_t *)calloc(n, sizeof(int64_t));
    unsigned char *pruned = (unsigned char *)calloc(n, 1);
    double *c = (double *)calloc(m, sizeof(double));
    /* per level k (state k solves LP(k-1), qd = m-k+1): two chains, the parent's right-hand side in both coordinate
     * systems, the real centre, the base of this level's own coordinate, the side we descended through */
    Chain (*ch)[2] = (Chain (*)[2])calloc(n + 3, sizeof(Chain[2]));
    double **cparL = (double **)calloc(n + 3, sizeof(double *)), **chatL = (double **)calloc(n + 3, sizeof(double *));
    double *xc = (double *)calloc(n + 3, sizeof(double)), *gbase = (double *)calloc(n + 3, sizeof(double));
    int *curside = (int *)calloc(n + 3, sizeof(int));
    for (int k = 0; k <= n + 2; ++k) {
        cparL[k] = (double *)calloc(m, sizeof(double)); chatL[k] = (double *)calloc(ld, sizeof(double));
        for (int s = 0; s < 2; ++s) {
            ch[k][s].H = (double *)calloc((size_t)ld * ld, sizeof(double));
            ch[k][s].zrow = (int *)calloc(ld, sizeof(int)); ch[k][s].sg = (signed char *)calloc(m, 1);
        }
    }
    int k = n;
    double xlast = 0.0, d = 0.0;

    /* initial solve LP(n): closest.cc:206; only x(n-1) is used */
    {
        double cmax = 0.0;
        for (int j = 0; j < m; ++j) if (fabs(p[j]) > cmax) cmax = fabs(p[j]);
        if (hist) hist[n]++;
        nodes++;
        const int qd = m - n;
        double vq = 0.0;
        if (qd > 0) {
            for (int i = 0; i < qd; ++i) lp->chat[i] = g.QTp[m - 1 - i];
            int bad = lp_init1(lp);
            for (int i = 1; i < qd && !bad; ++i) {
                lp_extend(lp);
                lp_values(lp);
                bad = lp_drop_artificial(lp, lp->val[i] > 0.0 ? 1 : -1);
            }
            if (bad || lp_solve(lp, INFINITY, cmax, &d) != NODE_EXPLORE) { rc = BLT_STATUS_LP_FAIL; goto done; }
            vq = lp->v[qd];
            chain_store(&ch[n + 1][0], lp);
            memcpy(chatL[n], lp->chat, qd * sizeof(double));
        }
        curside[n + 1] = 0;
        gbase[n] = g.QTp[n - 1];
        xlast = (gbase[n] - vq) / g.Rm[(size_t)(n - 1) * n + (n - 1)];
    }
    u[n - 1] = (int64_t)round(xlast);          /* set_index, closest.cc:74-82 */
    step[n - 1] = sgn(xlast - (double)u[n - 1]);
    xc[n] = xlast;
    memcpy(cparL[n], p, (size_t)m * sizeof(double));

    for (;;) {
        if (max_nodes && nodes >= max_nodes) { rc = ORACLE_TRUNCATED; goto done; }
        /* right-hand side of this node: c = cpar - u[k-1] * B[k-1]  (D2) */
        double cmax = 0.0;
        {
            const double mu = -(double)u[k - 1];
            const double *Bk = B + (size_t)(k - 1) * m, *cp = cparL[k];
            for (int j = 0; j < m; ++j) { c[j] = fma(mu, Bk[j], cp[j]); if (fabs(c[j]) > cmax) cmax = fabs(c[j]); }
        }
        int status, side = 0;
        nodes++;
        if (hist) hist[k - 1]++;
#ifdef ORACLE_CENSUS
        if (g_census_size) {
            if (k >= g_census_level) g_census_cur = -1;
            else if (g_census_cur >= 0) { g_census_size[g_census_cur]++; if (k >= g_census_level - g_census_near) g_census_probe[g_census_cur]++; }
        }
#endif
        if (k == 1) { /* LP(0): plain max-abs */
            d = cmax;
            status = approximatelyLE(d, radius, DIST_EPS) ? NODE_EXPLORE : NODE_PRUNE;
        } else {
            const int kk = k - 1, qd = m - kk;
            const int dir = sgn(xc[k] - (double)u[k - 1]);
            side = dir > 0 ? 0 : 1;
            const double clast = fma(-(double)u[k - 1], g.Rm[(size_t)kk * n + kk], gbase[k]);
            int art = 0;
            if (!ch[k][side].init) {
                if (qd == 1) {
                    if (lp_init1(lp)) { rc = BLT_STATUS_LP_FAIL; goto done; }
                } else {
                    chain_load(lp, &ch[k + 1][curside[k + 1]], qd - 1);
                    lp_extend(lp);
                    art = dir;
                }
                ch[k][side].init = 1;
            } else chain_load(lp, &ch[k][side], qd);
            memcpy(lp->chat, chatL[k], (qd - 1) * sizeof(double));
            lp->chat[qd - 1] = clast;
            status = lp_solve_art(lp, radius_lp, cmax, &d, art);
            if (status == NODE_FAIL) status = lp_cold_solve(lp, qd, radius_lp, cmax, &d); /* warm start went bad: from scratch */
            if (status == NODE_FAIL) { rc = BLT_STATUS_LP_FAIL; goto done; }
            chain_store(&ch[k][side], lp);
            if (status == NODE_EXPLORE) { /* real centre of the next level: x_{kk-1} of the optimum */
                if (radius_lp != radius && !approximatelyLE(d, radius, DIST_EPS)) g_inband++;
                /* next coordinate of the rotated right-hand side: the part fixed by the levels above (32-lane order),
                 * then this level's own coefficient */
                double part[32];
                for (int l = 0; l < 32; ++l) part[l] = 0.0;
                for (int i = kk + 1; i < n; ++i) part[(i - kk - 1) & 31] = fma(-(double)u[i], g.Rm[(size_t)(kk - 1) * n + i], part[(i - kk - 1) & 31]);
                const double g2 = g.QTp[kk - 1] + butterfly32(part);
                const double gn = fma(-(double)u[kk], g.Rm[(size_t)(kk - 1) * n + kk], g2);
                gbase[k - 1] = gn;
                xlast = (gn - lp->v[qd]) / g.Rm[(size_t)(kk - 1) * n + (kk - 1)];
            }
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
#ifdef ORACLE_CENSUS
        if (g_census_size && k == g_census_level && g_census_count < g_census_cap) {
            g_census_cur = g_census_count++; g_census_size[g_census_cur] = 0; g_census_probe[g_census_cur] = 0;
        }
#endif
        if (k > 1) { /* descend, closest.cc:88-95 */
            curside[k] = side;
            memcpy(cparL[k - 1], c, (size_t)m * sizeof(double));
            memcpy(chatL[k - 1], lp->chat, (size_t)(m - k + 1) * sizeof(double));
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
    for (int kk = 0; kk <= n + 2; ++kk) {
        free(cparL[kk]); free(chatL[kk]);
        for (int s = 0; s < 2; ++s) { free(ch[kk][s].H); free(ch[kk][s].zrow); free(ch[kk][s].sg); }
    }
    free(ch); free(cparL); free(chatL); free(xc); free(gbase); free(curside);
    free(u); free(uhat); free(step); free(pruned); free(c); free(B); free(p);
    lp_free(lp); geo_free(&g);
    return rc;
}
