// lp_reg.cuh -- the node LP with the basis inverse held in REGISTERS (included by search.cu).
//
// Same algorithm, same operations in the same order as the lp_* functions of search.cu (and of oracle/blt_oracle.c):
// every result is bit-identical.  What changes is where H = A^-1 lives while a warp works on a node:
//   * lane p holds row p of H in QP registers (QP = 24, the largest basis of this path; unused columns and rows are
//     exact zeros, which leave every dot4 sum untouched);
//   * row 0 (the dual vector y) and the leaving row are mirrored in small shared-memory vectors for the steps that
//     need them on every lane (pricing, alpha/beta, the rank-one update);
//   * the OTHER side of the sibling loop (or the extended parent vertex a side starts from) is parked in one
//     shared-memory matrix per warp; going to the other side swaps registers and that matrix.
// All loops over the columns of H are fully unrolled: no index arithmetic, no loop control, 16-byte broadcasts from
// shared memory.  The pivot loop is straight-line code that every warp of an SM runs through again and again, so it
// is written to appear ONCE in the instruction stream and to stay small (instruction caches), and every shared-memory
// array is addressed as a compile-time offset from the warp's base (struct W below) so that the accesses compile to
// LDS/STS with immediate offsets.  Bases above 24 rows, cold restarts and refactorisations use the generic code of
// search.cu through a WS built on the spot (those functions are not inlined).
#pragma once

// What the register engine knows about its warp: the base of the shared-memory block (everything else is an
// immediate offset, SL<MC> in search.cu), Q in shared memory, the warp's global buffers, and a few scalars.
template <int MC>
struct W {
    typedef SL<MC> L;
    double *b;
    const double *Q;          // first QS_MAX + 2 columns of Qrev, row stride LDQ_C
    double *Wpar, *Wg_a, *Wg_b, *tmpA;  // global: parent vertex, generic-path matrices, refactorisation scratch
    int LDG, lane, n, m;
    __device__ __forceinline__ double *park() const { return b; }
    __device__ __forceinline__ double *chat() const { return b + L::o_vec; }
    __device__ __forceinline__ double *val() const { return b + L::o_vec + L::V; }
    __device__ __forceinline__ double *a() const { return b + L::o_vec + 3 * L::V; }
    __device__ __forceinline__ double *v() const { return b + L::o_vec + 4 * L::V; }
    __device__ __forceinline__ double *r() const { return b + L::o_vec + 5 * L::V; }
    __device__ __forceinline__ double *c() const { return b + L::o_vec + 6 * L::V; }
    __device__ __forceinline__ double *cpar() const { return b + L::o_vec + 7 * L::V; }
    __device__ __forceinline__ double *brow() const { return b + L::o_vec + 8 * L::V; }
    __device__ __forceinline__ double *ybuf() const { return b + L::o_sm; }
    __device__ __forceinline__ double *rbuf() const { return b + L::o_sm + L::SV; }
    __device__ __forceinline__ double *vbuf() const { return b + L::o_sm + 2 * L::SV; }
    __device__ __forceinline__ long long *u() const { return (long long *)(b + L::o_u); }
    __device__ __forceinline__ unsigned long long *ctr() const { return (unsigned long long *)(b + L::o_ctr); }
    __device__ __forceinline__ int *zrow(int side) const { return (int *)(b + L::o_int) + side * L::V; }
    __device__ __forceinline__ int *pzrow() const { return (int *)(b + L::o_int) + 2 * L::V; }
    __device__ __forceinline__ unsigned char *pos(int side) const { return (unsigned char *)(b + L::o_byte) + side * L::NB; }
    __device__ __forceinline__ signed char *sg(int side) const { return (signed char *)(b + L::o_byte) + (2 + side) * L::NB; }
    __device__ __forceinline__ signed char *psg() const { return (signed char *)(b + L::o_byte) + 4 * L::NB; }
    __device__ __forceinline__ ord_t *ord() const { return (ord_t *)((unsigned char *)(b + L::o_byte) + 5 * L::NB); }
    __device__ __forceinline__ void count(int k) const { if (lane == 0) ctr()[k]++; }
};
// the same block as a WS, for the generic functions of search.cu (rare paths; the WS lives on the stack of the caller)
template <int MC>
__device__ __forceinline__ void to_ws(const W<MC> &w, WS &s) {
    (void)carve_fixed<MC>(s, w.b);
    s.Wpar = w.Wpar; s.Wg_a = w.Wg_a; s.Wg_b = w.Wg_b; s.tmpA = w.tmpA;
    s.LDS = LDS_C; s.LDG = w.LDG; s.QS = QS_MAX;
    s.Qsm = w.Q; s.LDQsm = LDQ_C; s.Qs = w.Q; s.LDQ = LDQ_C;
    s.n = w.n; s.m = w.m; s.qd = 0; s.lane = w.lane; s.repairs = 0; s.failcode = 0;
    s.W = s.park; s.LD = LDS_C;
    s.zrow = s.zrow_a; s.pos = s.pos_a; s.sg = s.sg_a;
}

// a / b with operands the compiler must take as they are: without the barrier it divides the unguarded values and
// selects afterwards, and a zero numerator or denominator sends the whole warp through the slow division routine
__device__ __forceinline__ double safe_div(double a, double b) {
    asm volatile("" : "+d"(a), "+d"(b));
    return a / b;
}

// dot4 of this lane's row with a shared-memory vector (16-byte aligned, QP entries readable)
template <int QP>
__device__ __forceinline__ double rdot(const double (&h)[QP], const double *x) {
    double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;
#pragma unroll
    for (int e = 0; e < QP; e += 4) {
        const double2 x0 = *(const double2 *)(x + e), x1 = *(const double2 *)(x + e + 2);
        a0 = fma(h[e], x0.x, a0); a1 = fma(h[e + 1], x0.y, a1); a2 = fma(h[e + 2], x1.x, a2); a3 = fma(h[e + 3], x1.y, a3);
    }
    return (a0 + a1) + (a2 + a3);
}
template <int QP>
__device__ __forceinline__ void row_to_smem(const double (&h)[QP], double *dst) {
#pragma unroll
    for (int e = 0; e < QP; e += 2) *(double2 *)(dst + e) = make_double2(h[e], h[e + 1]);
}
template <int QP>
__device__ __forceinline__ void row_from_smem(double (&h)[QP], const double *src) {
#pragma unroll
    for (int e = 0; e < QP; e += 2) { const double2 x = *(const double2 *)(src + e); h[e] = x.x; h[e + 1] = x.y; }
}
// registers <- parked matrix (rows and columns beyond q become zeros)
template <int QP>
__device__ __forceinline__ void park_load(double (&h)[QP], const double *park, int q, int lane) {
    const double *src = park + (size_t)(lane < QP ? lane : 0) * LDS_C;
#pragma unroll
    for (int e = 0; e < QP; e += 2) {
        const double2 x = *(const double2 *)(src + e);
        h[e] = (lane < q && e < q) ? x.x : 0.0;
        h[e + 1] = (lane < q && e + 1 < q) ? x.y : 0.0;
    }
}
// registers <-> parked matrix (both q x q)
template <int QP>
__device__ __forceinline__ void park_swap(double (&h)[QP], double *park, int q, int lane) {
    if (lane < QP) {
        double *row = park + (size_t)lane * LDS_C;
#pragma unroll
        for (int e = 0; e < QP; e += 2) {
            const double2 x = *(const double2 *)(row + e);
            *(double2 *)(row + e) = make_double2(h[e], h[e + 1]);
            h[e] = (lane < q && e < q) ? x.x : 0.0;
            h[e + 1] = (lane < q && e + 1 < q) ? x.y : 0.0;
        }
    }
}
template <int QP, int MC>
__device__ __forceinline__ void publish_y(const W<MC> &w, const double (&h)[QP]) {
    if (w.lane == 0) row_to_smem<QP>(h, w.ybuf());
    __syncwarp();
}
// lp_flip: t came out negative, the whole vertex changes sign
template <int QP, int MC>
__device__ __forceinline__ void r_flip(const W<MC> &w, int side, double (&h)[QP], double &myval) {
    if (w.lane == 0) {
#pragma unroll
        for (int e = 0; e < QP; ++e) h[e] = -h[e];
        myval = -myval;
        w.val()[0] = myval;
    }
    if (w.lane < QP) w.ybuf()[w.lane] = -w.ybuf()[w.lane];
    signed char *sg = w.sg(side);
#pragma unroll 1
    for (int j = w.lane; j < w.m; j += 32) sg[j] = (signed char)-sg[j];
    __syncwarp();
}
// lp_alpha_beta: alpha_j = n_j . y, beta_j = n_j . H[pe] of this lane's rows j = lane + 32 r, kept in registers.
// y and the leaving row are read from their shared-memory mirrors (ybuf, rbuf).  Returns |N y|_1.
template <int QP, int MC>
__device__ __forceinline__ double r_alpha_beta(const W<MC> &w, int side, double (&al)[MC / 32], double (&be)[MC / 32], double &bmax,
                                               bool &need_repair) {
    const int m = w.m;
    const unsigned char *pos = w.pos(side);
    const signed char *sg = w.sg(side);
    const double *yb = w.ybuf(), *rb = w.rbuf();
    double part = 0.0, bm = 0.0;
    bool bad = false;
#pragma unroll
    for (int r = 0; r < MC / 32; ++r) {
        const int j = w.lane + 32 * r;
        double a = 0.0, b = 0.0;
        if (j < m && !pos[j]) {
            const double *nj = w.Q + (size_t)j * LDQ_C;
            double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0, b0 = 0.0, b1 = 0.0, b2 = 0.0, b3 = 0.0;
#pragma unroll
            for (int e = 0; e < QP; e += 4) {
                const double2 x0 = *(const double2 *)(nj + e), x1 = *(const double2 *)(nj + e + 2);
                const double2 y0 = *(const double2 *)(yb + e), y1 = *(const double2 *)(yb + e + 2);
                const double2 z0 = *(const double2 *)(rb + e), z1 = *(const double2 *)(rb + e + 2);
                a0 = fma(x0.x, y0.x, a0); b0 = fma(x0.x, z0.x, b0);
                a1 = fma(x0.y, y0.y, a1); b1 = fma(x0.y, z0.y, b1);
                a2 = fma(x1.x, y1.x, a2); b2 = fma(x1.x, z1.x, b2);
                a3 = fma(x1.y, y1.y, a3); b3 = fma(x1.y, z1.y, b3);
            }
            a = (a0 + a1) + (a2 + a3);
            b = (b0 + b1) + (b2 + b3);
            part += fabs(a);
            const double fb = fabs(b);
            bm = fb > bm ? fb : bm;
            bad = bad || ((sg[j] > 0 ? a : -a) < -HARRIS_DELTA);
        }
        al[r] = a; be[r] = b;
    }
    bmax = bm;
    need_repair = __any_sync(FULLMASK, bad);
    return butterfly(part);
}
// lp_repair on the register-resident H (rare): rows whose sg*alpha fell below -HARRIS_DELTA go to the other bound.
template <int QP, int MC>
__device__ __forceinline__ int r_repair(const W<MC> &w, int side, int qd, double (&h)[QP], const double (&al)[MC / 32], bool need) {
    const int m = w.m, lane = w.lane;
    if (!need) return 0;
    const unsigned char *pos = w.pos(side);
    signed char *sg = w.sg(side);
    double *rr = w.r(), *aa = w.a();
    __syncwarp();
#pragma unroll
    for (int r = 0; r < MC / 32; ++r) {
        const int j = lane + 32 * r;
        if (j < m) {
            double rj = 0.0;
            if (!pos[j] && (double)sg[j] * al[r] < -HARRIS_DELTA) rj = -2.0 * (double)sg[j];
            rr[j] = rj;
        }
    }
    __syncwarp();
    if (lane < QP) {  // Delta = N^T r (entries beyond qd stay zero)
        const double *Qc = w.Q + lane;
        aa[lane] = lane < qd ? dot4(m, [&](int e) { return rr[e]; }, [&](int e) { return Qc[(size_t)e * LDQ_C]; }) : 0.0;
    }
    __syncwarp();
    const double wi = rdot<QP>(h, aa);
    const double den = 1.0 + __shfl_sync(FULLMASK, wi, 0);
    {   // H[i][c] -= (w_i / den) y[c] with the OLD row 0 (still in ybuf)
        const double f = wi / den;
        const double *yb = w.ybuf();
#pragma unroll
        for (int e = 0; e < QP; e += 2) {
            const double2 y = *(const double2 *)(yb + e);
            h[e] = fma(-f, y.x, h[e]); h[e + 1] = fma(-f, y.y, h[e + 1]);
        }
    }
    __syncwarp();
#pragma unroll 1
    for (int j = lane; j < m; j += 32) if (rr[j] != 0.0) sg[j] = (signed char)-sg[j];
    publish_y<QP>(w, h);
    return 1;
}
// lp_ratio with alpha/beta in registers; rbuf holds the leaving row (for the scale)
template <int QP, int MC>
__device__ __forceinline__ int r_ratio(const W<MC> &w, int side, const double (&al)[MC / 32], const double (&be)[MC / 32], int se, int pe,
                                       bool bland, double bmax) {
    constexpr int RPL = MC / 32;
    const int m = w.m, lane = w.lane;
    const unsigned char *pos = w.pos(side);
    const signed char *sg = w.sg(side);
    const bool artificial = w.zrow(side)[pe] < 0;
    double scale = bmax;
    if (lane < QP) { const double a = fabs(w.rbuf()[lane]); scale = a > scale ? a : scale; }
    scale = warp_max(scale);
    const double tiny = PIV_TOL * scale;
    int best = -1; double bd = 0.0;
    // per-row quantities that do not depend on the tier
    double den_[RPL], sa_[RPL];
    bool act_[RPL];
#pragma unroll
    for (int r = 0; r < RPL; ++r) {
        const int j = lane + 32 * r;
        act_[r] = j < m && !pos[j];
        const bool sp = act_[r] && sg[j] > 0;
        den_[r] = ((se > 0) == sp) ? -be[r] : be[r];  // -(se sg_j) beta_j
        sa_[r] = sp ? al[r] : -al[r];                 // sg_j alpha_j
    }
#pragma unroll 1
    for (int tier = 0; tier < 2 && best < 0; ++tier) {
        const double dmin = tier == 0 ? PIV_REL * scale : 0.0;
        // Every lane divides, rows out of the running with operands 1/1 and zero numerators set aside (0/den = +0):
        // a quotient that is zero, infinite or not a number sends the whole warp through the slow division routine.
        double tmax = INFINITY;
#pragma unroll
        for (int r = 0; r < RPL; ++r) {
            const bool ok = act_[r] && den_[r] > dmin;
            const double num = sa_[r] + HARRIS_DELTA;
            const bool pp = ok && num > 0.0;
            const double q = safe_div(pp ? num : 1.0, pp ? den_[r] : 1.0);
            const double rt = pp ? q : 0.0;
            if (ok && rt < tmax) tmax = rt;
        }
        tmax = warp_min(tmax);
        bd = 0.0; best = -1;
#pragma unroll
        for (int r = 0; r < RPL; ++r) {
            const bool ok = act_[r] && den_[r] > dmin;
            const bool pp = ok && sa_[r] > 0.0;
            const double q = safe_div(pp ? sa_[r] : 1.0, pp ? den_[r] : 1.0);
            const double rt = pp ? q : 0.0;
            if (ok && rt <= tmax && (bland ? best < 0 : den_[r] > bd)) { bd = den_[r]; best = lane + 32 * r; }
        }
        if (bland) { double key = best >= 0 ? -(double)best : 0.0; warp_argmax(key, best); bd = __shfl_sync(FULLMASK, bd, 0); }
        else warp_argmax_nn(bd, best);
    }
    if (artificial) return (best >= 0 && bd > 1.0e-3 * tiny) ? best : -1;
    if (best >= 0) return best;
    double ba = -1.0;
#pragma unroll
    for (int r = 0; r < RPL; ++r) {
        if (!act_[r] || fabs(be[r]) > tiny) continue;
        if (fabs(al[r]) > ba) { ba = fabs(al[r]); best = lane + 32 * r; }
    }
    warp_argmax_nn(ba, best);
    return best;
}
// lp_replace: basic position pe leaves with sign se, row l enters.  On entry rbuf holds row pe and ybuf row 0; on exit
// ybuf holds the new row 0.  The leaving row is transformed element-wise across the lanes in shared memory (one
// element per lane) instead of by its owner alone; the owner reloads it where it needs it in registers.
template <int QP, int MC>
__device__ __forceinline__ int r_replace(const W<MC> &w, int side, double (&h)[QP], int pe, int l, int se) {
    const int lane = w.lane;
    int *zrow = w.zrow(side);
    unsigned char *pos = w.pos(side);
    signed char *sg = w.sg(side);
    double *rb = w.rbuf(), *vb = w.vbuf();
    const double *yb = w.ybuf();
    const int e = zrow[pe];
    double adj = 0.0;
    if (lane < QP) {  // s += se * n_e
        adj = rb[lane];
        if (e >= 0) adj = fma(-(double)se, yb[lane], adj);   // == Hp - se*y: the product is exact
        rb[lane] = adj;
    }
    __syncwarp();
    if (lane == pe) row_from_smem<QP>(h, rb);
    const double wd = rdot<QP>(h, w.Q + (size_t)l * LDQ_C);
    const double wp = __shfl_sync(FULLMASK, wd, pe);
    if (wp == 0.0) return -1;
    double vc = 0.0;
    if (lane < QP) {  // a zero numerator would take the slow division path: such lanes divide 1 / wp instead
        const bool nz = adj != 0.0;
        const double q = safe_div(nz ? adj : 1.0, wp);
        vc = nz ? q : (wp > 0.0 ? adj : -adj);   // adj is +-0 here: 0 / wp has the sign of adj * sign(wp)
        vb[lane] = vc;
    }
    __syncwarp();
    if (lane != pe) {
        const double mw = -wd;
#pragma unroll
        for (int c = 0; c < QP; c += 2) {
            const double2 v = *(const double2 *)(vb + c);
            h[c] = fma(mw, v.x, h[c]); h[c + 1] = fma(mw, v.y, h[c + 1]);
        }
    }
    publish_y<QP>(w, h);
    if (lane < QP) rb[lane] = fma((double)sg[l], yb[lane], vc);   // s -= sg_l * n_l: v + sl*y, product exact
    __syncwarp();
    if (lane == pe) row_from_smem<QP>(h, rb);
    if (lane == 0) {
        if (e >= 0) { sg[e] = (signed char)se; pos[e] = 0; }
        zrow[pe] = l; pos[l] = (unsigned char)pe;
    }
    __syncwarp();
    w.count(C_PIVOTS);
    return 0;
}
// rows of H <- a matrix in global memory with leading dimension ld (generic-path results)
template <int QP>
__device__ __forceinline__ void rows_from_global_ld(double (&h)[QP], const double *src, int ld, int q, int lane) {
    const double *row = src + (size_t)(lane < q ? lane : 0) * ld;
#pragma unroll
    for (int e = 0; e < QP; ++e) h[e] = (lane < q && e < q) ? row[e < q ? e : 0] : 0.0;
}
// refactorisation (rare): the generic routine rebuilds H from Z and the signs in the warp's global buffer
template <int QP, int MC>
__device__ __forceinline__ int r_refactor(const W<MC> &w, int side, int qd, double (&h)[QP]) {
    w.count(C_REFAC);
    const int rc = lp_refactor_impl(w.Wg_a, w.LDG, w.tmpA, w.Q, LDQ_C, qd, w.m, w.pos(side), w.sg(side), w.zrow(side), w.lane);
    rows_from_global_ld<QP>(h, w.Wg_a, w.LDG, qd, w.lane);
    publish_y<QP>(w, h);
    return rc;
}
// lp_image: the full residual vector and its image, v[i] = sum_j Q[j][i] r_j for i < nv
template <int MC>
__device__ __forceinline__ void r_image(const W<MC> &w, int side, int nv) {
    const int m = w.m, lane = w.lane;
    const unsigned char *pos = w.pos(side);
    const signed char *sg = w.sg(side);
    const double *val = w.val();
    double *rr = w.r();
    const double t = val[0];
#pragma unroll 1
    for (int j = lane; j < m; j += 32) rr[j] = pos[j] ? val[pos[j]] : (double)sg[j] * t;
    __syncwarp();
    if (lane < nv) {  // nv <= QS_MAX + 1 <= 32: one lane per coordinate
        const double *Qi = w.Q + lane;
        w.v()[lane] = dot4(m, [&](int e) { return Qi[(size_t)e * LDQ_C]; }, [&](int e) { return rr[e]; });
    }
    __syncwarp();
}

// lp_solve on the register-resident H (same steps, same order; see lp_solve in search.cu for the commentary).
// On entry ybuf mirrors row 0.  art != 0: the artificial column at position qd-1 leaves first with that sign.
// Written as one loop in which every step appears ONCE in the instruction stream (values, pricing, alpha/beta, ratio
// test, replacement).
template <int QP, int MC>
__device__ __forceinline__ int r_solve(const W<MC> &w, int side, int qd, double (&h)[QP], double radius, double cmax, double &d_out, int art,
                                       int &failcode) {
    const int m = w.m, lane = w.lane;
    const int cap = 100 * m + 1000;
    bool refactored = false, bland = false;
    bool chk_flip = true;        // the sign of t is looked at after the first evaluation, an artificial's exit, a repair, a refactorisation
    bool art_repaired = false;   // the artificial's alpha/beta were recomputed after a repair: no second repair
    int want_refactor = 0;       // failure code of the refactorisation asked for (0 = none)
    int iter = 0, stall = 0;
    double tbest = -1.0, myval = 0.0;
#pragma unroll 1
    for (;;) {
        int pe, se;
        if (want_refactor) {
            if (refactored) { failcode = want_refactor; return NODE_FAIL; }
            refactored = true;
            if (r_refactor<QP>(w, side, qd, h)) { failcode = want_refactor + 1; return NODE_FAIL; }
            want_refactor = 0; chk_flip = true; ++iter;
        }
        if (art) { pe = qd - 1; se = art; }
        else {
            myval = rdot<QP>(h, w.chat());
            if (lane < QP) w.val()[lane] = myval;
            __syncwarp();
            double t = __shfl_sync(FULLMASK, myval, 0);
            if (chk_flip) { chk_flip = false; if (t < 0.0) { r_flip<QP>(w, side, h, myval); t = -t; } }
            if (!approxLE(t, radius)) { if (iter == 0) w.count(C_QUICK); return NODE_PRUNE; }
            if (t > tbest * (1.0 + 1.0e-12)) { tbest = t; stall = 0; } else if (++stall >= 2 * STALL_MAX) bland = true;
            const bool dantzig = stall >= STALL_MAX;
            const double tol = EPS_OPT * (cmax + t);
            pe = -1;
            const double a = fabs(myval);
            const bool viol = lane >= 1 && lane < qd && a > t + tol;
            if (!bland) {
                double best = -1.0;
                if (viol) {
                    const double msp = myval > 0.0 ? -1.0 : 1.0;
                    const double *yb = w.ybuf();
                    double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;
#pragma unroll
                    for (int e = 0; e < QP; e += 4) {
                        const double2 y0 = *(const double2 *)(yb + e), y1 = *(const double2 *)(yb + e + 2);
                        const double x0 = fma(msp, y0.x, h[e]), x1 = fma(msp, y0.y, h[e + 1]);      // h - sp*y, product exact
                        const double x2 = fma(msp, y1.x, h[e + 2]), x3 = fma(msp, y1.y, h[e + 3]);
                        a0 = fma(x0, x0, a0); a1 = fma(x1, x1, a1); a2 = fma(x2, x2, a2); a3 = fma(x3, x3, a3);
                    }
                    const double nr = (a0 + a1) + (a2 + a3);
                    const double vi = a - t;
                    const double score = dantzig ? vi : vi * vi / nr;
                    if (score > best) { best = score; pe = lane; }
                }
                warp_argmax_nn(best, pe);
            } else {
                int bp = -1; double key = 0.0;
                if (viol) { key = -(double)w.zrow(side)[lane]; bp = lane; }
                warp_argmax(key, bp);
                pe = bp;
            }
            if (pe < 0) {
                r_image<MC>(w, side, qd < m ? qd + 1 : qd);
                double err = 0.0;
                if (lane < qd) err = fabs(w.v()[lane] - w.chat()[lane]);
                err = warp_max(err);
                if (err <= 1.0e-8 * ((double)m * cmax + t)) { d_out = t; return NODE_EXPLORE; }
                want_refactor = 1;
                continue;
            }
            if (iter >= cap) { failcode = 3; return NODE_FAIL; }
            se = __shfl_sync(FULLMASK, myval, pe) > 0.0 ? 1 : -1;
        }
        if (lane == pe) row_to_smem<QP>(h, w.rbuf());
        __syncwarp();
        double al[MC / 32], be[MC / 32];
        double bmax;
        bool need;
        const double nrm1 = r_alpha_beta<QP>(w, side, al, be, bmax, need);
        if (!art_repaired && r_repair<QP>(w, side, qd, h, al, need)) {
            if (art) art_repaired = true;             // same leaving row, alpha/beta once more
            else { chk_flip = true; ++iter; }         // the vertex moved a little: price again
            continue;
        }
        if (!art && fabs(nrm1 - 1.0) > 1.0e-5) { want_refactor = 4; continue; }
        const int l = r_ratio<QP>(w, side, al, be, se, pe, bland, bmax);
        if (l < 0) { failcode = art ? 6 : 7; return NODE_FAIL; }
        if (r_replace<QP>(w, side, h, pe, l, se)) { failcode = 8; return NODE_FAIL; }
        if (art) { art = 0; chk_flip = true; art_repaired = false; }
        else ++iter;
    }
}

// ---- generic-path helpers (not inlined; they get a WS built by to_ws) --------------------------------------------
__device__ __forceinline__ void side_arrays(WS &s, int side) {
    s.zrow = side ? s.zrow_b : s.zrow_a; s.pos = side ? s.pos_b : s.pos_a; s.sg = side ? s.sg_b : s.sg_a;
}
// The parent's vertex (packed qpar x qpar at dH) extended by the level's coordinate, built in the parked matrix; the
// arrays of `side` receive its rows and signs.
__device__ __noinline__ void build_extended_parent(WS &s, int side, const double *dH, const int *dI, const signed char *dB, int qpar) {
    side_arrays(s, side);
    s.W = s.park; s.LD = LDS_C;
    chain_load(s, dH, dI, dB, qpar);
    lp_extend(s);
}
// a stored vertex (packed q x q) into the parked matrix and the arrays of `side`
__device__ __noinline__ void load_vertex_parked(WS &s, int side, const double *dH, const int *dI, const signed char *dB, int q) {
    side_arrays(s, side);
    s.W = s.park; s.LD = LDS_C;
    chain_load(s, dH, dI, dB, q);
}
// Last resort after a failed warm start, on the warp's global buffer: rebuild the vertex coordinate by coordinate from
// nothing, every intermediate problem solved to optimality (chat holds all qd0 entries).  The result is in Wg_a.
__device__ __noinline__ int cold_solve_global(WS &s, int side, int qd0, double radius, double cmax, double *d_out) {
    side_arrays(s, side);
    s.W = s.Wg_a; s.LD = s.LDG;
    int status = NODE_FAIL;
    double d = 0.0;
    if (!lp_init1(s)) {
#pragma unroll 1
        for (;;) {
            int art = 0;
            double rad = radius;
            if (s.qd < qd0) {
                lp_extend(s);
                lp_values<false>(s);
                art = s.val[s.qd - 1] > 0.0 ? 1 : -1;
                if (s.qd < qd0) rad = INFINITY;
            }
            status = lp_solve<false>(s, rad, cmax, d, art);
            if (status == NODE_FAIL) break;
            if (s.qd == qd0) break;
            if (status != NODE_EXPLORE) { status = NODE_FAIL; break; }
        }
    }
    *d_out = d;
    return status;
}
// a side that lost its vertex restarts from the parent's, extended: rebuilt in Wg_a
__device__ __noinline__ void rebuild_parent_global(WS &s, int side, int qpar) {
    side_arrays(s, side);
    s.W = s.Wg_a; s.LD = s.LDG;
    chain_load(s, s.Wpar, s.pzrow, s.psg, qpar);
    lp_extend(s);
}

// ---- rows of H between registers and packed q x q matrices in global memory (records) --------------------------
template <int QP>
__device__ __forceinline__ void rows_to_global(const double (&h)[QP], double *dst, int q, int lane) {
    if (lane < q) {
        double *row = dst + (size_t)lane * q;
#pragma unroll
        for (int e = 0; e < QP; ++e) if (e < q) row[e] = h[e];
    }
}
// packed q x q matrix in global memory <- the parked matrix
__device__ __forceinline__ void park_to_global(const double *park, double *dst, int q, int lane) {
    if (lane < q) {
        const double *src = park + lane;
        double *d = dst + lane;
#pragma unroll 1
        for (int i = 0; i < q; ++i) d[(size_t)i * q] = src[(size_t)i * LDS_C];
    }
}
// zrow / sg / pos of `side` <- a stored vertex (the matrix itself goes elsewhere)
template <int MC>
__device__ __forceinline__ void side_arrays_load(const W<MC> &w, int side, const int *dI, const signed char *dB, int q) {
    int *zrow = w.zrow(side);
    unsigned char *pos = w.pos(side);
    signed char *sg = w.sg(side);
#pragma unroll 1
    for (int j = w.lane; j < w.m; j += 32) { sg[j] = dB[j]; pos[j] = 0; }
#pragma unroll 1
    for (int p = w.lane; p < q; p += 32) zrow[p] = dI[p];
    __syncwarp();
#pragma unroll 1
    for (int p = 1 + w.lane; p < q; p += 32) if (zrow[p] >= 0) pos[zrow[p]] = (unsigned char)p;
    __syncwarp();
}

// One task (the sibling loop below one explored node) with bases of 2..QP rows: twin of expand_impl in search.cu, node
// LPs solved by r_solve.
template <int QP, int MC>
__device__ __forceinline__ int expand_reg(const Params &P, const W<MC> &w, const unsigned char *R, int pool_slot, unsigned char *next,
                                          Arena &A, int gwarp, unsigned int *hist) {
    const int n = P.n, m = P.m, lane = w.lane;
    const RecLayout RL = rec_layout(n, m);
    const int k0 = ((const int *)R)[0];
    const int rflags = ((const int *)R)[1];
    const double xc0 = ((const double *)R)[1];
    const double gb = ((const double *)R)[2];
    const bool is_top = (rflags & REC_TOP) != 0, is_cont = (rflags & REC_CONT) != 0;
    const int emax = (rflags & REC_DIVE) ? 1 : P.eager_max;
    const int c_step = ((const int *)R)[6], c_state = ((const int *)R)[7];
    const int kk = k0 - 1, qd0 = m - kk, qpar = m - k0;  // 2 <= qd0 <= QP, k0 > 1

    double h[QP];
    int inited = is_cont ? ((c_state >> 1) & 3) : 0;
    int in_regs = -2;  // -2: registers hold the start state common to both sides; otherwise the side whose matrix they hold
    int lost = 0;      // sides whose matrix was lost to a failed solve: they restart from the parent's vertex
    long long *u = w.u();
    ord_t *ord = w.ord();
    double *chat = w.chat(), *cpar = w.cpar(), *brow = w.brow(), *cc = w.c();
    // ---- take what is needed out of the record; afterwards the record is dead ----
#pragma unroll 1
    for (int e = lane; e < n; e += 32) { u[e] = ((const long long *)(R + RL.off_u))[e]; ord[e] = ((const ord_t *)(R + RL.off_ord))[e]; }
    {
        signed char *psg = w.psg();
        int *pzrow = w.pzrow();
#pragma unroll 1
        for (int j = lane; j < m; j += 32) { cpar[j] = ((const double *)(R + RL.off_cpar))[j]; psg[j] = ((const signed char *)(R + RL.off_sg))[j]; }
#pragma unroll 1
        for (int p = lane; p < qpar; p += 32) pzrow[p] = ((const int *)(R + RL.off_zrow))[p];
    }
    copy_words(w.Wpar, R + RL.off_H, (size_t)qpar * qpar * 8, lane);  // kept for parked loops and for a side that loses its vertex
    {
        WS t;
        to_ws<MC>(w, t);
        const double *pH = (const double *)(R + RL.off_H);
        const int *pI = (const int *)(R + RL.off_zrow);
        const signed char *pB = (const signed char *)(R + RL.off_sg);
        if (inited != 0) {
            // resumed loop.  Both sides under way: side 0 is parked, side 1 goes to the registers.  One side under way:
            // it goes to the registers, the other starts from the extended parent (parked).
            const ContLayout CL = cont_layout(RL, m, k0);
            const int sd = inited == 1 ? 0 : 1;
            if (inited == 3) load_vertex_parked(t, 0, (const double *)(R + CL.off_HA), (const int *)(R + CL.off_zA), (const signed char *)(R + CL.off_sA), qd0);
            else build_extended_parent(t, 1 - sd, pH, pI, pB, qpar);
            side_arrays_load<MC>(w, sd, (const int *)(R + (sd ? CL.off_zB : CL.off_zA)), (const signed char *)(R + (sd ? CL.off_sB : CL.off_sA)), qd0);
            rows_from_global_ld<QP>(h, (const double *)(R + (sd ? CL.off_HB : CL.off_HA)), qd0, qd0, lane);
            in_regs = sd;
        } else {  // fresh loop: both sides start from the extended parent
            build_extended_parent(t, 0, pH, pI, pB, qpar);
            // the byte arrays are padded to whole 8-byte words (SL<MC>::NB)
#pragma unroll 1
            for (int j = lane; j < (m + 7) / 8; j += 32) {
                ((unsigned long long *)w.sg(1))[j] = ((const unsigned long long *)w.sg(0))[j];
                ((unsigned long long *)w.pos(1))[j] = ((const unsigned long long *)w.pos(0))[j];
            }
#pragma unroll 1
            for (int p = lane; p < qd0; p += 32) w.zrow(1)[p] = w.zrow(0)[p];
            __syncwarp();
            park_load<QP>(h, w.park(), qd0, lane);
        }
    }
    // chat: the parent's coordinates, this level's coordinate (set per node), zeros up to QP
    if (lane < QP) chat[lane] = lane < qpar ? ((const double *)(R + RL.off_chat))[lane] : 0.0;
    __syncwarp();
    if (pool_slot >= 0) {
        __threadfence();
        if (lane == 0) ring_push(P.q, &P.q->ftail, P.fring, P.ring_mask, pool_slot);  // the record slot is free again
    }

    double d = 0.0, xlast = 0.0;
    // ---- constants of this level ----
#pragma unroll 1
    for (int j = lane; j < m; j += 32) brow[j] = P.B[(size_t)kk * m + j];
    const double Rkk = P.Rm[(size_t)kk * n + kk];
    double Rk1 = 1.0, Rk1k = 0.0, g2 = 0.0;
    {
        const double *Rr = P.Rm + (size_t)(kk - 1) * n;   // kk >= 1 on this path
        Rk1 = Rr[kk - 1]; Rk1k = Rr[kk];
        double part = 0.0;
#pragma unroll 1
        for (int i = kk + 1 + lane; i < n; i += 32) part = fma(-(double)u[i], Rr[i], part);
        g2 = P.QTp[kk - 1] + butterfly(part);
    }
    long long ucur = (long long)round(xc0);
    int step = sgn_d(xc0 - (double)ucur);
    bool pruned = false;
    int nkept = 0, nexp = 0;
    if (is_cont) { ucur = u[kk]; step = c_step; pruned = (c_state & 1) != 0; }
    else if (lane == 0) { u[kk] = ucur; ord[kk] = 0; }
    __syncwarp();
    const bool forced = (n - k0) < P.static_depth;
    unsigned int wd = 0;

    for (;;) {
        {
            int ab = 0;
            if (lane == 0) {
                ab = *(volatile int *)&P.q->abort;
                if (!ab && (++wd & 63u) == 0) ab = overdue(P.q, 1) ? 1 : 0;
            }
            if (__shfl_sync(FULLMASK, ab, 0)) return nkept;
        }
        double cmax = 0.0;
        {
            const double mu = -(double)ucur;
#pragma unroll 1
            for (int j = lane; j < m; j += 32) { const double cj = fma(mu, brow[j], cpar[j]); cc[j] = cj; const double a = fabs(cj); cmax = a > cmax ? a : cmax; }
            cmax = warp_max(cmax);
            __syncwarp();
        }
        if (lane == 0) { w.ctr()[C_NODES]++; hist[kk]++; }
        int status;
        double gn = 0.0;
        const int dir = sgn_d(xc0 - (double)ucur);
        const int side = dir > 0 ? 0 : 1;
        if (in_regs == -2) in_regs = side;
        else if (in_regs != side) { park_swap<QP>(h, w.park(), qd0, lane); in_regs = side; __syncwarp(); }
        if (lane == 0) chat[qd0 - 1] = fma(-(double)ucur, Rkk, gb);
        const bool first = !((inited >> side) & 1);
        if ((lost >> side) & 1) {  // the side lost its vertex: the parent's, extended, rebuilt in the warp's global buffer
            WS t;
            to_ws<MC>(w, t);
            rebuild_parent_global(t, side, qpar);
            rows_from_global_ld<QP>(h, w.Wg_a, w.LDG, qd0, lane);
            lost &= ~(1 << side);
        }
        inited |= 1 << side;
        publish_y<QP>(w, h);
        int failcode = 0;
        status = r_solve<QP>(w, side, qd0, h, P.radius, cmax, d, first ? dir : 0, failcode);
        if (status == NODE_FAIL) {
            WS t;
            to_ws<MC>(w, t);
            status = cold_solve_global(t, side, qd0, P.radius, cmax, &d);
            if (status != NODE_FAIL) rows_from_global_ld<QP>(h, w.Wg_a, w.LDG, qd0, lane);
            else failcode = t.failcode;
        }
        if (status == NODE_FAIL) {
            // Remember the failure and go on as if the node had been pruned: another warp may find a point inside
            // the box, and then the failure is irrelevant (the sequential order might never have come here).
            if (lane == 0 && atomicCAS(&P.q->fail_info, 0, failcode | (k0 << 8) | (qd0 << 16)) == 0) { for (int e = 0; e < n; ++e) P.q->fail_u[e] = u[e]; }
            inited &= ~(1 << side);
            lost |= 1 << side;
            status = NODE_PRUNE;
        }
        if (status == NODE_EXPLORE) {  // real centre of the next level: x_{kk-1} of the optimum
            gn = fma(-(double)ucur, Rk1k, g2);
            xlast = (gn - w.v()[qd0]) / Rk1;
        }
        if (status == NODE_PRUNE) {
            if (is_top) return nkept;  // closest.cc:280-286: first prune at the top level ends the search
            w.count(C_NPRUNE);
            if (pruned) return nkept;  // second prune: the level is left
            pruned = true;
        } else {
            // explored inner node: its children are one more task
            w.count(C_NITER);
            unsigned char *C = nullptr;
            int slot = -1, dest = 0;  // dest: 0 next, 1 arena, 2 queue, 3 static-split output
            const size_t need = rec_bytes(RL, m, k0 - 1);
            if (forced) {
                if (lane == 0) slot = ring_pop(P.q, &P.q->fhead, &P.q->ftail, P.fring, P.ring_mask);
                slot = __shfl_sync(FULLMASK, slot, 0);
                if (slot < 0) { fail_search(P, lane, BLTCU_STATUS_ERROR); return nkept; }
                C = P.pool + (size_t)slot * RL.stride;
                dest = ((n - (k0 - 1)) >= P.static_depth) ? 3 : 2;
            } else if (nkept == 0) { C = next; dest = 0; }
            else {
                const size_t top = (size_t)A.offs[A.acount];
                if (A.acount + 1 < A.slots && top + need <= A.cap) { C = A.base + top; dest = 1; }
                else {
                    if (lane == 0) slot = ring_pop(P.q, &P.q->fhead, &P.q->ftail, P.fring, P.ring_mask);
                    slot = __shfl_sync(FULLMASK, slot, 0);
                    if (slot < 0) { fail_search(P, lane, BLTCU_STATUS_ERROR); return nkept; }
                    C = P.pool + (size_t)slot * RL.stride;
                    dest = 2;
                }
            }
            // the child's record: header, path, right-hand side, coordinates, vertex
            if (lane == 0) { ((int *)C)[0] = k0 - 1; ((int *)C)[1] = ((rflags & REC_DIVE) && dest == 0) ? REC_DIVE : 0; ((double *)C)[1] = xlast; ((double *)C)[2] = gn; ((int *)C)[6] = 0; ((int *)C)[7] = 0; }
#pragma unroll 1
            for (int e = lane; e < n; e += 32) { ((long long *)(C + RL.off_u))[e] = u[e]; ((ord_t *)(C + RL.off_ord))[e] = ord[e]; }
#pragma unroll 1
            for (int j = lane; j < m; j += 32) ((double *)(C + RL.off_cpar))[j] = cc[j];
            if (lane < qd0) ((double *)(C + RL.off_chat))[lane] = chat[lane];
            if (k0 - 1 > 1) {
                rows_to_global<QP>(h, (double *)(C + RL.off_H), qd0, lane);
                const int *zr = w.zrow(side);
                const signed char *sgs = w.sg(side);
                if (lane < qd0) ((int *)(C + RL.off_zrow))[lane] = zr[lane];
#pragma unroll 1
                for (int j = lane; j < m; j += 32) ((signed char *)(C + RL.off_sg))[j] = sgs[j];
            }
            __syncwarp();
            nexp++;
            if (dest == 0) nkept = 1;
            else if (dest == 1) { if (lane == 0) A.offs[A.acount + 1] = (int)(A.offs[A.acount] + need); A.acount++; __syncwarp(); }
            else {
                __threadfence();
                if (lane == 0) {
                    if (dest == 3) { const unsigned int o = atomicAdd(&P.q->out_count, 1u); P.out_list[o] = slot; }
                    else { atomicAdd(&P.q->pending, 1); ring_push(P.q, &P.q->tail, P.ring, P.ring_mask, slot); }
                }
                w.count(C_DONATED);
            }
        }
        // next sibling, closest.cc:118-121
        ucur += step; step = -step - (step > 0 ? 1 : -1);
        if (lane == 0) { u[kk] = ucur; if (ord[kk] < ORD_MAX) ord[kk]++; }
        __syncwarp();
        // wide level or little stack space left: park the rest of this loop and go down into the first child
        if (status == NODE_EXPLORE && !forced && (nexp >= emax || (size_t)A.offs[A.acount] * 2 > A.cap)) {
            const ContLayout CL = cont_layout(RL, m, k0);
            unsigned char *C = nullptr;
            int slot = -1;
            const size_t top = (size_t)A.offs[A.acount];
            const bool in_arena = A.acount + 1 < A.slots && top + CL.bytes <= A.cap;
            if (in_arena) C = A.base + top;
            else {
                if (lane == 0) slot = ring_pop(P.q, &P.q->fhead, &P.q->ftail, P.fring, P.ring_mask);
                slot = __shfl_sync(FULLMASK, slot, 0);
                if (slot < 0) { fail_search(P, lane, BLTCU_STATUS_ERROR); return nkept; }
                C = P.pool + (size_t)slot * RL.stride;
            }
            const int live = inited & ~lost;  // sides whose vertex exists (a lost side restarts from the parent on resume)
            if (lane == 0) {
                ((int *)C)[0] = k0; ((int *)C)[1] = (is_top ? REC_TOP : 0) | REC_CONT; ((double *)C)[1] = xc0; ((double *)C)[2] = gb;
                ((int *)C)[6] = step; ((int *)C)[7] = (pruned ? 1 : 0) | (live << 1); ((int *)C)[8] = (int)CL.bytes;
            }
            {
                const signed char *psg = w.psg();
                const int *pzrow = w.pzrow();
#pragma unroll 1
                for (int e = lane; e < n; e += 32) { ((long long *)(C + RL.off_u))[e] = u[e]; ((ord_t *)(C + RL.off_ord))[e] = ord[e]; }
#pragma unroll 1
                for (int j = lane; j < m; j += 32) { ((double *)(C + RL.off_cpar))[j] = cpar[j]; ((signed char *)(C + RL.off_sg))[j] = psg[j]; }
#pragma unroll 1
                for (int e = lane; e < qpar; e += 32) { ((double *)(C + RL.off_chat))[e] = chat[e]; ((int *)(C + RL.off_zrow))[e] = pzrow[e]; }
            }
            copy_words(C + RL.off_H, w.Wpar, (size_t)qpar * qpar * 8, lane);
            for (int sd = 0; sd < 2; ++sd) {
                if (!((live >> sd) & 1)) continue;
                double *dH = (double *)(C + (sd ? CL.off_HB : CL.off_HA));
                if (sd == in_regs) rows_to_global<QP>(h, dH, qd0, lane);
                else park_to_global(w.park(), dH, qd0, lane);
                const int *zr = w.zrow(sd);
                const signed char *sgs = w.sg(sd);
                int *dI = (int *)(C + (sd ? CL.off_zB : CL.off_zA));
                signed char *dB = (signed char *)(C + (sd ? CL.off_sB : CL.off_sA));
#pragma unroll 1
                for (int p = lane; p < qd0; p += 32) dI[p] = zr[p];
#pragma unroll 1
                for (int j = lane; j < m; j += 32) dB[j] = sgs[j];
            }
            __syncwarp();
            if (in_arena) { if (lane == 0) A.offs[A.acount + 1] = (int)(top + CL.bytes); A.acount++; __syncwarp(); }
            else {
                __threadfence();
                if (lane == 0) { atomicAdd(&P.q->pending, 1); ring_push(P.q, &P.q->tail, P.ring, P.ring_mask, slot); }
                w.count(C_DONATED);
            }
            return nkept;
        }
    }
}
