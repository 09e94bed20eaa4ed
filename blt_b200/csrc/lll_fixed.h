// lll_fixed.h -- the exact integral LLL of host/lll.cpp (Cohen Alg. 2.6.7 as NTL::LLL(det2,B,U) runs it, delta = 3/4,
// dependent rows moved to the front; reference seam: libblt/ntl_wrapper.cc:37-52) on FIXED-CAPACITY multiprecision
// integers, written once for host and device.  One problem is reduced by one thread: the batched kernel in
// lll_batch.cu runs thousands of small problems side by side (where the
// reduction and not the search is the time-to-decision).  A number that does not fit its W limbs raises the problem's
// overflow flag; the caller then reduces that problem with the arbitrary-precision host implementation.
//
// Same sequence of operations as host/lll.cpp, hence the same reduced basis and the same U, bit for bit
// (tests/test_host_layer.py compares the two on the host, tests/test_gpu_lll.py on the device).
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define LF_HD __host__ __device__
#else
#define LF_HD
#endif

namespace lllfix {

// sign-magnitude, base 2^32, little endian, no leading zeros; len == 0 is zero
template <int W>
struct Num {
    uint32_t d[W];
    int len;
    int neg;
};

template <int W>
struct Arith {
    typedef Num<W> N;
    typedef Num<2 * W + 2> Wd;  // products and sums of two products
    int ovf;

    LF_HD Arith() : ovf(0) {}

    template <int A>
    LF_HD static void set_i64(Num<A> &r, int64_t v) {
        r.neg = v < 0;
        uint64_t a = r.neg ? (uint64_t)0 - (uint64_t)v : (uint64_t)v;
        r.len = 0;
        while (a) { r.d[r.len++] = (uint32_t)a; a >>= 32; }
        if (r.len == 0) r.neg = 0;
    }
    template <int A>
    LF_HD static int64_t to_i64(const Num<A> &x) {  // low 64 bits with sign, like NTL's to_long
        uint64_t a = 0;
        if (x.len > 0) a |= x.d[0];
        if (x.len > 1) a |= (uint64_t)x.d[1] << 32;
        return (int64_t)(x.neg ? (uint64_t)0 - a : a);
    }
    template <int A, int B>
    LF_HD static int cmp_mag(const Num<A> &a, const Num<B> &b) {
        if (a.len != b.len) return a.len < b.len ? -1 : 1;
        for (int i = a.len; i-- > 0;)
            if (a.d[i] != b.d[i]) return a.d[i] < b.d[i] ? -1 : 1;
        return 0;
    }
    template <int A, int B>
    LF_HD static int cmp(const Num<A> &a, const Num<B> &b) {
        if (a.neg != b.neg) return a.neg ? -1 : 1;
        const int c = cmp_mag(a, b);
        return a.neg ? -c : c;
    }
    // r = a (narrowing or widening copy)
    template <int A, int B>
    LF_HD void copy(Num<A> &r, const Num<B> &a) {
        if (a.len > A) { ovf = 1; r.len = 0; r.neg = 0; return; }
        for (int i = 0; i < a.len; ++i) r.d[i] = a.d[i];
        r.len = a.len; r.neg = a.neg;
    }
    // r = a * b
    template <int A, int B>
    LF_HD void mul(Wd &r, const Num<A> &a, const Num<B> &b) {
        if (a.len == 0 || b.len == 0) { r.len = 0; r.neg = 0; return; }
        const int rl = a.len + b.len;
        if (rl > 2 * W + 2) { ovf = 1; r.len = 0; r.neg = 0; return; }
        for (int i = 0; i < rl; ++i) r.d[i] = 0;
        for (int i = 0; i < a.len; ++i) {
            uint64_t c = 0;
            const uint64_t ai = a.d[i];
            for (int j = 0; j < b.len; ++j) {
                c += ai * b.d[j] + r.d[i + j];
                r.d[i + j] = (uint32_t)c; c >>= 32;
            }
            r.d[i + b.len] = (uint32_t)c;
        }
        r.len = rl;
        while (r.len > 0 && r.d[r.len - 1] == 0) --r.len;
        r.neg = a.neg != b.neg;
    }
    // r = a + sb * b  (sb = +1 or -1), all wide
    LF_HD void addsub(Wd &r, const Wd &a, const Wd &b, int sb) {
        const int bneg = b.len ? (sb > 0 ? b.neg : !b.neg) : 0;
        if (a.neg == bneg || a.len == 0 || b.len == 0) {
            if (b.len == 0) { r = a; return; }
            if (a.len == 0) { r = b; r.neg = bneg; return; }
            const Wd &x = a.len >= b.len ? a : b, &y = a.len >= b.len ? b : a;
            uint64_t c = 0;
            int i = 0;
            for (; i < x.len; ++i) {
                c += (uint64_t)x.d[i] + (i < y.len ? y.d[i] : 0);
                r.d[i] = (uint32_t)c; c >>= 32;
            }
            int len = x.len;
            if (c) {
                if (len >= 2 * W + 2) { ovf = 1; r.len = 0; r.neg = 0; return; }
                r.d[len++] = (uint32_t)c;
            }
            r.len = len; r.neg = a.neg;
            return;
        }
        const int c = cmp_mag(a, b);
        if (c == 0) { r.len = 0; r.neg = 0; return; }
        const Wd &x = c > 0 ? a : b, &y = c > 0 ? b : a;
        int64_t br = 0;
        for (int i = 0; i < x.len; ++i) {
            br += (int64_t)x.d[i] - (i < y.len ? y.d[i] : 0);
            r.d[i] = (uint32_t)br; br >>= 32;
        }
        r.len = x.len;
        while (r.len > 0 && r.d[r.len - 1] == 0) --r.len;
        r.neg = c > 0 ? a.neg : bneg;
    }
    // magnitude division (Knuth D): q = |a| / |b|, rem = |a| % |b|; b != 0
    template <int B>
    LF_HD void divmod_mag(const Wd &a, const Num<B> &b, Wd &q, Wd &rem) {
        q.len = 0; q.neg = 0; rem.len = 0; rem.neg = 0;
        if (b.len == 0) { ovf = 1; return; }  // only after an earlier overflow truncated a divisor
        if (cmp_mag(a, b) < 0) { for (int i = 0; i < a.len; ++i) rem.d[i] = a.d[i]; rem.len = a.len; return; }
        if (b.len == 1) {
            const uint64_t dv = b.d[0];
            uint64_t r = 0;
            for (int i = a.len; i-- > 0;) { r = (r << 32) | a.d[i]; q.d[i] = (uint32_t)(r / dv); r %= dv; }
            q.len = a.len;
            while (q.len > 0 && q.d[q.len - 1] == 0) --q.len;
            if (r) { rem.d[0] = (uint32_t)r; rem.len = 1; }
            return;
        }
#if defined(__CUDA_ARCH__)
        const int s = __clz((int)b.d[b.len - 1]);
#else
        const int s = __builtin_clz(b.d[b.len - 1]);
#endif
        const int n = b.len, m = a.len - n;
        uint32_t u[2 * W + 3], v[B];
        for (int i = n; i-- > 0;) v[i] = (b.d[i] << s) | ((s && i) ? (b.d[i - 1] >> (32 - s)) : 0);
        u[a.len] = s ? (a.d[a.len - 1] >> (32 - s)) : 0;
        for (int i = a.len; i-- > 0;) u[i] = (a.d[i] << s) | ((s && i) ? (a.d[i - 1] >> (32 - s)) : 0);
        for (int j = m + 1; j-- > 0;) {
            const uint64_t num = ((uint64_t)u[j + n] << 32) | u[j + n - 1];
            uint64_t qh = num / v[n - 1], rh = num % v[n - 1];
            while (qh >= (1ull << 32) || qh * v[n - 2] > ((rh << 32) | u[j + n - 2])) {
                --qh; rh += v[n - 1];
                if (rh >= (1ull << 32)) break;
            }
            int64_t borrow = 0;
            uint64_t carry = 0;
            for (int i = 0; i < n; ++i) {
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
                for (int i = 0; i < n; ++i) { c += (uint64_t)u[i + j] + v[i]; u[i + j] = (uint32_t)c; c >>= 32; }
                u[j + n] += (uint32_t)c;
            }
            q.d[j] = (uint32_t)qh;
        }
        q.len = m + 1;
        while (q.len > 0 && q.d[q.len - 1] == 0) --q.len;
        for (int i = 0; i < n; ++i) rem.d[i] = (u[i] >> s) | (s ? (uint32_t)((uint64_t)u[i + 1] << (32 - s)) : 0);
        rem.len = n;
        while (rem.len > 0 && rem.d[rem.len - 1] == 0) --rem.len;
    }
    // q = floor(a / d), r = a - q d  (d > 0)
    template <int B>
    LF_HD void floordiv(const Wd &a, const Num<B> &d, Wd &q, Wd &r) {
        divmod_mag(a, d, q, r);
        if (!a.neg) return;
        if (r.len == 0) { q.neg = q.len ? 1 : 0; return; }
        // q = -(q + 1), r = d - r
        Wd one, t;
        set_i64(one, 1);
        q.neg = 0;
        addsub(t, q, one, +1);
        q = t; q.neg = 1;
        Wd dw;
        copy(dw, d); dw.neg = 0;
        r.neg = 0;
        addsub(t, dw, r, -1);
        r = t;
    }
    // r = floor((a*b + s*(c*d)) / e), e > 0, narrowed to W limbs
    template <int A1, int A2, int A3, int A4, int A5>
    LF_HD void muladd_div(N &r, const Num<A1> &a, const Num<A2> &b, int s, const Num<A3> &c, const Num<A4> &d, const Num<A5> &e) {
        Wd p1, p2, sum, q, rem;
        mul(p1, a, b);
        mul(p2, c, d);
        addsub(sum, p1, p2, s);
        floordiv(sum, e, q, rem);
        copy(r, q);
    }
    template <int A1, int A2>
    LF_HD void fdiv(N &r, const Num<A1> &a, const Num<A2> &e) {
        Wd aw, q, rem;
        copy(aw, a);
        floordiv(aw, e, q, rem);
        copy(r, q);
    }
    // nearest integer to a/d (d > 0), ties towards zero
    LF_HD void baldiv(N &res, const N &a, const N &d) {
        Wd aw, q, r, r2, dw, one, t;
        copy(aw, a);
        floordiv(aw, d, q, r);
        addsub(r2, r, r, +1);
        copy(dw, d);
        const int c = cmp(r2, dw);
        if (c > 0 || (c == 0 && q.neg)) { set_i64(one, 1); addsub(t, q, one, +1); q = t; }
        copy(res, q);
    }
    // r = a - q*b
    LF_HD void submul(N &r, const N &a, const N &q, const N &b) {
        Wd p, aw, t;
        mul(p, q, b);
        copy(aw, a);
        addsub(t, aw, p, -1);
        copy(r, t);
    }
    // r = x*a + y*b
    LF_HD void lin2(N &r, const N &x, const N &a, const N &y, const N &b) {
        Wd p1, p2, t;
        mul(p1, x, a);
        mul(p2, y, b);
        addsub(t, p1, p2, +1);
        copy(r, t);
    }
    // g = a*x + b*y >= 0 : classical Euclid on magnitudes, signs restored afterwards
    LF_HD void xgcd(const N &a, const N &b, N &g, N &x, N &y) {
        N u = a, v = b, u1, v1, u2, v2, qn, t;
        u.neg = 0; v.neg = 0;
        set_i64(u1, 1); set_i64(v1, 0); set_i64(u2, 0); set_i64(v2, 1);
        while (v.len != 0) {
            Wd uw, q, r;
            copy(uw, u);
            floordiv(uw, v, q, r);
            copy(qn, q);
            u = v; copy(v, r);
            submul(t, u1, qn, u2); u1 = u2; u2 = t;
            submul(t, v1, qn, v2); v1 = v2; v2 = t;
        }
        if (a.neg && u1.len) u1.neg = !u1.neg;
        if (b.neg && v1.len) v1.neg = !v1.neg;
        g = u; x = u1; y = v1;
    }
};

// Who runs the reduction: one thread (host, or a single device thread), or the 32 lanes of a warp.  In the warp form the
// loops over a row of B, U or lam (the bulk of the work: size reduction and the swap updates) are dealt to the lanes,
// scalar updates are done by lane 0, and every decision is computed by all lanes from the same memory, so control flow
// stays uniform.  Both forms perform the same operations on the same numbers.
struct OneLane {
    LF_HD int lane() const { return 0; }
    LF_HD int width() const { return 1; }
    LF_HD void sync() const {}
    LF_HD int any(int x) const { return x; }
};
#if defined(__CUDACC__)
struct WarpLanes {
    __device__ int lane() const { return (int)(threadIdx.x & 31); }
    __device__ int width() const { return 32; }
    __device__ void sync() const { __syncwarp(); }
    __device__ int any(int x) const { return __any_sync(0xffffffffu, x); }
};
#endif

// The reduction itself.  Storage is supplied by the caller (global memory on the device):
//   B: m x n, U: m x m, lam: (m+1) x (m+1) (1-based like host/lll.cpp), D: m+1, P: m+2 ints (P[m+1] = rank so far).
// Returns the rank, or -1 when a number overflowed its W limbs or the step limit was hit.
template <int W, class Lanes>
struct Reducer {
    typedef Num<W> N;
    typedef Arith<W> A;
    int m, n;
    N *B, *U, *lam, *D;
    int *P;
    A ar;
    Lanes L;

    LF_HD N &b(int i, int t) { return B[(long)i * n + t]; }
    LF_HD N &u(int i, int t) { return U[(long)i * m + t]; }
    LF_HD N &l(int k, int j) { return lam[(long)k * (m + 1) + j]; }
    LF_HD int &rank_so_far() { return P[m + 1]; }

    LF_HD void inner(N &r, int a, int c) {
        typename A::Wd acc, p, t;
        acc.len = 0; acc.neg = 0;
        for (int x = 0; x < n; ++x) { ar.mul(p, b(a - 1, x), b(c - 1, x)); ar.addsub(t, acc, p, +1); acc = t; }
        ar.copy(r, acc);
    }
    // a chain of dependent steps: lane 0
    LF_HD void incremental_gs(int k) {
        if (L.lane() == 0) {
            N uu;
            for (int j = 1; j < k; ++j) {
                const int posj = P[j];
                if (posj == 0) continue;
                inner(uu, k, j);
                for (int i = 1; i < posj; ++i) { N t; ar.muladd_div(t, D[i], uu, -1, l(k, i), l(j, i), D[i - 1]); uu = t; }
                l(k, posj) = uu;
            }
            inner(uu, k, k);
            const int s = rank_so_far();
            for (int i = 1; i <= s; ++i) { N t; ar.muladd_div(t, D[i], uu, -1, l(k, i), l(k, i), D[i - 1]); uu = t; }
            if (uu.len == 0) P[k] = 0;
            else { rank_so_far() = s + 1; P[k] = s + 1; D[s + 1] = uu; }
        }
        L.sync();
    }
    LF_HD void reduce(int k, int ll) {
        const int pl = P[ll];
        if (pl == 0) return;
        {
            typename A::Wd x, t, dw;
            ar.copy(x, l(k, pl));
            ar.addsub(t, x, x, +1);
            t.neg = 0;
            ar.copy(dw, D[pl]);
            if (A::cmp(t, dw) <= 0) return;
        }
        N r;
        ar.baldiv(r, l(k, pl), D[pl]);  // every lane computes the same multiplier
        L.sync();                       // ... before lane 0 changes lam[k][pl] below
        for (int t = L.lane(); t < n; t += L.width()) { N x; ar.submul(x, b(k - 1, t), r, b(ll - 1, t)); b(k - 1, t) = x; }
        for (int t = L.lane(); t < m; t += L.width()) { N x; ar.submul(x, u(k - 1, t), r, u(ll - 1, t)); u(k - 1, t) = x; }
        for (int j = 1 + L.lane(); j < ll; j += L.width()) {
            const int pj = P[j];
            if (pj != 0) { N x; ar.submul(x, l(k, pj), r, l(ll, pj)); l(k, pj) = x; }
        }
        if (L.lane() == 0) { N x; ar.submul(x, l(k, pl), r, D[pl]); l(k, pl) = x; }
        L.sync();
    }
    LF_HD void swap_num(N &x, N &y) { const N t = x; x = y; y = t; }
    LF_HD void swap_rows_data(int k) {  // rows k-1 and k of B, U and lam (columns of the rows above)
        for (int t = L.lane(); t < n; t += L.width()) swap_num(b(k - 2, t), b(k - 1, t));
        for (int t = L.lane(); t < m; t += L.width()) swap_num(u(k - 2, t), u(k - 1, t));
        for (int j = 1 + L.lane(); j <= k - 2; j += L.width()) { const int pj = P[j]; if (pj) swap_num(l(k - 1, pj), l(k, pj)); }
        L.sync();
    }
    LF_HD int swap_rows(int k, int max_k) {
        if (P[k] != 0) {
            swap_rows_data(k);
            const int pk = P[k];
            for (int i = k + 1 + L.lane(); i <= max_k; i += L.width()) {
                N t1, t2;
                ar.muladd_div(t1, l(i, pk - 1), l(k, pk - 1), +1, l(i, pk), D[pk - 2], D[pk - 1]);
                ar.muladd_div(t2, l(i, pk - 1), D[pk], -1, l(i, pk), l(k, pk - 1), D[pk - 1]);
                l(i, pk - 1) = t1; l(i, pk) = t2;
            }
            L.sync();
            if (L.lane() == 0) {
                N nd;
                ar.muladd_div(nd, D[pk], D[pk - 2], +1, l(k, pk - 1), l(k, pk - 1), D[pk - 1]);
                D[pk - 1] = nd;
            }
            L.sync();
            return 0;
        }
        const int pk1 = P[k - 1];
        if (l(k, pk1).len != 0) {
            N e, x, y, t1, t2, t3;  // the same in every lane
            ar.xgcd(l(k, pk1), D[pk1], e, x, y);
            ar.fdiv(t1, l(k, pk1), e);
            ar.fdiv(t2, D[pk1], e);
            t3 = t2;
            if (t2.len) t2.neg = !t2.neg;
            L.sync();
            for (int t = L.lane(); t < n; t += L.width()) {
                const N c1 = b(k - 2, t), c2 = b(k - 1, t);
                ar.lin2(b(k - 2, t), t1, c1, t2, c2);
                ar.lin2(b(k - 1, t), y, c1, x, c2);
            }
            for (int t = L.lane(); t < m; t += L.width()) {
                const N c1 = u(k - 2, t), c2 = u(k - 1, t);
                ar.lin2(u(k - 2, t), t1, c1, t2, c2);
                ar.lin2(u(k - 1, t), y, c1, x, c2);
            }
            for (int j = 1 + L.lane(); j <= k - 2; j += L.width()) {
                const int pj = P[j];
                if (!pj) continue;
                const N c1 = l(k - 1, pj), c2 = l(k, pj);
                ar.lin2(l(k - 1, pj), t1, c1, t2, c2);
                ar.lin2(l(k, pj), y, c1, x, c2);
            }
            L.sync();
            if (L.lane() == 0) {  // the rare branch (dependent rows): the divisions stay with one lane
                typename A::Wd t2sw;
                ar.mul(t2sw, t2, t2);
                N t2s;
                ar.copy(t2s, t2sw);
                { N q; ar.fdiv(q, D[pk1], t2s); D[pk1] = q; }
                for (int i = k + 1; i <= max_k; ++i) {
                    const int pi = P[i];
                    if (!pi) continue;
                    { N q; ar.fdiv(q, D[pi], t2s); D[pi] = q; }
                    for (int j = i + 1; j <= max_k; ++j) { N q; ar.fdiv(q, l(j, pi), t2s); l(j, pi) = q; }
                }
                for (int i = k + 1; i <= max_k; ++i) { N q; ar.fdiv(q, l(i, pk1), t3); l(i, pk1) = q; }
                const int tp = P[k - 1]; P[k - 1] = P[k]; P[k] = tp;
            }
            L.sync();
            return 1;
        }
        swap_rows_data(k);
        if (L.lane() == 0) { const int tp = P[k - 1]; P[k - 1] = P[k]; P[k] = tp; }
        L.sync();
        return 0;
    }
    // delta = a/b
    LF_HD int run(int a, int bq, long max_steps) {
        int k = 1, max_k = 0, force_reduce = 1;
        long steps = 0;
        while (k <= m) {
            if (L.any(ar.ovf) || ++steps > max_steps) return -1;
            if (k > max_k) { incremental_gs(k); max_k = k; }
            if (k == 1) { force_reduce = 1; ++k; continue; }
            if (force_reduce) for (int j = k - 1; j >= 1; --j) reduce(k, j);
            bool do_swap = false;
            if (P[k - 1] != 0) {
                if (P[k] == 0) do_swap = true;
                else {
                    const int pk = P[k];
                    // a * D[pk-1]^2  >  b * (D[pk] D[pk-2] + lam[k][pk-1]^2)
                    typename A::Wd d2, p1, p2, sum, lhs, rhs, aw, bw;
                    ar.mul(d2, D[pk - 1], D[pk - 1]);
                    ar.mul(p1, D[pk], D[pk - 2]);
                    ar.mul(p2, l(k, pk - 1), l(k, pk - 1));
                    ar.addsub(sum, p1, p2, +1);
                    // small multipliers: repeated addition keeps everything inside the wide type
                    lhs.len = 0; lhs.neg = 0; rhs.len = 0; rhs.neg = 0;
                    for (int i = 0; i < a; ++i) { ar.addsub(aw, lhs, d2, +1); lhs = aw; }
                    for (int i = 0; i < bq; ++i) { ar.addsub(bw, rhs, sum, +1); rhs = bw; }
                    do_swap = A::cmp(lhs, rhs) > 0;
                }
            }
            L.sync();  // every lane has read what it needs for the decision
            if (do_swap) { force_reduce = swap_rows(k, max_k); --k; }
            else { force_reduce = 1; ++k; }
        }
        return L.any(ar.ovf) ? -1 : rank_so_far();
    }
};

// One problem, storage carved from `work` (bytes_needed<W>(r, c) bytes).  B: r x c in/out, U: r x r out.
template <int W>
LF_HD inline long bytes_needed(int r, int c) {
    return (long)sizeof(Num<W>) * ((long)r * c + (long)r * r + (long)(r + 1) * (r + 1) + (r + 1)) + (long)sizeof(int) * (r + 4);
}
template <int W, class Lanes>
LF_HD inline int reduce_one(int64_t *B, int64_t *U, int r, int c, void *work) {
    typedef Num<W> N;
    Reducer<W, Lanes> R;
    R.m = r; R.n = c;
    N *p = (N *)work;
    R.B = p; p += (long)r * c;
    R.U = p; p += (long)r * r;
    R.lam = p; p += (long)(r + 1) * (r + 1);
    R.D = p; p += r + 1;
    R.P = (int *)p;
    const int lane = R.L.lane(), width = R.L.width();
    for (int i = lane; i < r; i += width) {
        for (int j = 0; j < c; ++j) Arith<W>::set_i64(R.b(i, j), B[(long)i * c + j]);
        for (int j = 0; j < r; ++j) Arith<W>::set_i64(R.u(i, j), i == j ? 1 : 0);
    }
    for (long i = lane; i < (long)(r + 1) * (r + 1); i += width) { R.lam[i].len = 0; R.lam[i].neg = 0; }
    for (int i = lane; i <= r + 1; i += width) R.P[i] = 0;
    for (int i = lane; i <= r; i += width) { R.D[i].len = 0; R.D[i].neg = 0; }
    R.L.sync();
    if (lane == 0) Arith<W>::set_i64(R.D[0], 1);
    R.L.sync();
    const int rank = R.run(3, 4, 4000000L);
    if (rank < 0) return -1;
    for (int i = lane; i < r; i += width) {
        for (int j = 0; j < c; ++j) B[(long)i * c + j] = Arith<W>::to_i64(R.b(i, j));
        if (U) for (int j = 0; j < r; ++j) U[(long)i * r + j] = Arith<W>::to_i64(R.u(i, j));
    }
    R.L.sync();
    return rank;
}

}  // namespace lllfix
