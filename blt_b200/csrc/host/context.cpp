// context.cpp -- Context::check() and friends (reference: libblt/context.cc, libblt/expr.cc, libblt/parser.cc).
// Plain C++17, no Boost.  The two hot seams go to the GPU library: bltcu_lll() and bltcu_closest_ex().
#include "context.h"

#include <ctype.h>
#include <math.h>
#include <stdlib.h>

#include <algorithm>
#include <chrono>
#include <cstring>
#include <iostream>
#include <sstream>

#include "../../../include/bltcu.h"
#include "../../../include/codes.h"

#define ERROR(x) do { std::cerr << "Error [" << __FILE__ << ":" << __LINE__ << "]: " << x << std::endl; } while (0)
#define WARN(x) do { std::cerr << "Warning: " << x << std::endl; } while (0)

// ---- integer helpers (boost::math::gcd / lcm semantics on int64; products wrap like the reference's) ----------
static int64 gcd64(int64 a, int64 b) {
    uint64_t x = a < 0 ? (uint64_t)0 - (uint64_t)a : (uint64_t)a, y = b < 0 ? (uint64_t)0 - (uint64_t)b : (uint64_t)b;
    while (y) { const uint64_t t = x % y; x = y; y = t; }
    return (int64)x;
}
static int64 lcm64(int64 a, int64 b) {
    if (a == 0 || b == 0) return 0;
    const int64 g = gcd64(a, b);
    const uint64_t x = a < 0 ? (uint64_t)0 - (uint64_t)a : (uint64_t)a, y = b < 0 ? (uint64_t)0 - (uint64_t)b : (uint64_t)b;
    return (int64)((x / (uint64_t)g) * y);  // wraps on overflow; callers test for it the way the reference does
}

// ---- rat64 ----------------------------------------------------------------------------------------------------
static rat64 make_rat(__int128 n, __int128 d) {
    rat64 r;
    if (d == 0) { r.num = 0; r.den = 1; return r; }  // the reference throws bad_rational here
    if (d < 0) { n = -n; d = -d; }
    __int128 a = n < 0 ? -n : n, b = d;
    while (b) { const __int128 t = a % b; a = b; b = t; }
    if (a > 1) { n /= a; d /= a; }
    r.num = (int64)n; r.den = (int64)d;
    return r;
}
rat64::rat64(int64 n, int64 d) { *this = make_rat(n, d); }
rat64 operator+(const rat64 &a, const rat64 &b) { return make_rat((__int128)a.num * b.den + (__int128)b.num * a.den, (__int128)a.den * b.den); }
rat64 operator-(const rat64 &a, const rat64 &b) { return make_rat((__int128)a.num * b.den - (__int128)b.num * a.den, (__int128)a.den * b.den); }
rat64 operator*(const rat64 &a, const rat64 &b) { return make_rat((__int128)a.num * b.num, (__int128)a.den * b.den); }
bool operator<=(const rat64 &a, const rat64 &b) { return (__int128)a.num * b.den <= (__int128)b.num * a.den; }
bool operator==(const rat64 &a, const rat64 &b) { return a.num == b.num && a.den == b.den; }

// ---- Expr (libblt/expr.cc) ---------------------------------------------------------------------------------------
rat64 Expr::operator[](int idx) const {
    auto it = coeff_map_.find(idx);
    return it == coeff_map_.end() ? rat64(0) : it->second;
}
Expr &Expr::operator+=(const Expr &rhs) {
    for (auto &kv : rhs.coeff_map_) {
        coeff_map_[kv.first] = (*this)[kv.first] + kv.second;
        indices_.insert(kv.first);
    }
    const_ = const_ + rhs.const_;
    return *this;
}
Expr &Expr::operator*=(const rat64 &s) {
    for (auto &kv : coeff_map_) kv.second = kv.second * s;
    const_ = const_ * s;
    return *this;
}
Expr Expr::rounded() const {  // expr.cc:120-136: coefficients with denominator 1 are kept, others go through double
    Expr r = *this;
    for (int idx : indices_) {
        const rat64 c = r[idx];
        if (c.denominator() == 1) continue;
        r.coeff_map_[idx] = rat64((int64)::round(c.to_double()));
    }
    r.const_ = rat64((int64)::round(r.const_.to_double()));
    return r;
}

// ---- Context ----------------------------------------------------------------------------------------------------
int Context::nvars() const { return (int)std::count(mask.begin(), mask.end(), true); }

Expr Context::mkVar(const std::string &name) {  // context.cc:51-63
    int idx;
    auto it = std::find(names.begin(), names.end(), name);
    if (it == names.end()) {
        idx = (int)names.size();
        names.push_back(name);
        indices[name] = idx;
        mask.resize(names.size());
    } else idx = indices[name];
    return Expr(idx);
}
bool Context::varActive(const std::string &name) const {
    auto it = indices.find(name);
    return it != indices.end() && mask[it->second];
}
int Context::assume(const rat64 &lower, const Expr &expr, const rat64 &upper) {  // context.cc:113-131
    if (lower <= upper) {
        try {
            assumptions.push_back(Atom{lower, upper, expr});
            for (int i : expr.indices()) mask[i] = true;
            return BLT_STATUS_OK;
        } catch (std::bad_alloc &e) {
            ERROR("Context::assume: Out of memory" << std::endl << e.what());
            return BLT_STATUS_ERROR;
        }
    }
    return BLT_CHECK_UNSAT;
}
void Context::update_mask() {  // context.cc:379-393
    mask.assign(names.size(), false);
    for (auto &a : assumptions) for (int vi : a.expr.indices()) mask[vi] = true;
}
int Context::backtrack(SolvState st) {  // context.cc:318-330
    if (st.pos <= assumptions.size()) {
        assumptions.resize(st.pos);
        update_mask();
        return BLT_STATUS_OK;
    }
    return BLT_STATUS_ERROR;
}
int Context::get_model(const std::string &name, int64 *val) const {  // context.cc:286-293
    auto it = model.find(name);
    if (it != model.end()) { *val = it->second; return BLT_STATUS_OK; }
    return BLT_STATUS_INPUT_ERROR;
}

int convert_to_integral(const std::vector<rat64> &L0, const std::vector<rat64> &C0, size_t nrows, size_t ncols,
                        std::vector<int64> &L, std::vector<int64> &C) {  // context.cc:405-444
    L.assign(nrows * ncols, 0);
    C.assign(ncols * 2, 0);
    for (size_t c = 0; c < ncols; ++c) {
        int64 mult = 1;
        for (size_t r = 0; r < nrows; ++r) {
            const int64 t = lcm64(mult, L0[r * ncols + c].denominator());
            if (t < mult) { ERROR("integer overflow detected!"); return BLT_STATUS_ERROR; }
            mult = t;
        }
        mult = lcm64(mult, C0[2 * c].denominator());
        mult = lcm64(mult, C0[2 * c + 1].denominator());
        for (size_t r = 0; r < nrows; ++r) {
            const rat64 &q = L0[r * ncols + c];
            L[r * ncols + c] = (int64)((uint64_t)(mult / q.denominator()) * (uint64_t)q.numerator());
        }
        for (int t = 0; t < 2; ++t) {
            const rat64 &q = C0[2 * c + t];
            C[2 * c + t] = (int64)((uint64_t)(mult / q.denominator()) * (uint64_t)q.numerator());
        }
    }
    return BLT_STATUS_OK;
}

// BLT_CHECKED_STRETCH=1: the same scaling with every product checked in 128 bits; an lcm or a scaled entry beyond int64
// is BLT_STATUS_INPUT_ERROR with a message instead of a silent wrap (the reference's heuristics at context.cc:470-474
// miss some overflows: test_95.bilp needs a 73-bit lcm).  Off by default: wrap-faithful like the reference.
static int stretch_to_cube_checked(std::vector<int64> &L, std::vector<int64> &C, size_t nrows, size_t ncols) {
    typedef __int128 i128;
    const i128 lim = (i128)INT64_MAX;
    i128 l = 1;
    for (size_t i = 0; i < ncols; ++i) {
        const i128 r = (i128)C[2 * i + 1] - (i128)C[2 * i];
        if (r <= 0) { ERROR("upper and lower bound coincide or conflict"); return BLT_STATUS_INPUT_ERROR; }
        i128 a = l, b = r;
        while (b) { const i128 t = a % b; a = b; b = t; }
        l = l / a * r;
        if (l > lim) { ERROR("stretch_to_cube: the common width (lcm of the row widths) does not fit 64 bits"); return BLT_STATUS_INPUT_ERROR; }
    }
    for (size_t i = 0; i < ncols; ++i) {
        const i128 mult = l / ((i128)C[2 * i + 1] - (i128)C[2 * i]);
        const i128 lo = (i128)C[2 * i] * mult, hi = (i128)C[2 * i + 1] * mult;
        if (lo > lim || lo < -lim || hi > lim || hi < -lim) { ERROR("stretch_to_cube: a scaled bound does not fit 64 bits"); return BLT_STATUS_INPUT_ERROR; }
        C[2 * i] = (int64)lo; C[2 * i + 1] = (int64)hi;
        for (size_t j = 0; j < nrows; ++j) {
            const i128 v = (i128)L[j * ncols + i] * mult;
            if (v > lim || v < -lim) { ERROR("stretch_to_cube: a scaled coefficient does not fit 64 bits"); return BLT_STATUS_INPUT_ERROR; }
            L[j * ncols + i] = (int64)v;
        }
    }
    return BLT_STATUS_OK;
}

int stretch_to_cube(std::vector<int64> &L, std::vector<int64> &C, size_t nrows, size_t ncols) {  // context.cc:455-501
    if (const char *e = getenv("BLT_CHECKED_STRETCH")) if (e[0] == '1') return stretch_to_cube_checked(L, C, nrows, ncols);
    int64 l = 1;
    for (size_t i = 0; i < ncols; ++i) {
        const int64 r = (int64)((uint64_t)C[2 * i + 1] - (uint64_t)C[2 * i]);
        if (r <= 0) { ERROR("upper and lower bound coincide or conflict"); return BLT_STATUS_INPUT_ERROR; }
        const int64 tmpl = lcm64(l, r);
        if (tmpl == 0 || tmpl % l != 0 || tmpl % r != 0) { ERROR("integer overflow in lcm detected"); return BLT_STATUS_INPUT_ERROR; }
        l = tmpl;
    }
    for (size_t i = 0; i < ncols; ++i) {
        const int64 mult = l / (C[2 * i + 1] - C[2 * i]);
        C[2 * i] = (int64)((uint64_t)C[2 * i] * (uint64_t)mult);
        C[2 * i + 1] = (int64)((uint64_t)C[2 * i + 1] * (uint64_t)mult);
        for (size_t j = 0; j < nrows; ++j) L[j * ncols + i] = (int64)((uint64_t)L[j * ncols + i] * (uint64_t)mult);
    }
    return BLT_STATUS_OK;
}

bool check_soln_host(const std::vector<int64> &lat, const std::vector<int64> &con, size_t n, size_t m,
                     const std::vector<int64> &u) {  // closest.cc:330-340
    for (size_t j = 0; j < m; ++j) {
        uint64_t z = 0;
        for (size_t i = 0; i < n; ++i) z += (uint64_t)u[i] * (uint64_t)lat[i * m + j];
        const int64 zz = (int64)z;
        if (con[2 * j] > zz || zz > con[2 * j + 1]) return false;
    }
    return true;
}

// Context::check() in three parts, so that the batch front-end below can run the two GPU seams for many contexts at
// once: prepare (context.cc:139-174), reduced (context.cc:176-208), finish (context.cc:246-279).
int Context::prepare(Prepared &p) {
    p.m = (size_t)nvars();          // lattice vectors (rows) = active variables
    p.n = assumptions.size();       // constraints = dimension of the ambient space
    const size_t m = p.m, n = p.n;
    if (n == 0 || m == 0) {
        WARN("Context::check: zero-dimensional input lattice or constraint set");
        return BLT_STATUS_INPUT_ERROR;
    }
    // compute_lattice / compute_constr, context.cc:336-377
    std::vector<rat64> L0(m * n), C0(n * 2);
    for (size_t col = 0; col < n; ++col) {
        const Expr &e = assumptions[col].expr;
        size_t row = 0;
        for (size_t idx = 0; idx < mask.size(); ++idx)
            if (mask[idx]) L0[(row++) * n + col] = e[(int)idx];
        const rat64 c = e.get_const();
        C0[2 * col] = assumptions[col].lower - c;
        C0[2 * col + 1] = assumptions[col].upper - c;
    }
    int rc = convert_to_integral(L0, C0, m, n, p.L, p.C);
    if (rc != BLT_STATUS_OK) return rc;
    rc = stretch_to_cube(p.L, p.C, m, n);
    if (rc != BLT_STATUS_OK) return rc;
    p.U.assign(m * m, 0);
    return BLT_STATUS_OK;
}
// after rank = LLL(L, U): the reduced basis is the last `rank` rows
int Context::reduced(Prepared &p, int rank) {
    last.rank = rank;
    if (rank == 0) return BLT_STATUS_INPUT_ERROR;
    p.rank = rank;
    p.newL.assign((size_t)rank * p.n, 0);
    for (size_t i = 0; i < (size_t)rank; ++i)
        for (size_t j = 0; j < p.n; ++j) p.newL[i * p.n + j] = p.L[(i + p.m - rank) * p.n + j];
    // context.cc:210-228 loads libyices.so here when sound mode was requested.  This build has no Yices: sound mode is
    // the search kernel's own (bltcu_opts.sound, a prune needs a bound clear of the radius), selected in check().
    p.presoln.assign((size_t)rank, 0);
    return BLT_STATUS_OK;
}
int Context::finish(const Prepared &p, uint64_t niter, uint64_t nprune) {
    const size_t m = p.m, n = p.n, rank = (size_t)p.rank;
    last.niter = niter; last.nprune = nprune;
    // back-substitution soln = [0 .. 0 | presoln_b] * U, context.cc:246-254 (int64 products wrap like ublas prod)
    std::vector<int64> soln(m, 0);
    for (size_t j = 0; j < m; ++j) {
        uint64_t acc = 0;
        for (size_t i = m - rank; i < m; ++i) acc += (uint64_t)p.presoln[i - (m - rank)] * (uint64_t)p.U[i * m + j];
        soln[j] = (int64)acc;
    }
    model.clear();
    for (size_t i = 0, j = 0; i < mask.size(); ++i)
        if (mask[i]) model[names[i]] = soln[j++];
    return check_soln_host(p.newL, p.C, rank, n, p.presoln) ? BLT_CHECK_SAT : BLT_CHECK_UNSAT;  // context.cc:276-279
}

int Context::check() {  // context.cc:139-280
    using clk = std::chrono::steady_clock;
    Prepared p;
    int rc = prepare(p);
    if (rc != BLT_STATUS_OK) return rc;
    // reduction seam: rank = LLL(L, U), context.cc:176
    int rank = 0;
    auto t0 = clk::now();
    rc = bltcu_lll(p.L.data(), p.U.data(), (int)p.m, (int)p.n, &rank);
    last.lll_ms = std::chrono::duration<double, std::milli>(clk::now() - t0).count();
    if (rc != BLT_STATUS_OK) return rc;
    rc = reduced(p, rank);
    if (rc != BLT_STATUS_OK) return rc;
    // search seam: closest(newL, C, yok, presoln_b, stats), context.cc:236
    bltcu_stats st;
    bltcu_opts opts;
    memset(&opts, 0, sizeof opts);
    opts.sound = enable_yices_ ? 1 : 0;  // blt_init(enable_yices), run_blt -y
    t0 = clk::now();
    rc = bltcu_closest_ex(p.newL.data(), p.C.data(), rank, (int)p.n, &opts, p.presoln.data(), &st);
    last.search_ms = std::chrono::duration<double, std::milli>(clk::now() - t0).count();
    if (rc != BLT_STATUS_OK) return rc;
    return finish(p, st.niter, st.nprune);
}

// Many contexts at once (no counterpart in the reference): contexts of equal shape share
// one bltcu_lll_batch() call, reduced problems of equal shape one bltcu_closest_batch() call.  results[i] is what
// ctx[i]->check() would return; a model is left in every context that is SAT.
int check_batch(Context **ctx, int count, int *results) {
    if (count < 0 || (count > 0 && (!ctx || !results))) return BLT_STATUS_INPUT_ERROR;
    std::vector<Prepared> P((size_t)count);
    std::vector<char> live((size_t)count, 0);
    for (int i = 0; i < count; ++i) {
        if (!ctx[i]) { results[i] = BLT_STATUS_INPUT_ERROR; continue; }
        if (ctx[i]->sound()) { results[i] = ctx[i]->check(); continue; }  // the batch kernels have no sound mode: one at a time
        results[i] = ctx[i]->prepare(P[i]);
        live[i] = results[i] == BLT_STATUS_OK;
    }
    // ---- reduction, grouped by (m, n) ----
    std::map<std::pair<size_t, size_t>, std::vector<int>> groups;
    for (int i = 0; i < count; ++i) if (live[i]) groups[{P[i].m, P[i].n}].push_back(i);
    for (auto &g : groups) {
        const size_t m = g.first.first, n = g.first.second, k = g.second.size();
        std::vector<int64> B(k * m * n), U(k * m * m);
        std::vector<int> rank(k, 0);
        for (size_t j = 0; j < k; ++j) std::copy(P[g.second[j]].L.begin(), P[g.second[j]].L.end(), B.begin() + j * m * n);
        const int rc = bltcu_lll_batch(B.data(), U.data(), (int)k, (int)m, (int)n, rank.data(), 0, nullptr);
        for (size_t j = 0; j < k; ++j) {
            const int i = g.second[j];
            if (rc != BLT_STATUS_OK) { results[i] = rc; live[i] = 0; continue; }
            std::copy(B.begin() + j * m * n, B.begin() + (j + 1) * m * n, P[i].L.begin());
            std::copy(U.begin() + j * m * m, U.begin() + (j + 1) * m * m, P[i].U.begin());
            const int r2 = ctx[i]->reduced(P[i], rank[j]);
            if (r2 != BLT_STATUS_OK) { results[i] = r2; live[i] = 0; }
        }
    }
    // ---- search, grouped by (rank, n) ----
    groups.clear();
    for (int i = 0; i < count; ++i) if (live[i]) groups[{(size_t)P[i].rank, P[i].n}].push_back(i);
    for (auto &g : groups) {
        const size_t r = g.first.first, n = g.first.second, k = g.second.size();
        std::vector<int64> lat(k * r * n), con(k * n * 2), sol(k * r);
        std::vector<int> status(k, 0), found(k, 0);
        std::vector<uint64_t> niter(k, 0), nprune(k, 0);
        for (size_t j = 0; j < k; ++j) {
            std::copy(P[g.second[j]].newL.begin(), P[g.second[j]].newL.end(), lat.begin() + j * r * n);
            std::copy(P[g.second[j]].C.begin(), P[g.second[j]].C.end(), con.begin() + j * n * 2);
        }
        const int rc = bltcu_closest_batch(lat.data(), con.data(), (int)k, (int)r, (int)n, nullptr, sol.data(), status.data(),
                                           niter.data(), nprune.data(), found.data(), nullptr);
        for (size_t j = 0; j < k; ++j) {
            const int i = g.second[j];
            if (rc != BLT_STATUS_OK) { results[i] = rc; continue; }
            if (status[j] != BLT_STATUS_OK) { results[i] = status[j]; continue; }
            std::copy(sol.begin() + j * r, sol.begin() + (j + 1) * r, P[i].presoln.begin());
            results[i] = ctx[i]->finish(P[i], niter[j], nprune[j]);
        }
    }
    return BLT_STATUS_OK;
}

// ---- Boost ublas matrix text, "[r,c]((a,b,..),(..),..)" (read_lattice, libblt/parser.cc:43-55): the legacy .lat/.con
// files of test/data -------------------------------------------------------------------------------------------------
int read_lattice(std::istream &in, std::vector<int64> &M, size_t &nr, size_t &nc) {
    if (!in) { ERROR("could not open input stream for parsing"); return 1; }
    auto skip = [&]() { while (in && isspace(in.peek())) in.get(); };
    auto expect = [&](char c) { skip(); if (in.peek() != c) return false; in.get(); return true; };
    long long r = 0, c = 0;
    if (!expect('[') || !(in >> r) || !expect(',') || !(in >> c) || !expect(']') || r <= 0 || c <= 0 || !expect('(')) {
        ERROR("malformed matrix header, expected [rows,cols](("); return 1;
    }
    nr = (size_t)r; nc = (size_t)c;
    M.assign(nr * nc, 0);
    for (size_t i = 0; i < nr; ++i) {
        if (i && !expect(',')) { ERROR("malformed matrix: missing row separator"); return 1; }
        if (!expect('(')) { ERROR("malformed matrix: missing row"); return 1; }
        for (size_t j = 0; j < nc; ++j) {
            if (j && !expect(',')) { ERROR("malformed matrix: short row"); return 1; }
            long long v = 0;
            if (!(in >> v)) { ERROR("malformed matrix: expecting an integer"); return 1; }
            M[i * nc + j] = (int64)v;
        }
        if (!expect(')')) { ERROR("malformed matrix: long row"); return 1; }
    }
    if (!expect(')')) { ERROR("malformed matrix: missing end"); return 1; }
    return 0;
}

// solve_raw() of libblt/test_regression.cc:36-61: LLL on the lattice as given (no stretch), closest(), exact check.
// lat: r x c generating rows, con: c x 2.  x (r entries, may be NULL) receives the coefficients of the point with
// respect to the ORIGINAL rows (sol * U).  Returns BLT_CHECK_SAT / BLT_CHECK_UNSAT or a status code.
int solve_raw(const std::vector<int64> &lat, size_t r, size_t c, const std::vector<int64> &con, int sound, std::vector<int64> *x,
              CheckStats *stats) {
    using clk = std::chrono::steady_clock;
    if (r == 0 || c == 0 || lat.size() != r * c || con.size() != 2 * c) return BLT_STATUS_INPUT_ERROR;
    std::vector<int64> M = lat, U(r * r, 0);
    int rank = 0;
    auto t0 = clk::now();
    int rc = bltcu_lll(M.data(), U.data(), (int)r, (int)c, &rank);
    const double lll_ms = std::chrono::duration<double, std::milli>(clk::now() - t0).count();
    if (rc != BLT_STATUS_OK) return rc;
    if (rank == 0) return BLT_STATUS_INPUT_ERROR;
    std::vector<int64> newL(M.begin() + (r - (size_t)rank) * c, M.end()), sol((size_t)rank, 0);
    bltcu_stats st;
    bltcu_opts opts;
    memset(&opts, 0, sizeof opts);
    opts.sound = sound ? 1 : 0;
    t0 = clk::now();
    rc = bltcu_closest_ex(newL.data(), con.data(), rank, (int)c, &opts, sol.data(), &st);
    const double search_ms = std::chrono::duration<double, std::milli>(clk::now() - t0).count();
    if (stats) { stats->niter = st.niter; stats->nprune = st.nprune; stats->rank = rank; stats->lll_ms = lll_ms; stats->search_ms = search_ms; }
    if (rc != BLT_STATUS_OK) return rc;
    if (x) {
        x->assign(r, 0);
        for (size_t j = 0; j < r; ++j) {
            uint64_t acc = 0;
            for (size_t i = r - (size_t)rank; i < r; ++i) acc += (uint64_t)sol[i - (r - (size_t)rank)] * (uint64_t)U[i * r + j];
            (*x)[j] = (int64)acc;
        }
    }
    return check_soln_host(newL, con, (size_t)rank, c, sol) ? BLT_CHECK_SAT : BLT_CHECK_UNSAT;
}

// ---- parser (libblt/parser.cc:72-166) -------------------------------------------------------------------------------
static int parse_row(std::istream &in, std::vector<int64> &row) {
    std::vector<std::string> toks;
    std::string w;
    while (in >> w) toks.push_back(w);
    const size_t nt = toks.size();
    if (nt < 5 || toks[1] != "<=" || toks[nt - 2] != "<=") { ERROR("malformed row: " << w); return 1; }
    try {
        row.resize(nt - 2);
        row[0] = std::stoll(toks[0]);
        row[nt - 3] = std::stoll(toks[nt - 1]);
        for (size_t i = 2; i < nt - 2; ++i) row[i - 1] = std::stoll(toks[i]);
    } catch (const std::exception &e) {
        ERROR("parser failed, expecting integer");
        ERROR(e.what());
        return 1;
    }
    return 0;
}
int read_bilp(std::istream &in, std::vector<int64> &L, std::vector<int64> &C, size_t &nr, size_t &nc) {
    if (!in) { ERROR("could not open input stream for parsing"); return 1; }
    nr = nc = 0;
    L.clear(); C.clear();
    std::string line;
    while (std::getline(in, line)) {
        std::vector<int64> row;
        std::istringstream iss(line);
        if (parse_row(iss, row) != 0 || (nc != 0 && row.size() - 2 != nc)) { ERROR("parse error on input line " << nr + 1); return 1; }
        if (nc == 0) nc = row.size() - 2;
        C.push_back(row.front());
        C.push_back(row.back());
        for (size_t i = 0; i < nc; ++i) L.push_back(row[i + 1]);
        ++nr;
    }
    if (nr == 0 || nc == 0) { ERROR("input stream is empty"); return 1; }
    return 0;
}
