// context.h -- host side of the drop-in boundary: BLT's problem model without Boost.
// Mirrors libblt/context.h, libblt/expr.h and libblt/types.h of the reference (same names, same behaviour);
// the reduction and the search are delegated to include/bltcu.h.
#pragma once

#include <stdint.h>

#include <istream>
#include <map>
#include <set>
#include <string>
#include <vector>

typedef int64_t int64;

// rational over int64, always normalised (gcd 1, positive denominator) -- boost::rational<int64_t> in the reference
struct rat64 {
    int64 num, den;
    rat64() : num(0), den(1) {}
    rat64(int64 n) : num(n), den(1) {}
    rat64(int64 n, int64 d);
    int64 numerator() const { return num; }
    int64 denominator() const { return den; }
    double to_double() const { return (double)num / (double)den; }  // boost::rational_cast<double>
};
rat64 operator+(const rat64 &, const rat64 &);
rat64 operator-(const rat64 &, const rat64 &);
rat64 operator*(const rat64 &, const rat64 &);
bool operator<=(const rat64 &, const rat64 &);
bool operator==(const rat64 &, const rat64 &);
inline bool operator!=(const rat64 &a, const rat64 &b) { return !(a == b); }

// affine expression a1 x1 + ... + an xn + C (libblt/expr.h:14-72)
class Expr {
    std::map<int, rat64> coeff_map_;
    rat64 const_;
    std::set<int> indices_;

public:
    Expr() : const_(0) {}
    explicit Expr(int idx) : const_(0) { coeff_map_[idx] = rat64(1); indices_.insert(idx); }
    static Expr constant(rat64 c) { Expr e; e.const_ = c; return e; }
    const std::set<int> &indices() const { return indices_; }
    rat64 operator[](int idx) const;
    rat64 get_const() const { return const_; }
    Expr &operator+=(const Expr &);
    Expr &operator*=(const rat64 &);
    Expr operator+(const Expr &o) const { Expr r = *this; r += o; return r; }
    Expr operator*(const rat64 &s) const { Expr r = *this; r *= s; return r; }
    Expr rounded() const;  // round(): libblt/expr.cc:120-136
};

// lower <= expr <= upper (libblt/context.h:17-31)
struct Atom {
    rat64 lower, upper;
    Expr expr;
};

struct SolvState {
    size_t pos;
};

// Statistics of the last check(); not part of blt.h (the reference keeps them private).
struct CheckStats {
    uint64_t niter = 0, nprune = 0;
    int rank = 0;
    double lll_ms = 0, search_ms = 0;
};

class Context {
    std::vector<std::string> names;
    std::vector<bool> mask;
    std::map<std::string, int> indices;
    std::vector<Atom> assumptions;
    std::map<std::string, int64> model;
    bool enable_yices_;

    void update_mask();

public:
    CheckStats last;

    explicit Context(int flag = 0) : enable_yices_(flag != 0) {}
    bool sound() const { return enable_yices_; }  // blt_init(enable_yices): sound pruning requested
    int nvars() const;
    int nassumpts() const { return (int)assumptions.size(); }
    Expr mkVar(const std::string &);
    bool varActive(const std::string &) const;
    int assume(const rat64 &lower, const Expr &expr, const rat64 &upper);
    SolvState save_state() const { return SolvState{assumptions.size()}; }
    int backtrack(SolvState);
    int get_model(const std::string &, int64 *) const;
    int check();

    // the parts of check(), used one after the other by check() and in bulk by check_batch()
    struct Prepared {
        size_t m = 0, n = 0;
        int rank = 0;
        std::vector<int64> L, C, U, newL, presoln;
    };
    int prepare(Prepared &p);
    int reduced(Prepared &p, int rank);
    int finish(const Prepared &p, uint64_t niter, uint64_t nprune);
};
typedef Context::Prepared Prepared;
int check_batch(Context **ctx, int count, int *results);
int read_lattice(std::istream &in, std::vector<int64> &M, size_t &nr, size_t &nc);   // libblt/parser.cc:43-55
int solve_raw(const std::vector<int64> &lat, size_t r, size_t c, const std::vector<int64> &con, int sound, std::vector<int64> *x,
              CheckStats *stats);                                                        // libblt/test_regression.cc:36-61

// libblt/context.cc:405-501, on row-major vectors: L is nrows x ncols (rows = variables), C is ncols x 2
int convert_to_integral(const std::vector<rat64> &L0, const std::vector<rat64> &C0, size_t nrows, size_t ncols,
                        std::vector<int64> &L, std::vector<int64> &C);
int stretch_to_cube(std::vector<int64> &L, std::vector<int64> &C, size_t nrows, size_t ncols);
bool check_soln_host(const std::vector<int64> &lat, const std::vector<int64> &con, size_t n, size_t m,
                     const std::vector<int64> &u);

// libblt/parser.cc:72-166: rows `l <= a1 ... an <= u`; returns 0 on success
int read_bilp(std::istream &in, std::vector<int64> &L, std::vector<int64> &C, size_t &nrows, size_t &ncols);
