// blt_api.cpp -- the twelve functions of blt.h (reference: libblt/blt.cc:22-120), same ownership rules:
// every CExpr is owned by the library and freed by blt_free(); callers never free.
#include <fstream>
#include <map>
#include <mutex>
#include <string>
#include <vector>

#include "../../../include/blt.h"
#include "../../../include/codes.h"
#include "context.h"

// The reference's pools are unlocked globals (blt.cc:22-23), which forbids contexts on different threads; here the
// maps are guarded, so distinct contexts may be built, checked and freed concurrently (one context is still
// single-threaded, like every object of the API).
static std::mutex pool_mu;
static std::map<CContext, std::vector<Expr *>> e_pool;
static std::map<CContext, std::vector<SolvState *>> s_pool;

static CExpr keep(CContext ctx, Expr *e) {
    std::lock_guard<std::mutex> lk(pool_mu);
    e_pool[ctx].push_back(e);
    return reinterpret_cast<CExpr>(e);
}

extern "C" {

CContext blt_init(int enable_yices) { return reinterpret_cast<CContext>(new Context(enable_yices)); }

CExpr blt_var(CContext ctx, const char *name) {
    Context *cp = reinterpret_cast<Context *>(ctx);
    return keep(ctx, new Expr(cp->mkVar(name)));
}

CExpr blt_const(CContext ctx, int64 n, int64 d) { return keep(ctx, new Expr(Expr::constant(rat64(n, d)))); }

CExpr blt_add(CContext ctx, const CExpr lhs, const CExpr rhs) {
    if (!lhs || !rhs) return nullptr;  // the reference asserts (disabled by -DNDEBUG, its default)
    return keep(ctx, new Expr(*reinterpret_cast<Expr *>(lhs) + *reinterpret_cast<Expr *>(rhs)));
}

CExpr blt_smul(CContext ctx, int64 n, int64 d, CExpr ce) {
    if (d == 0 || !ce) return nullptr;
    return keep(ctx, new Expr(*reinterpret_cast<Expr *>(ce) * rat64(n, d)));
}

CExpr blt_tof(CContext ctx, int64 s, CExpr ce) {
    if (!ce) return nullptr;
    return keep(ctx, new Expr((*reinterpret_cast<Expr *>(ce) * rat64(s)).rounded()));
}

int blt_assume(CContext ctx, int64 ln, int64 ld, const CExpr ce, int64 un, int64 ud) {
    if (!ce) return BLT_STATUS_ERROR;
    return reinterpret_cast<Context *>(ctx)->assume(rat64(ln, ld), *reinterpret_cast<Expr *>(ce), rat64(un, ud));
}

int blt_check(CContext ctx) { return reinterpret_cast<Context *>(ctx)->check(); }

int blt_check_batch(CContext *ctxs, int count, int *results) {
    return check_batch(reinterpret_cast<Context **>(ctxs), count, results);
}

int blt_model(const CContext ctx, const char *name, int64 *val) {
    return reinterpret_cast<Context *>(ctx)->get_model(std::string(name), val);
}

CSolvState blt_save(CContext ctx) {
    SolvState *s = new SolvState(reinterpret_cast<Context *>(ctx)->save_state());
    std::lock_guard<std::mutex> lk(pool_mu);
    s_pool[ctx].push_back(s);  // the reference leaks these (blt.cc:96-99); pooling them is invisible to callers
    return reinterpret_cast<CSolvState>(s);
}

int blt_backtrack(CContext ctx, CSolvState sav) {
    return reinterpret_cast<Context *>(ctx)->backtrack(*reinterpret_cast<SolvState *>(sav));
}

void blt_free(CContext ctx) {
    std::vector<Expr *> es;
    std::vector<SolvState *> ss;
    {
        std::lock_guard<std::mutex> lk(pool_mu);
        es.swap(e_pool[ctx]); e_pool.erase(ctx);
        ss.swap(s_pool[ctx]); s_pool.erase(ctx);
    }
    for (Expr *e : es) delete e;
    for (SolvState *s : ss) delete s;
    delete reinterpret_cast<Context *>(ctx);
}

// Extension (include/blt.h): solve_raw() of libblt/test_regression.cc:36-61 on two Boost-format matrix files.
int blt_b200_solve_raw_files(const char *lat_path, const char *con_path, int sound, int64 *x, int x_cap, int *rows, uint64_t *niter,
                             uint64_t *nprune) {
    if (!lat_path || !con_path) return BLT_STATUS_INPUT_ERROR;
    std::ifstream fl(lat_path), fc(con_path);
    std::vector<int64> M, C, sol;
    size_t r = 0, c = 0, cr = 0, cc = 0;
    if (read_lattice(fl, M, r, c) != 0 || read_lattice(fc, C, cr, cc) != 0) return BLT_STATUS_INPUT_ERROR;
    if (cc != 2 || cr != c) return BLT_STATUS_INPUT_ERROR;  // closest.cc:173-177
    CheckStats st;
    const int rc = solve_raw(M, r, c, C, sound, &sol, &st);
    if (rows) *rows = (int)r;
    if (niter) *niter = st.niter;
    if (nprune) *nprune = st.nprune;
    if (x && (rc == BLT_CHECK_SAT || rc == BLT_CHECK_UNSAT))
        for (size_t i = 0; i < r && (int)i < x_cap; ++i) x[i] = sol[i];
    return rc;
}

// Not in blt.h (the reference keeps SolverStats private): counters of the context's last blt_check().
int blt_b200_last_stats(CContext ctx, uint64_t *niter, uint64_t *nprune, int *rank, double *lll_ms, double *search_ms) {
    if (!ctx) return BLT_STATUS_INPUT_ERROR;
    const CheckStats &s = reinterpret_cast<Context *>(ctx)->last;
    if (niter) *niter = s.niter;
    if (nprune) *nprune = s.nprune;
    if (rank) *rank = s.rank;
    if (lll_ms) *lll_ms = s.lll_ms;
    if (search_ms) *search_ms = s.search_ms;
    return BLT_STATUS_OK;
}
}
