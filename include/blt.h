/* blt.h -- the BLT C API, unchanged (reference: libblt/blt.h:16-134).  libblt.so of this repository exports these
 * twelve functions with the reference's signatures, ownership rules and return codes; behind blt_check() the
 * reduction and the search run through include/bltcu.h on the GPU. */
#ifndef BLT_H
#define BLT_H

#include <stdint.h>

#define __BLT_VERSION "1.0.0"

/* machine word size, as the reference defines it (libblt/blt.h:18-26); its C client test keys on ENVIRONMENT_64 */
#if INTPTR_MAX == INT32_MAX
#define ENVIRONMENT_32
#elif INTPTR_MAX == INT64_MAX
#define ENVIRONMENT_64
#else
#error "Platform is not 32 or 64-bit."
#endif

typedef int64_t int64;

typedef struct Context_internal_ *CContext;     /* a problem context */
typedef struct Expr_internal_ *CExpr;           /* an affine expression, owned by the library */
typedef struct SolvState_internal_ *CSolvState; /* a bookmark in the assumption stack */

#ifdef __cplusplus
extern "C" {
#endif

/* libblt/blt.cc:25-28.  enable_yices != 0 asks for sound mode (run_blt -y).  The reference then consults Yices on every
 * LP prune (closest.cc:258-267); this build needs no libyices.so: the GPU search prunes a layer only when the LP's dual
 * bound is clear of the radius by a margin far above its rounding error, and keeps it otherwise (bltcu_opts.sound). */
CContext blt_init(int enable_yices);
/* libblt/blt.cc:30-35: a fresh (or the existing) variable of that name as an expression. */
CExpr blt_var(CContext, const char *);
/* libblt/blt.cc:37-42: the constant n/d. */
CExpr blt_const(CContext, int64 n, int64 d);
/* libblt/blt.cc:44-52 */
CExpr blt_add(CContext, const CExpr, const CExpr);
/* libblt/blt.cc:54-65: (n/d) * expr; NULL when d == 0. */
CExpr blt_smul(CContext, int64 n, int64 d, CExpr expr);
/* libblt/blt.cc:67-74: round(s * expr) coefficient-wise. */
CExpr blt_tof(CContext, int64 s, CExpr expr);
/* libblt/blt.cc:76-83: push n1/d1 <= expr <= n2/d2.  BLT_STATUS_OK, BLT_STATUS_ERROR, or BLT_CHECK_UNSAT when the
 * bounds cross. */
int blt_assume(CContext, int64 n1, int64 d1, const CExpr expr, int64 n2, int64 d2);
/* libblt/blt.cc:85-88: BLT_CHECK_SAT / BLT_CHECK_UNSAT / BLT_STATUS_* */
int blt_check(CContext);
/* libblt/blt.cc:90-94 */
int blt_model(const CContext, const char *name, int64 *x0);
/* libblt/blt.cc:96-99 */
CSolvState blt_save(CContext);
/* libblt/blt.cc:101-105 */
int blt_backtrack(CContext, CSolvState);
/* libblt/blt.cc:109-120 */
void blt_free(CContext);

/* Extension, no counterpart in the reference (its globals forbid many checks in flight, libblt/blt.cc:22-23):
 * blt_check() for `count` contexts at once.  Contexts of equal shape share one batched LLL and one batched search on
 * the GPU (bltcu_lll_batch, bltcu_closest_batch); results[i] receives what blt_check(ctxs[i]) returns, and
 * blt_model() works on every context afterwards as after blt_check().  Returns BLT_STATUS_OK unless the arguments
 * are unusable. */
int blt_check_batch(CContext *ctxs, int count, int *results);

/* Extension, no counterpart in the reference (SolverStats stays private there, libblt/context.cc:235-238, and run_blt
 * prints XXX for the counters, run_blt.cc:173-175): counters and phase times of the context's last blt_check().
 * niter/nprune are SolverStats of libblt/closest.h:14-17, rank is what LLL() returned, lll_ms and search_ms are wall
 * times of the two seams.  Any pointer may be NULL. */
int blt_b200_last_stats(CContext ctx, uint64_t *niter, uint64_t *nprune, int *rank, double *lll_ms, double *search_ms);

/* Extension, no counterpart in blt.h: the reference's own regression driver solve_raw() (libblt/test_regression.cc:36-61)
 * as a library call, for the legacy lattice/constraint pairs of test/data (test_95, test_50, test_50_8, test_4x4), whose
 * .bilp forms cannot pass stretch_to_cube (73-bit lcm).  Reads two matrices in Boost ublas text format (read_lattice,
 * libblt/parser.cc:43-55): lat_path r x c generating rows, con_path c x 2 bounds; exact LLL, closest(), exact check; no
 * stretch.  x (x_cap entries, may be NULL) receives the coefficients of the point with respect to the rows of the file.
 * Returns BLT_CHECK_SAT / BLT_CHECK_UNSAT or a BLT_STATUS_* code.  run_blt --latcon <lat> <con> calls it. */
int blt_b200_solve_raw_files(const char *lat_path, const char *con_path, int sound, int64 *x, int x_cap, int *rows, uint64_t *niter,
                             uint64_t *nprune);

#ifdef __cplusplus
}
#endif
#endif
