/*
 * bltcu.h -- C ABI of libbltcu.so, the CUDA (sm_100a) implementation of BLT's lattice-search hot path.
 *
 * These are the two seams that BLT's Context::check() crosses (reference paths relative to /root/reference):
 *     rank = LLL(L, U)                                  libblt/context.cc:176   -> bltcu_lll()
 *     rc   = closest(newL, C, yok, presoln_b, stats)    libblt/context.cc:236   -> bltcu_closest()
 * Plain pointers and sizes only; caller-owned host buffers unless a function says "device"; blocking; no
 * exceptions cross the boundary.  Return values are BLT's own codes (libblt/codes.h:5-24).
 *
 * There is no CPU fallback anywhere behind this header: without a usable CUDA device every compute entry point
 * returns BLT_STATUS_ERROR and prints the CUDA error on stderr.
 */
#ifndef BLTCU_H
#define BLTCU_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* libblt/codes.h:5-24 (values are fixed: the Haskell binding hard-codes them, src/BLT/Internal.hsc:40-45) */
#define BLTCU_STATUS_OK 0
#define BLTCU_STATUS_ERROR 1
#define BLTCU_STATUS_LP_FAIL 2
#define BLTCU_STATUS_INPUT_ERROR 3

#define BLTCU_MAX_DIM 96 /* largest n (basis rows) and m (coordinates) the search kernel accepts */

/* Options of one search.  Zero-initialise, then set what you need. */
typedef struct bltcu_opts {
    int device;          /* CUDA device ordinal */
    int rank, world;     /* this process searches the subtrees  i % world == rank  of the static split (world<=1: all) */
    int split_depth;     /* static split: expand this many top levels into subtrees first (0 = none; needed for world>1) */
    int donate_min_level;/* dynamic split: only subtrees rooted at a level >= this are handed to idle warps (0 = default) */
    int no_early_exit;   /* 1 = keep enumerating after a solution was found (exhaustive count) */
    int warps_per_cta;   /* 0 = as many as shared memory allows */
    int ctas;            /* 0 = one per SM */
    int timeout_ms;      /* device-side watchdog: give up with BLT_STATUS_ERROR after this long (0 = env
                            BLTCU_TIMEOUT_MS, else none) */
    int eager;           /* explored children a warp collects on one level before it descends into the first (0 = default;
                            1 = each warp descends at once, the reference's depth-first order inside a subtree) */
    int sound;           /* 1 = sound pruning (the reference's -y mode, closest.cc:258-267, without Yices): a layer is pruned
                            only when the LP's dual bound clears the radius by a margin far above its floating-point
                            error; layers inside that margin are kept (counted in bltcu_stats.inband) */
    int gpus;            /* bltcu_closest_ex() with world <= 1: search on this many devices (device, device+1, ...) from
                            this one process, a host thread and a share of the subtrees each, results merged before the
                            call returns (0 = environment variable BLT_GPUS, else 1) */
    void *stop_flag;     /* world > 1: host address of 16 ints shared by all ranks of the search (any host memory the
                            ranks share: an array when they are threads, a POSIX shared-memory page when they are
                            processes), zeroed by the caller before the search.  Word 0 is the stop word: the rank that
                            finds a point sets it, every other rank stops within a millisecond, as closest() returns at
                            the first point (closest.cc:300-312).  Word 1 is a counter from which the ranks take the
                            last eighth of the split's subtrees in small chunks, whoever has idle warps (load balance
                            between GPUs; the rest is dealt i % world).  GPUs never access the block; each rank's host
                            thread relays between it and its device while the kernel runs.  NULL = static deal only,
                            each rank stops only on its own find.
                            A LARGER block (its size in bytes in word 2, words 3..1023 zeroed, word 7 an epoch number that
                            changes from search to search; at most 16 ranks) also holds one mailbox per rank behind the
                            first 4096 bytes; with BLTCU_SHARE=1 in the environment the ranks then pass subtree records
                            to each other through it while they search (a rank that runs dry asks, the warps of a busy
                            one hand over the oldest records of their stacks, every copy made by the host threads), and
                            no part of the split is dealt on demand.  Results do not depend on it.
                            blt_b200.dist.StopFlag and bltcu_closest_ex(gpus > 1) make such a block */
} bltcu_opts;

/* Statistics of one search.  niter/nprune are SolverStats of libblt/closest.h:14-17. */
typedef struct bltcu_stats {
    uint64_t niter, nprune;
    uint64_t nodes;        /* layer-distance solves = niter + nprune + 2 on a complete single-rank run */
    uint64_t pivots, quick_prunes, refactors, cold_starts;
    uint64_t donated;      /* subtrees handed over through the global queue */
    uint64_t hist[BLTCU_MAX_DIM + 1]; /* hist[kk] = solves with kk free coefficients (roofline input) */
    double kernel_ms;      /* CUDA-event time of the search kernel(s) */
    double total_ms;       /* wall time of the call including copies */
    int found;             /* 1 = a lattice point inside the box was found */
    int launches;          /* kernels launched by this call */
    int warps, ctas;       /* launch shape */
    uint64_t scratch_bytes;/* global scratch used (per-warp stacks + subtree pool) */
    double busy_frac;      /* share of warp time spent expanding records (the rest waits for work) */
    /* Position of the returned candidate in the reference's depth-first order when no point was found: the ordinal of
     * the sibling taken on every level, top level first.  A caller that splits one search over several ranks returns
     * the `sol` of the rank whose key is lexicographically last (closest.cc:300-303 keeps the last recorded leaf). */
    uint16_t best_key[BLTCU_MAX_DIM];
    int best_valid;        /* 1 = a leaf was recorded by this rank (best_key and sol are meaningful) */
    int stopped;           /* 1 = this rank ended early because another rank set the stop flag */
    int gpus;              /* devices that took part (bltcu_closest_ex with opts.gpus / BLT_GPUS) */
    uint64_t inband;       /* sound mode: layers kept only because of the safety margin */
} bltcu_stats;

/* closest(): libblt/closest.h:50-54, libblt/closest.cc:166-327.
 * lat: n x m row-major (rows = basis vectors), con: m x 2 (lower, upper).  sol: n coefficients w.r.t. lat (the
 * point found, or the last recorded candidate `uhat`, exactly as the reference returns).  Returns
 * BLT_STATUS_OK / LP_FAIL / INPUT_ERROR / ERROR.  Runs on device 0 with default options. */
int bltcu_closest(const int64_t *lat, const int64_t *con, int n, int m, int64_t *sol, uint64_t *niter,
                  uint64_t *nprune);

/* Same seam with options and statistics (opts and stats may be NULL). */
int bltcu_closest_ex(const int64_t *lat, const int64_t *con, int n, int m, const bltcu_opts *opts, int64_t *sol,
                     bltcu_stats *stats);

/* Same search with lat/con already resident in device memory (d_lat: n*m int64, d_con: 2*m int64, on
 * opts->device).  sol and stats stay host pointers.  Used to time the search without the input copies. */
int bltcu_closest_dev(const int64_t *d_lat, const int64_t *d_con, int n, int m, const bltcu_opts *opts,
                      int64_t *sol, bltcu_stats *stats);

/* Batch form of the same seam: `count` independent problems of one shape.
 * lat: count x n x m, con: count x m x 2, sol: count x n, status/found/niter/nprune: count entries each (niter, nprune,
 * found and stats may be NULL).  status[i] is what bltcu_closest() would return for problem i alone.  Each problem is
 * searched in the reference's own depth-first order (closest.cc:208-318), so for SAT problems too sol, niter and
 * nprune are those of the sequential search: a problem starts on one warp; warps that run out of problems take over
 * subtrees of the ones still running (counts do not depend on who searches what), and a problem that was shared and
 * turns out to have a point is searched once more by one warp alone (stats->donated = problems that were shared).  With opts->world > 1 only problems
 * [count*rank/world, count*(rank+1)/world) are searched and written.  stats->found = number of problems with a point;
 * the other counters are sums.  Returns BLT_STATUS_OK unless the call itself failed. */
int bltcu_closest_batch(const int64_t *lat, const int64_t *con, int count, int n, int m, const bltcu_opts *opts,
                        int64_t *sol, int *status, uint64_t *niter, uint64_t *nprune, int *found, bltcu_stats *stats);

/* LLL(): libblt/ntl_wrapper.h:20, libblt/ntl_wrapper.cc:37-52 (exact integral LLL, delta = 3/4, NTL::LLL(det2,B,U)).
 * B: r x c row-major, reduced in place: zero rows first, the `rank` reduced basis rows last.  U: r x r output with
 * U * B_in = B_out (may be NULL).  Returns BLT_STATUS_OK or BLT_STATUS_INPUT_ERROR.  Host routine (a single reduction
 * is one chain of dependent steps): first the integral algorithm's row operations with its decisions taken in fp64
 * against a margin and the operations themselves in exact int64 (60 x 60: a few ms); a decision inside the margin,
 * dependent rows or numbers beyond 2^53 hand over to multiprecision passes that perform the same operations. */
int bltcu_lll(int64_t *B, int64_t *U, int r, int c, int *rank);

/* Batched LLL() for many small problems of one shape (configs 2 and 5 are reduction-bound).
 * B: count x r x c, reduced in place; U: count x r x r out (may be NULL); rank: count.  One GPU warp reduces one problem,
 * the same operations in the same order as bltcu_lll(), so basis and U are identical to what a loop over bltcu_lll()
 * gives: first the fp64-decision pass (see bltcu_lll), then, for what it deferred (stats->donated), the host routine
 * on all cores when it is a handful or a second kernel with 512-bit exact integers when it is a crowd (dependent
 * generating sets always are); numbers beyond 512 bits end in the host routine (stats->found = problems reduced on the
 * host).  Returns BLT_STATUS_OK / INPUT_ERROR / ERROR. */
int bltcu_lll_batch(int64_t *B, int64_t *U, int count, int r, int c, int *rank, int device, bltcu_stats *stats);

/* yprune(): libblt/yices_wrapper.h:26-34, libblt/yices_wrapper.cc:47-157 -- the exact decision the reference's sound
 * mode asks of Yices: does the layer with coefficients u[k..n) fixed meet the cube?  Decided on the host by
 * Fourier-Motzkin elimination in exact 128-bit integers (no Yices).  *decision = BLT_CHECK_SAT (keep) / BLT_CHECK_UNSAT
 * (prune); returns LP_OK.  Layers with more than BLTCU_YPRUNE_MAX_K free coefficients, or whose elimination outgrows
 * its limits, give BLT_CHECK_UNKNOWN and LP_ERROR like a solver failure in the reference. */
#define BLTCU_YPRUNE_MAX_K 6
int bltcu_yprune(const int64_t *lat, int n, int m, const int64_t *u, int k, const int64_t *con, int *decision);

/* check_soln(): libblt/closest.cc:330-340, evaluated on the device (int64, wraps like the reference). */
int bltcu_check_soln(const int64_t *lat, const int64_t *con, int n, int m, const int64_t *u, int *ok);

/* infdist(): libblt/glpk_wrapper.h:26-34, one cold layer-distance solve on the device (test seam for the
 * reference's LP known answers, libblt/test_glpk.cc).  B: n x m doubles, u: n, p: m, x: k outputs. */
int bltcu_infdist(const double *B, int n, int m, const int64_t *u, int k, const double *p, double *x,
                  double *dist);

/* No counterpart in the reference.  closest() for callers that must not allocate: a plugin inside a host
 * application that owns the GPU heap, a real-time loop that cannot afford a device-synchronising cudaMalloc, a
 * process that has handed all free HBM to its own arena.  Calls no cudaMalloc/cudaFree and creates no stream or
 * event; all kernel scratch lives in one fixed 20 MiB array inside the library image (held by every GPU that loads
 * the library until the process exits).  n, m <= BLTCU_NOALLOC_MAX_DIM, otherwise BLT_STATUS_INPUT_ERROR and a
 * message on stderr.  The first call on a device resolves the array's device address: make it before any stream
 * capture.  All calls share the array, so calls that overlap on one GPU race: serialise them.  Pointers are host
 * pointers with the meaning they have in bltcu_closest(); found receives 1 if a lattice point was found.
 * Each warp's record stack inside the array has the worst-case size of a depth-first path, and the number of CTAs is
 * reduced until everything fits (about 90 at 20 x 20), so a hard instance cannot run out of room.  There is no
 * timeout argument: the environment variable BLTCU_TIMEOUT_MS is the only watchdog of this entry point. */
#define BLTCU_NOALLOC_MAX_DIM 20
int bltcu_closest_noalloc(const int64_t *lat, const int64_t *con, int n, int m, int device, int64_t *sol,
                          uint64_t *niter, uint64_t *nprune, int *found);

/* Measurement helper (no counterpart in the reference): FP64 FMA throughput of `device` in TFLOP/s, best of
 * `reps` launches.  bench.py uses it as the roofline denominator of the search kernel. */
int bltcu_fp64_peak(int device, int reps, double *tflops);
/* The same for the INT32 multiply-add pipe (Tera-operations per second, 2 per IMAD): the second roofline north_star names. */
int bltcu_int32_peak(int device, int reps, double *tiops);

/* Releases the cached device workspace of every device this process used. */
void bltcu_release(void);

/* Version string of the library. */
const char *bltcu_version(void);

#ifdef __cplusplus
}
#endif
#endif
