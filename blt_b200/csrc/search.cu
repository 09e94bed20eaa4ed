// search.cu -- sm_100a implementation of BLT's bounded lattice-point search (closest(), libblt/closest.cc:166-327)
// and of its per-node layer-distance solve (infdist(), libblt/glpk_wrapper.cc:33-232).
//
// Execution model (DESIGN.md sections 3-5):
//   * ONE persistent kernel per search, one CTA per SM, up to 12 warps per CTA.  A warp owns one task at a time; its
//     32 lanes cooperate on that task's current node.
//   * Geometry once per problem: B^T = Q R (Gram-Schmidt applied twice).  A node with kk free coefficients is the LP
//         min |r|_inf   s.t.   N^T r = chat,      N = the last q = m - kk columns of Q,  chat = N^T c,
//     i.e. a dual simplex whose basis is the set Z of the q-1 rows that are strictly inside the box, not the kk+1 rows
//     on its boundary.  H = A^-1 (q x q) lives in shared memory while q <= QS_MAX = 24 (the template parameter BIG
//     selects the home at compile time) and in an L2-resident per-warp buffer above that.  Row 0 of H is the dual
//     vector y; sg_j (n_j . y) >= 0 is the invariant, so t = y . chat is a valid lower bound at every step and a node
//     is pruned as soon as the bound crosses the radius.  Pricing is dual steepest edge with exact weights (the
//     columns of N are orthonormal), the ratio test is Harris' two-pass test.
//   * Going one level down adds one coordinate: H is bordered with an artificial column which is then pivoted out.
//   * Warm starts ("side chains"): a node starts from the final vertex of the previous sibling on the same side of
//     the level's real centre, or, for the first node of a side, from its parent's optimal vertex extended by the
//     level's coordinate.  Both chains of the level being expanded live in shared memory.
//   * A task is ONE sibling loop: all children of an explored node are solved back to back (the set of siblings
//     visited depends only on their own LP outcomes, closest.cc:110-122), every explored child becomes a record: the
//     first one is expanded next by the same warp, the others go on the warp's private stack in global memory.  A
//     loop that has produced EAGER_MAX explored children is parked as a continuation record; on the first descent of
//     the search (REC_DIVE) it is parked after the first one, so that the first leaf is reached in n node latencies.
//   * Load balance: a global MPMC queue of records.  A warp whose stack is non-empty while other warps are idle
//     moves its OLDEST record (the largest subtree) to the queue.  Every node is solved exactly once, from the same
//     starting vertex as in the sequential order, so niter and nprune do not depend on scheduling.
//   * The shared-memory layout of a warp is a compile-time function of the size class of m (SL<MC>), so every array
//     is addressed off one base register; per-warp counters live in shared memory too.
//   * bltcu_batch_kernel: many small problems, one warp each, in the reference's own depth-first order.
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

#include <algorithm>
#include <chrono>
#include <mutex>
#include <thread>
#include <vector>

#include "../../include/bltcu.h"

#define FULLMASK 0xffffffffu
#define DIST_EPS 1.0e-4  // libblt/glpk_wrapper.h:13
// Sound mode (bltcu_opts.sound; the reference asks Yices about every LP prune,
// closest.cc:258-267): a layer is pruned only when the dual bound clears the radius by this relative margin on top of
// the 1e-4 band -- orders of magnitude above the floating-point error of the bound -- and kept otherwise.
#define SOUND_MARGIN 1.0e-6
#define EPS_OPT 1.0e-9
#define PIV_TOL 1.0e-11
#define HARRIS_DELTA 1.0e-10
#define PIV_REL 1.0e-9
#define STALL_MAX 40
#define MAXD BLTCU_MAX_DIM
#define QS_MAX 24  // largest basis kept in shared memory (larger ones live in the per-warp global buffers)
#define LDS_C 26   // row stride of the shared-memory matrices: ld_vec(QS_MAX)
typedef unsigned short ord_t;  // sibling ordinal on one level of a path key (saturates at ORD_MAX)
#define ORD_MAX 65535
#define LDQ_C 26   // row stride of the shared-memory copy of the first QS_MAX + 1 columns of Q: ld_vec(QS_MAX + 1)

enum { NODE_EXPLORE = 0, NODE_PRUNE = 1, NODE_FAIL = -1 };
enum { C_NITER = 0, C_NPRUNE, C_NODES, C_DONATED, C_PIVOTS, C_QUICK, C_REFAC, C_COLDS, C_IDLE_NS, C_BUSY_NS, C_INBAND, C_COUNT = 12 };
#define CTR(s, k) do { if ((s).lane == 0) (s).ctr[k]++; } while (0)
enum { REC_TOP = 1, REC_CONT = 2, REC_DIVE = 4 };  // REC_CONT: a suspended sibling loop (resume it), not a fresh explored node
// REC_DIVE: this record is on the very first descent of the search (the reference's first dive towards the Babai
// point): its loop is parked after the first explored child so that the first leaf is reached in n node latencies
#define SHARE_EXPORT_SLOTS 64  // pool slots a rank keeps for records on their way out
#define EXP_TICKETS 4096  // records one rank can hand to other GPUs in one search
#define EAGER_MAX 8  // explored children a sibling loop produces before it is suspended in favour of its first child
// REC_TOP: the record's children are the top level (first prune ends the search, closest.cc:280-286)

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
    int fail_info;              // failcode | level << 8 of the first LP failure (diagnostics)
    int dive_state;             // first descent of the search: 0 not started, 1 under way (the other warps wait), 2 over
    // feed: subtrees the rank's host thread hands in while the kernel runs (multi-rank searches: the ranks take the
    // last eighth of the static split's subtrees from a shared counter, whoever runs dry first)
    unsigned int feed_head;     // next entry of Params::feed to take (device)
    unsigned int feed_tail;     // entries written so far (host, through the copy engine)
    int feed_open;              // host: more entries may come
    // export: records handed to OTHER GPUs of a multi-rank search.  The host grants tickets (slot of ticket t in
    // Params::exp_slots[t]); a warp that takes ticket t copies the oldest record of its stack there and marks exp_list[t].
    int export_want;            // host: tickets granted so far
    int export_claim;           // device: tickets taken
    unsigned long long deadline_ns;  // %globaltimer value after which every loop gives up (0 = none)
    unsigned long long niter, nprune, nodes, pivots, quick, refactors, colds, donated, inband;
    unsigned long long dive_end_ns;  // %globaltimer when the first descent ended (diagnostics)
    unsigned long long idle_ns, busy_ns;  // summed over warps: time spent waiting for work / expanding records
    unsigned long long hist[MAXD + 1];
    long long fail_u[MAXD];     // coefficients of the node whose LP failed first (diagnostics)
    // batch kernel only: a problem's control block is shared by its owner warp and the warps that help it
    int lock;                   // guards the problem's shared best candidate (uhat)
    int owner_inst;             // index of the problem this block belongs to
    int helped;                 // subtrees of this problem were handed to other warps
    int active;                 // global block only: problems not finished yet
    double radius_b;            // the problem's radius, for the helpers
};

struct Params {
    int n, m, LDQ, QS, LDS, LDG;
    const double *B;        // n x m (double)lat
    const double *p;        // m centre
    const double *Qrev;     // m x m: Qrev[j*m+i] = q_{m-1-i}[j]
    const double *Rm;       // n x n upper triangular, Rm[j*n+i] = q_j . b_i
    const double *QTp;      // m
    double radius;
    double radius_lp;       // the radius the node LPs prune against: radius, or radius * (1 + SOUND_MARGIN) in sound mode
    const int64_t *lat;     // n x m
    const int64_t *con;     // m x 2
    QCtl *q;
    int *ring;              // work ring (record indices, -1 = empty)
    int *fring;             // free-slot ring
    unsigned int ring_mask; // capacity - 1
    unsigned char *pool;    // subtree records
    size_t rec_stride;
    int pool_cap;
    double *wbuf;           // per-warp doubles: parent H (m x m), two side matrices for q > QS (m x LDG each), m x m scratch
    size_t wbuf_stride;
    unsigned char *nextrec; // per-warp: two record buffers for the child the warp expands next
    unsigned char *arena;   // per-warp private stack of pending records, followed by its table of offsets
    size_t arena_stride;    // bytes per warp (stack + table)
    size_t arena_bytes;     // bytes of the stack part
    int arena_slots;        // entries of the offset table
    int64_t *sol;           // n, written by the finder
    int64_t *best_u;        // per warp: n (last recorded candidate)
    ord_t *best_key;        // per warp: n
    int *best_valid;        // per warp
    int *out_list;          // static split output (record indices)
    int *feed;              // record indices handed in by the host while the kernel runs (nullptr = none)
    int static_depth;       // records shallower than this emit every child; children at this depth are collected
    int dynamic;            // 1 = donate to idle warps
    int donate_min_level;
    int lowwater;
    int no_early_exit;
    int dive_hold;          // 1 = the other warps wait while the first descent is under way
    int *exp_slots;         // multi-rank searches with a work-sharing block: pool slot of each export ticket (host-written)
    int *exp_list;          // ... 1 once the record of that ticket is in its slot (device-written, host-polled)
    int eager_max;          // explored children a sibling loop produces before it is parked (1 = the reference's depth-first order)
    int warps_per_cta;
    size_t warp_smem_doubles;
};

// record field offsets (bytes), computed identically on host and device
struct RecLayout {
    size_t off_u, off_ord, off_cpar, off_chat, off_sg, off_zrow, off_H, stride;
};
__host__ __device__ inline size_t al8(size_t x) { return (x + 7) / 8 * 8; }
// leading dimension (in doubles) of a shared-memory matrix whose rows are read with 16-byte loads, one row per lane:
// the smallest value >= x that is 2 mod 4 (even => rows stay 16-byte aligned; half of it odd => eight consecutive
// rows start in eight different 16-byte bank groups)
__host__ __device__ inline int ld_vec(int x) { return (x + 1) / 4 * 4 + 2; }
__host__ __device__ inline RecLayout rec_layout(int n, int m) {
    RecLayout L;
    size_t o = 48;  // header: int k, int flags, double xc, double gbase, int step, int state, int bytes, pad
    L.off_u = o; o += (size_t)8 * n;
    L.off_ord = o; o += al8(sizeof(ord_t) * (size_t)n);
    L.off_cpar = o; o += (size_t)8 * m;
    L.off_chat = o; o += (size_t)8 * (m + 1);
    L.off_sg = o; o += al8((size_t)m);
    L.off_zrow = o; o += al8((size_t)4 * (m + 1));
    L.off_H = o;
    // the largest record is a suspended sibling loop: parent vertex + two side vertices + their rows and signs
    o += (size_t)8 * 3 * m * m + 2 * (al8((size_t)m) + al8((size_t)4 * (m + 1)));
    L.stride = (o + 255) / 256 * 256;
    return L;
}
// bytes a record for children level k really uses (its basis has size m - k; leaf-level records carry no basis)
__host__ __device__ inline size_t rec_bytes(const RecLayout &L, int m, int k) {
    const size_t q = (size_t)(m - k);
    return k <= 1 ? L.off_sg : (L.off_H + 8 * q * q + 15) / 16 * 16;
}
// a suspended sibling loop at children level k: the record it was started from, then both sides
struct ContLayout { size_t off_HA, off_HB, off_zA, off_zB, off_sA, off_sB, bytes; };
__host__ __device__ inline ContLayout cont_layout(const RecLayout &L, int m, int k) {
    const size_t qpar = (size_t)(m - k), qd = qpar + 1;
    ContLayout C;
    size_t o = L.off_H + 8 * qpar * qpar;
    C.off_HA = o; o += 8 * qd * qd;
    C.off_HB = o; o += 8 * qd * qd;
    C.off_zA = o; o += al8(4 * (size_t)(m + 1));
    C.off_zB = o; o += al8(4 * (size_t)(m + 1));
    C.off_sA = o; o += al8((size_t)m);
    C.off_sB = o; o += al8((size_t)m);
    C.bytes = (o + 15) / 16 * 16;
    return C;
}

// ------------------------------------------------------------------------------------------------------------
// Per-warp state and primitives
// ------------------------------------------------------------------------------------------------------------
struct WS {
    double *W;  // current home of H (shared memory while qd <= QS, else the per-warp global buffer)
    int LD;
    double *Ws, *Wg, *tmpA;  // homes of the side being worked on
    int LDS, LDG, QS;
    double *chat, *val, *w, *a, *v, *r, *c, *cpar, *chatpar, *brow;
    int *zrow;
    long long *u;
    unsigned char *pos;
    ord_t *ord;
    signed char *sg;
    const double *Qs;  // Qrev in shared memory, leading dimension LDQ
    int LDQ, n, m, qd, lane;
    double *al, *be;  // alpha_j = n_j . y, beta_j = n_j . H[pe] (shared memory, m each)
    unsigned long long *ctr;  // per-warp counters in shared memory (C_* below): they would cost twenty registers otherwise
    int failcode;  // why the last NODE_FAIL happened (diagnostics)
    unsigned int repairs;
    // the two sides of the level being expanded, and the parent's vertex
    double *Ws_a, *Ws_b, *Wg_a, *Wg_b;
    const double *Qsm;  // the first QS+1 columns of Qrev in shared memory (leading dimension LDQsm); larger bases read Qrev itself
    int LDQsm;
    int *zrow_a, *zrow_b;
    unsigned char *pos_a, *pos_b;
    signed char *sg_a, *sg_b;
    double *Wpar;
    int *pzrow;
    signed char *psg;
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
#pragma unroll 1
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
// two dot4 products against the same left operand (one load of it, two fma streams); same bits as two dot4 calls
template <class FA, class FB, class FC>
__device__ __forceinline__ void dot4x2(int len, FA fa, FB fb, FC fc, double &r1, double &r2) {
    double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0, b0 = 0.0, b1 = 0.0, b2 = 0.0, b3 = 0.0;
    int e = 0;
#pragma unroll 1
    for (; e + 3 < len; e += 4) {
        const double x0 = fa(e), x1 = fa(e + 1), x2 = fa(e + 2), x3 = fa(e + 3);
        a0 = fma(x0, fb(e), a0); b0 = fma(x0, fc(e), b0);
        a1 = fma(x1, fb(e + 1), a1); b1 = fma(x1, fc(e + 1), b1);
        a2 = fma(x2, fb(e + 2), a2); b2 = fma(x2, fc(e + 2), b2);
        a3 = fma(x3, fb(e + 3), a3); b3 = fma(x3, fc(e + 3), b3);
    }
    if (e < len) { const double x = fa(e); a0 = fma(x, fb(e), a0); b0 = fma(x, fc(e), b0); ++e; }
    if (e < len) { const double x = fa(e); a1 = fma(x, fb(e), a1); b1 = fma(x, fc(e), b1); ++e; }
    if (e < len) { const double x = fa(e); a2 = fma(x, fb(e), a2); b2 = fma(x, fc(e), b2); }
    r1 = (a0 + a1) + (a2 + a3);
    r2 = (b0 + b1) + (b2 + b3);
}

// The same sums in the same order with 16-byte loads (both operands 16-byte aligned): half the load instructions.
__device__ __forceinline__ double dot4v(int len, const double *a, const double *b) {
    double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;
    const double *ae = a + (len & ~3);  // pointers are bumped so that the loop carries no index arithmetic
#pragma unroll 1
    for (; a != ae; a += 4, b += 4) {
        const double2 x0 = ((const double2 *)a)[0], x1 = ((const double2 *)a)[1];
        const double2 y0 = ((const double2 *)b)[0], y1 = ((const double2 *)b)[1];
        a0 = fma(x0.x, y0.x, a0); a1 = fma(x0.y, y0.y, a1); a2 = fma(x1.x, y1.x, a2); a3 = fma(x1.y, y1.y, a3);
    }
    const int t = len & 3;
    if (t > 0) a0 = fma(a[0], b[0], a0);
    if (t > 1) a1 = fma(a[1], b[1], a1);
    if (t > 2) a2 = fma(a[2], b[2], a2);
    return (a0 + a1) + (a2 + a3);
}
__device__ __forceinline__ void dot4x2v(int len, const double *a, const double *b, const double *c, double &r1, double &r2) {
    double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0, b0 = 0.0, b1 = 0.0, b2 = 0.0, b3 = 0.0;
    const double *ae = a + (len & ~3);
#pragma unroll 1
    for (; a != ae; a += 4, b += 4, c += 4) {
        const double2 x0 = ((const double2 *)a)[0], x1 = ((const double2 *)a)[1];
        const double2 y0 = ((const double2 *)b)[0], y1 = ((const double2 *)b)[1];
        const double2 z0 = ((const double2 *)c)[0], z1 = ((const double2 *)c)[1];
        a0 = fma(x0.x, y0.x, a0); b0 = fma(x0.x, z0.x, b0);
        a1 = fma(x0.y, y0.y, a1); b1 = fma(x0.y, z0.y, b1);
        a2 = fma(x1.x, y1.x, a2); b2 = fma(x1.x, z1.x, b2);
        a3 = fma(x1.y, y1.y, a3); b3 = fma(x1.y, z1.y, b3);
    }
    const int t = len & 3;
    if (t > 0) { const double x = a[0]; a0 = fma(x, b[0], a0); b0 = fma(x, c[0], b0); }
    if (t > 1) { const double x = a[1]; a1 = fma(x, b[1], a1); b1 = fma(x, c[1], b1); }
    if (t > 2) { const double x = a[2]; a2 = fma(x, b[2], a2); b2 = fma(x, c[2], b2); }
    r1 = (a0 + a1) + (a2 + a3);
    r2 = (b0 + b1) + (b2 + b3);
}

// Reductions of NON-NEGATIVE doubles (every caller passes |x|, a ratio of non-negatives or +inf): the bit patterns order
// like the values, so two 32-bit redux.sync steps (high word, then low word among the lanes that hold the winning high
// word) give exactly the value a shuffle tree would.
__device__ __forceinline__ double warp_max(double v) {
    const unsigned hi = (unsigned)__double2hiint(v);
    const unsigned mh = __reduce_max_sync(FULLMASK, hi);
    const unsigned ml = __reduce_max_sync(FULLMASK, hi == mh ? (unsigned)__double2loint(v) : 0u);
    return __hiloint2double((int)mh, (int)ml);
}
__device__ __forceinline__ double warp_min(double v) {
    const unsigned hi = (unsigned)__double2hiint(v);
    const unsigned mh = __reduce_min_sync(FULLMASK, hi);
    const unsigned ml = __reduce_min_sync(FULLMASK, hi == mh ? (unsigned)__double2loint(v) : 0xffffffffu);
    return __hiloint2double((int)mh, (int)ml);
}
__device__ __forceinline__ double butterfly(double v) {  // the oracle's butterfly32
    for (int off = 16; off >= 1; off >>= 1) v = v + __shfl_xor_sync(FULLMASK, v, off);
    return v;
}
// (largest value, smallest index); idx < 0 never wins
__device__ __forceinline__ void warp_argmax(double &v, int &idx) {
    for (int off = 16; off >= 1; off >>= 1) {
        const double ov = __shfl_xor_sync(FULLMASK, v, off);
        const int oi = __shfl_xor_sync(FULLMASK, idx, off);
        const bool take = (oi >= 0) && (idx < 0 || ov > v || (ov == v && oi < idx));
        if (take) { v = ov; idx = oi; }
    }
}

// the same result for lanes whose value is non-negative whenever idx >= 0 (three redux.sync steps and a vote)
__device__ __forceinline__ void warp_argmax_nn(double &v, int &idx) {
    const bool valid = idx >= 0;
    const unsigned hi = valid ? (unsigned)__double2hiint(v) : 0u;
    const unsigned mh = __reduce_max_sync(FULLMASK, hi);
    const bool m1 = valid && hi == mh;
    const unsigned lo = m1 ? (unsigned)__double2loint(v) : 0u;
    const unsigned ml = __reduce_max_sync(FULLMASK, lo);
    const unsigned mi = __reduce_min_sync(FULLMASK, (m1 && lo == ml) ? (unsigned)idx : 0xffffffffu);
    if (mi != 0xffffffffu) { idx = (int)mi; v = __hiloint2double((int)mh, (int)ml); }
}

// ---- geometry: B^T = Q R by Gram-Schmidt applied twice, one warp (twin of geo_build in the oracle) ------------
__device__ __forceinline__ double treedot_g(const double *a, const double *b, int len, int lane) {
    double acc = 0.0;
#pragma unroll 1
    for (int e = lane; e < len; e += 32) acc = fma(a[e], b[e], acc);
    return butterfly(acc);
}
__device__ int geo_build_warp(const double *B, const double *p, int n, int m, double *Q, double *Qrev, double *Rm,
                              double *QTp, double *v, double *coef, int lane) {
#pragma unroll 1
    for (int e = lane; e < n * n; e += 32) Rm[e] = 0.0;
    __syncwarp();
    for (int t = 0; t < m; ++t) {
        if (t < n) { for (int e = lane; e < m; e += 32) v[e] = B[(size_t)t * m + e]; }
        else {  // completion: the unit vector with the largest part outside the current span (ties: lowest j)
            int best = -1; double bv = -1.0;
#pragma unroll 1
            for (int j = lane; j < m; j += 32) {
                double sres = 1.0;
                for (int i = 0; i < t; ++i) { const double x = Q[(size_t)i * m + j]; sres = fma(-x, x, sres); }
                if (sres > bv) { bv = sres; best = j; }
            }
            warp_argmax(bv, best);
#pragma unroll 1
            for (int e = lane; e < m; e += 32) v[e] = (e == best) ? 1.0 : 0.0;
        }
#pragma unroll 1
        for (int e = lane; e < m; e += 32) coef[e] = 0.0;
        __syncwarp();
        for (int pass = 0; pass < 2; ++pass)
            for (int j = 0; j < t; ++j) {
                const double *qj = Q + (size_t)j * m;
                const double h = treedot_g(qj, v, m, lane);
#pragma unroll 1
                for (int e = lane; e < m; e += 32) v[e] = fma(-h, qj[e], v[e]);
                if (lane == 0) coef[j] += h;
                __syncwarp();
            }
        const double nrm = sqrt(treedot_g(v, v, m, lane));
        if (!(nrm > 0.0)) return -1;
#pragma unroll 1
        for (int e = lane; e < m; e += 32) Q[(size_t)t * m + e] = v[e] / nrm;
        if (t < n) {
#pragma unroll 1
            for (int j = lane; j < t; j += 32) Rm[(size_t)j * n + t] = coef[j];
            if (lane == 0) Rm[(size_t)t * n + t] = nrm;
        }
        __syncwarp();
    }
#pragma unroll 1
    for (int e = lane; e < m * m; e += 32) { const int j = e / m, i = e - j * m; Qrev[e] = Q[(size_t)(m - 1 - i) * m + j]; }
    for (int t = 0; t < m; ++t) { const double x = treedot_g(Q + (size_t)t * m, p, m, lane); if (lane == 0) QTp[t] = x; }
    __syncwarp();
    return 0;
}

// ---- LP engine (twin of the oracle's lp_* functions) -----------------------------------------------------------
// V: H rows, the rows of Qs and chat are 16-byte aligned (the shared-memory layout), so dot products use 16-byte loads
template <bool V>
__device__ __forceinline__ void lp_values(WS &s) {
    const int qd = s.qd;
#pragma unroll 1
    for (int p = s.lane; p < qd; p += 32) {
        const double *Hp = s.W + (size_t)p * s.LD, *ch = s.chat;
        s.val[p] = V ? dot4v(qd, Hp, ch) : dot4(qd, [&](int e) { return Hp[e]; }, [&](int e) { return ch[e]; });
    }
    __syncwarp();
}
__device__ __forceinline__ void lp_flip(WS &s) {
#pragma unroll 1
    for (int c = s.lane; c < s.qd; c += 32) s.W[c] = -s.W[c];
#pragma unroll 1
    for (int j = s.lane; j < s.m; j += 32) s.sg[j] = (signed char)-s.sg[j];
    if (s.lane == 0) s.val[0] = -s.val[0];
    __syncwarp();
}
// alpha, beta of every non-basic row; returns |N y|_1.  bmax: this lane's largest |beta| (for the ratio test's scale);
// need_repair: some row's sg*alpha fell below -HARRIS_DELTA (warp-uniform).
template <bool V>
__device__ __forceinline__ double lp_alpha_beta(WS &s, int pe, double &bmax, bool &need_repair) {
    const int qd = s.qd, m = s.m;
    const double *y = s.W, *rho = s.W + (size_t)pe * s.LD;
    double part = 0.0, bm = 0.0;
    bool bad = false;
#pragma unroll 1
    for (int j = s.lane; j < m; j += 32) {
        double a = 0.0, b = 0.0;
        if (!s.pos[j]) {
            const double *nj = s.Qs + (size_t)j * s.LDQ;
            if (V) dot4x2v(qd, nj, y, rho, a, b);
            else dot4x2(qd, [&](int e) { return nj[e]; }, [&](int e) { return y[e]; }, [&](int e) { return rho[e]; }, a, b);
            part += fabs(a);
            const double fb = fabs(b);
            bm = fb > bm ? fb : bm;
            bad = bad || ((s.sg[j] > 0 ? a : -a) < -HARRIS_DELTA);
        }
        s.al[j] = a; s.be[j] = b;
    }
    bmax = bm;
    need_repair = __any_sync(FULLMASK, bad);
    return butterfly(part);
}
// sign repair (twin of the oracle's lp_repair): rows whose sg_j*alpha_j fell below -HARRIS_DELTA go to the other bound;
// column 0 of A changes by Delta = sum -2 sg_j n_j, a rank-one correction of H.  Returns 1 if anything was flipped.
__device__ __forceinline__ int lp_repair(WS &s, bool need) {
    const int m = s.m, qd = s.qd, LD = s.LD;
    if (!need) return 0;
    __syncwarp();
#pragma unroll 1
    for (int j = s.lane; j < m; j += 32) {
        double rj = 0.0;
        if (!s.pos[j] && (double)s.sg[j] * s.al[j] < -HARRIS_DELTA) rj = -2.0 * (double)s.sg[j];
        s.r[j] = rj;
    }
    __syncwarp();
#pragma unroll 1
    for (int c = s.lane; c < qd; c += 32) {
        const double *r = s.r, *Qc = s.Qs + c;
        const int LDQ = s.LDQ;
        s.a[c] = dot4(m, [&](int e) { return r[e]; }, [&](int e) { return Qc[(size_t)e * LDQ]; });
    }
    __syncwarp();
#pragma unroll 1
    for (int i = s.lane; i < qd; i += 32) {
        const double *Hi = s.W + (size_t)i * LD, *a = s.a;
        s.w[i] = dot4(qd, [&](int e) { return Hi[e]; }, [&](int e) { return a[e]; });
    }
#pragma unroll 1
    for (int c = s.lane; c < qd; c += 32) s.v[c] = s.W[c];
    __syncwarp();
    const double den = 1.0 + s.w[0];
#pragma unroll 1
    for (int c = s.lane; c < qd; c += 32) {
        const double vc = s.v[c];
        double *Hc = s.W + c;
#pragma unroll 1
        for (int i = 0; i < qd; ++i) { const double f = s.w[i] / den; Hc[(size_t)i * LD] = fma(-f, vc, Hc[(size_t)i * LD]); }
    }
#pragma unroll 1
    for (int j = s.lane; j < m; j += 32) if (s.r[j] != 0.0) s.sg[j] = (signed char)-s.sg[j];
    __syncwarp();
    s.repairs++;
    return 1;
}
__device__ __forceinline__ int lp_ratio(WS &s, int se, int pe, bool bland, double bmax) {
    const int m = s.m, qd = s.qd;
    const bool artificial = s.zrow[pe] < 0;
    double scale = bmax;  // this lane's largest |beta| over the non-basic rows (lp_alpha_beta)
#pragma unroll 1
    for (int c = s.lane; c < qd; c += 32) { const double a = fabs(s.W[(size_t)pe * s.LD + c]); scale = a > scale ? a : scale; }
    scale = warp_max(scale);
    const double tiny = PIV_TOL * scale;
    int best = -1; double bd = 0.0;
    // tier 0 only accepts pivots that are not tiny next to the row's scale; tier 1 accepts any positive pivot
#pragma unroll 1
    for (int tier = 0; tier < 2 && best < 0; ++tier) {
        const double dmin = tier == 0 ? PIV_REL * scale : 0.0;
        double tmax = INFINITY;
#pragma unroll 1
        for (int j = s.lane; j < m; j += 32) {
            if (s.pos[j]) continue;
            const double bj = s.be[j]; const double den = ((se > 0) == (s.sg[j] > 0)) ? -bj : bj;  // -(se sg_j) beta_j
            if (!(den > dmin)) continue;
            double num = (s.sg[j] > 0 ? s.al[j] : -s.al[j]) + HARRIS_DELTA;  // rows already below zero get less room, not more
            if (num < 0.0) num = 0.0;
            const double rt = num / den;
            if (rt < tmax) tmax = rt;
        }
        tmax = warp_min(tmax);
        bd = 0.0; best = -1;
#pragma unroll 1
        for (int j = s.lane; j < m; j += 32) {
            if (s.pos[j]) continue;
            const double bj = s.be[j]; const double den = ((se > 0) == (s.sg[j] > 0)) ? -bj : bj;  // -(se sg_j) beta_j
            if (!(den > dmin)) continue;
            double num = s.sg[j] > 0 ? s.al[j] : -s.al[j];
            if (num < 0.0) num = 0.0;
            if (num / den <= tmax && (bland ? best < 0 : den > bd)) { bd = den; best = j; }
        }
        if (bland) { double key = best >= 0 ? -(double)best : 0.0; warp_argmax(key, best); bd = __shfl_sync(FULLMASK, bd, 0); }  // lowest row index
        else warp_argmax_nn(bd, best);
    }
    if (artificial) return (best >= 0 && bd > 1.0e-3 * tiny) ? best : -1;  // its pivot is den itself: refuse a vanishing one
    if (best >= 0) return best;
    double ba = -1.0;
#pragma unroll 1
    for (int j = s.lane; j < m; j += 32) {
        if (s.pos[j] || fabs(s.be[j]) > tiny) continue;
        if (fabs(s.al[j]) > ba) { ba = fabs(s.al[j]); best = j; }
    }
    warp_argmax_nn(ba, best);
    return best;
}
template <bool V>
__device__ __forceinline__ int lp_replace(WS &s, int pe, int l, int se) {
    const int qd = s.qd, LD = s.LD;
    double *H = s.W, *Hp = s.W + (size_t)pe * LD;
    const int e = s.zrow[pe];
    if (e >= 0) { for (int c = s.lane; c < qd; c += 32) Hp[c] = Hp[c] - (double)se * H[c]; }
    __syncwarp();
    const double *nl = s.Qs + (size_t)l * s.LDQ;
#pragma unroll 1
    for (int i = s.lane; i < qd; i += 32) {
        const double *Hi = H + (size_t)i * LD;
        s.w[i] = V ? dot4v(qd, Hi, nl) : dot4(qd, [&](int c) { return Hi[c]; }, [&](int c) { return nl[c]; });
    }
    __syncwarp();
    const double wp = s.w[pe];
    if (wp == 0.0) return -1;
#pragma unroll 1
    for (int c = s.lane; c < qd; c += 32) s.v[c] = Hp[c] / wp;
    __syncwarp();
    // H[i][c] -= w[i] v[c] for every row but pe.  Row pe is rewritten from v below, so it needs no exception here.
    if (qd <= 16) {  // lanes 0-15 take the even rows, lanes 16-31 the odd ones; one lane per column
        const int c = s.lane & 15;
        if (c < qd) {
            const double vc = s.v[c];
            double *Hc = H + c;
#pragma unroll 1
            for (int i = s.lane >> 4; i < qd; i += 2) Hc[(size_t)i * LD] = fma(-s.w[i], vc, Hc[(size_t)i * LD]);
        }
    } else if (qd <= 32) {  // one lane per column, rows in a loop: same arithmetic, no index division
        if (s.lane < qd) {
            const double vc = s.v[s.lane];
            double *Hc = H + s.lane;
#pragma unroll 1
            for (int i = 0; i < qd; ++i) Hc[(size_t)i * LD] = fma(-s.w[i], vc, Hc[(size_t)i * LD]);
        }
    } else {
#pragma unroll 1
        for (int x = s.lane; x < qd * qd; x += 32) {
            const int i = x / qd, c = x - i * qd;
            if (i != pe) H[(size_t)i * LD + c] = fma(-s.w[i], s.v[c], H[(size_t)i * LD + c]);
        }
    }
    __syncwarp();
    const double sl = (double)s.sg[l];
#pragma unroll 1
    for (int c = s.lane; c < qd; c += 32) Hp[c] = s.v[c] + sl * H[c];
    __syncwarp();
    if (s.lane == 0) {
        if (e >= 0) { s.sg[e] = (signed char)se; s.pos[e] = 0; }
        s.zrow[pe] = l; s.pos[l] = (unsigned char)pe;
    }
    __syncwarp();
    CTR(s, C_PIVOTS);
    return 0;
}
__device__ __forceinline__ int lp_init1(WS &s) {
    const int m = s.m;
    double part = 0.0;
#pragma unroll 1
    for (int j = s.lane; j < m; j += 32) {
        const double x = s.Qs[(size_t)j * s.LDQ];
        s.sg[j] = x >= 0.0 ? 1 : -1; s.pos[j] = 0;
        part += fabs(x);
    }
    const double sum = butterfly(part);
    if (!(sum > 0.0)) return -1;
    if (s.lane == 0) { s.W[0] = 1.0 / sum; s.zrow[0] = -1; }
    s.qd = 1; CTR(s, C_COLDS);
    __syncwarp();
    return 0;
}
// one more coordinate (index qd): border H with an artificial basic column at position qd
__device__ __forceinline__ void lp_extend(WS &s) {
    const int m = s.m, qd = s.qd, LD = s.LD;
    double part = 0.0;
#pragma unroll 1
    for (int j = s.lane; j < m; j += 32) if (!s.pos[j]) part = fma((double)s.sg[j], s.Qs[(size_t)j * s.LDQ + qd], part);
    const double a0 = butterfly(part);
    if (s.lane == 0) s.a[0] = a0;
#pragma unroll 1
    for (int p = 1 + s.lane; p < qd; p += 32) s.a[p] = s.Qs[(size_t)s.zrow[p] * s.LDQ + qd];
    __syncwarp();
#pragma unroll 1
    for (int c = s.lane; c < qd; c += 32) {
        const double *a = s.a, *Hc = s.W + c;
        s.W[(size_t)qd * LD + c] = -dot4(qd, [&](int e) { return a[e]; }, [&](int e) { return Hc[(size_t)e * LD]; });
    }
#pragma unroll 1
    for (int p = s.lane; p < qd; p += 32) s.W[(size_t)p * LD + qd] = 0.0;
    if (s.lane == 0) { s.W[(size_t)qd * LD + qd] = 1.0; s.zrow[qd] = -1; }
    s.qd = qd + 1;
    __syncwarp();
}
// cold builds only (root solve, test seam): pivot the artificial at the last position out
__device__ __forceinline__ int lp_drop_artificial(WS &s, int sf) {
    const int pe = s.qd - 1;
    double bmax;
    bool need;
    lp_alpha_beta<false>(s, pe, bmax, need);
    if (lp_repair(s, need)) lp_alpha_beta<false>(s, pe, bmax, need);
    const int l = lp_ratio(s, sf, pe, false, bmax);
    if (l < 0) return -1;
    return lp_replace<false>(s, pe, l, sf);
}
// rare path: rebuild H = A^-1 from Z and the signs by Gauss-Jordan with row pivoting, one lane, the oracle's loops.
// Takes plain pointers so that the caller's state can stay in registers.
__device__ __noinline__ int lp_refactor_impl(double *H, int LD, double *A, const double *Qs, int LDQ, int qd, int m,
                                             const unsigned char *pos, const signed char *sg, const int *zrow, int lane) {
    int rc = 0;
    __syncwarp();
    if (lane == 0) {
        for (int c = 0; c < qd; ++c) {
            double sum = 0.0;
            for (int j = 0; j < m; ++j) if (!pos[j]) sum = fma((double)sg[j], Qs[(size_t)j * LDQ + c], sum);
            A[(size_t)c * qd] = sum;
            for (int p = 1; p < qd; ++p) A[(size_t)c * qd + p] = zrow[p] >= 0 ? Qs[(size_t)zrow[p] * LDQ + c] : (c == p ? 1.0 : 0.0);
        }
        for (int i = 0; i < qd; ++i) for (int c = 0; c < qd; ++c) H[(size_t)i * LD + c] = (i == c) ? 1.0 : 0.0;
        for (int col = 0; col < qd; ++col) {
            int pr = -1; double bv = 0.0;
            for (int i = col; i < qd; ++i) if (fabs(A[(size_t)i * qd + col]) > bv) { bv = fabs(A[(size_t)i * qd + col]); pr = i; }
            if (pr < 0) { rc = -1; break; }
            if (pr != col) {
                for (int c = 0; c < qd; ++c) {
                    double t = A[(size_t)pr * qd + c]; A[(size_t)pr * qd + c] = A[(size_t)col * qd + c]; A[(size_t)col * qd + c] = t;
                    t = H[(size_t)pr * LD + c]; H[(size_t)pr * LD + c] = H[(size_t)col * LD + c]; H[(size_t)col * LD + c] = t;
                }
            }
            const double pv = A[(size_t)col * qd + col];
            for (int c = 0; c < qd; ++c) { A[(size_t)col * qd + c] /= pv; H[(size_t)col * LD + c] /= pv; }
            for (int i = 0; i < qd; ++i) {
                if (i == col) continue;
                const double f = A[(size_t)i * qd + col];
                if (f == 0.0) continue;
                for (int c = 0; c < qd; ++c) {
                    A[(size_t)i * qd + c] = fma(-f, A[(size_t)col * qd + c], A[(size_t)i * qd + c]);
                    H[(size_t)i * LD + c] = fma(-f, H[(size_t)col * LD + c], H[(size_t)i * LD + c]);
                }
            }
        }
    }
    rc = __shfl_sync(FULLMASK, rc, 0);
    __syncwarp();
    return rc;
}
__device__ __forceinline__ int lp_refactor(WS &s) {
    CTR(s, C_REFAC);
    return lp_refactor_impl(s.W, s.LD, s.tmpA, s.Qs, s.LDQ, s.qd, s.m, s.pos, s.sg, s.zrow, s.lane);
}
__device__ __forceinline__ void lp_image(WS &s, int nv) {
    const int m = s.m;
    const double t = s.val[0];
#pragma unroll 1
    for (int j = s.lane; j < m; j += 32) s.r[j] = s.pos[j] ? s.val[s.pos[j]] : (double)s.sg[j] * t;
    __syncwarp();
#pragma unroll 1
    for (int i = s.lane; i < nv; i += 32) {
        const double *Qi = s.Qs + i, *r = s.r;
        const int LDQ = s.LDQ;
        s.v[i] = dot4(m, [&](int e) { return Qi[(size_t)e * LDQ]; }, [&](int e) { return r[e]; });
    }
    __syncwarp();
}
// Solve from the current dual-feasible vertex; radius = +inf disables pruning; on NODE_EXPLORE v[qd] is valid if
// qd < m.  art != 0: the basis still holds the artificial column of a fresh coordinate at position qd-1; it is pivoted
// out first with sign `art` (the oracle's lp_drop_artificial followed by lp_solve; merged so that the pivot code
// exists once in the instruction stream).
template <bool V>
__device__ __forceinline__ int lp_solve(WS &s, double radius, double cmax, double &d_out, int art) {
    const int qd = s.qd, m = s.m;
    if (!art) {
        lp_values<V>(s);
        if (s.val[0] < 0.0) lp_flip(s);
    }
    const int cap = 100 * m + 1000;
    bool refactored = false, bland = false;
    int iter = 0, stall = 0;
    double tbest = -1.0;
#pragma unroll 1
    for (;;) {
        int pe, se;
        if (art) { pe = qd - 1; se = art; }
        else {
            const double t = s.val[0];
            if (!approxLE(t, radius)) { if (iter == 0) CTR(s, C_QUICK); return NODE_PRUNE; }
            // degenerate stalling: after STALL_MAX steps without progress leave by largest violation, after twice as many
            // switch to lowest-index (Bland) choices
            if (t > tbest * (1.0 + 1.0e-12)) { tbest = t; stall = 0; } else if (++stall >= 2 * STALL_MAX) bland = true;
            const bool dantzig = stall >= STALL_MAX;
            const double tol = EPS_OPT * (cmax + t);
            // Pricing (twin of the oracle's): among the basic rows outside the box the one with the steepest dual edge
            // leaves, score = (|r_z| - t)^2 / |H[p] - se H[0]|^2 (N has orthonormal columns, so this is the exact norm of
            // the dual ray); ties: lowest position.
            pe = -1;
            if (!bland) {
                double best = -1.0;
                const double *H0 = s.W;
#pragma unroll 1
                for (int p = 1 + s.lane; p < qd; p += 32) {
                    const double vp = s.val[p], a = fabs(vp);
                    if (!(a > t + tol)) continue;
                    const double sp = vp > 0.0 ? 1.0 : -1.0;
                    const double *Hp = s.W + (size_t)p * s.LD;
                    double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;
                    int e = 0;
                    if (V) {
#pragma unroll 1
                        for (; e + 3 < qd; e += 4) {
                            const double2 h0 = *(const double2 *)(Hp + e), h1 = *(const double2 *)(Hp + e + 2);
                            const double2 y0 = *(const double2 *)(H0 + e), y1 = *(const double2 *)(H0 + e + 2);
                            const double x0 = h0.x - sp * y0.x, x1 = h0.y - sp * y0.y, x2 = h1.x - sp * y1.x, x3 = h1.y - sp * y1.y;
                            a0 = fma(x0, x0, a0); a1 = fma(x1, x1, a1); a2 = fma(x2, x2, a2); a3 = fma(x3, x3, a3);
                        }
                    } else {
#pragma unroll 1
                        for (; e + 3 < qd; e += 4) {
                            const double x0 = Hp[e] - sp * H0[e], x1 = Hp[e + 1] - sp * H0[e + 1];
                            const double x2 = Hp[e + 2] - sp * H0[e + 2], x3 = Hp[e + 3] - sp * H0[e + 3];
                            a0 = fma(x0, x0, a0); a1 = fma(x1, x1, a1); a2 = fma(x2, x2, a2); a3 = fma(x3, x3, a3);
                        }
                    }
                    if (e < qd) { const double x = Hp[e] - sp * H0[e]; a0 = fma(x, x, a0); ++e; }
                    if (e < qd) { const double x = Hp[e] - sp * H0[e]; a1 = fma(x, x, a1); ++e; }
                    if (e < qd) { const double x = Hp[e] - sp * H0[e]; a2 = fma(x, x, a2); }
                    const double nr = (a0 + a1) + (a2 + a3);
                    const double viol = a - t;
                    const double score = dantzig ? viol : viol * viol / nr;
                    if (score > best) { best = score; pe = p; }
                }
                warp_argmax_nn(best, pe);
            } else {  // the violated basic row with the lowest row index
                int bp = -1; double key = 0.0;
#pragma unroll 1
                for (int p = 1 + s.lane; p < qd; p += 32)
                    if (fabs(s.val[p]) > t + tol && (bp < 0 || -(double)s.zrow[p] > key)) { key = -(double)s.zrow[p]; bp = p; }
                warp_argmax(key, bp);
                pe = bp;
            }
            if (pe < 0) {
                const int nv = qd < m ? qd + 1 : qd;
                lp_image(s, nv);
                double err = 0.0;
#pragma unroll 1
                for (int i = s.lane; i < qd; i += 32) { const double e = fabs(s.v[i] - s.chat[i]); err = e > err ? e : err; }
                err = warp_max(err);
                if (err <= 1.0e-8 * ((double)m * cmax + t)) { d_out = t; return NODE_EXPLORE; }
                if (refactored) { s.failcode = 1; return NODE_FAIL; }
                refactored = true;
                if (lp_refactor(s)) { s.failcode = 2; return NODE_FAIL; }
                lp_values<V>(s);
                if (s.val[0] < 0.0) lp_flip(s);
                ++iter;
                continue;
            }
            if (iter >= cap) { s.failcode = 3; return NODE_FAIL; }
            se = s.val[pe] > 0.0 ? 1 : -1;
        }
        double bmax;
        bool need;
        double nrm1 = lp_alpha_beta<V>(s, pe, bmax, need);
        if (lp_repair(s, need)) {
            if (art) nrm1 = lp_alpha_beta<V>(s, pe, bmax, need);
            else {  // the vertex moved a little: price again
                lp_values<V>(s);
                if (s.val[0] < 0.0) lp_flip(s);
                ++iter;
                continue;
            }
        }
        if (!art && fabs(nrm1 - 1.0) > 1.0e-5) {
            if (refactored) { s.failcode = 4; return NODE_FAIL; }
            refactored = true;
            if (lp_refactor(s)) { s.failcode = 5; return NODE_FAIL; }
            lp_values<V>(s);
            if (s.val[0] < 0.0) lp_flip(s);
            ++iter;
            continue;
        }
        const int l = lp_ratio(s, se, pe, bland, bmax);
        if (l < 0) { s.failcode = art ? 6 : 7; return NODE_FAIL; }
        if (lp_replace<V>(s, pe, l, se)) { s.failcode = 8; return NODE_FAIL; }
        lp_values<V>(s);
        if (art) { art = 0; if (s.val[0] < 0.0) lp_flip(s); }
        else ++iter;
    }
}

// ------------------------------------------------------------------------------------------------------------
// Level store and records
// ------------------------------------------------------------------------------------------------------------
// chain head: H (q x q packed), basic rows (q ints), signs (m bytes)
__device__ __forceinline__ void chain_store(const WS &s, double *dH, int *dI, signed char *dB) {
    const int q = s.qd, LD = s.LD;
    if (q <= 32) {  // one lane per column: no index division
        if (s.lane < q) {
            const double *src = s.W + s.lane;
            double *dst = dH + s.lane;
#pragma unroll 1
            for (int i = 0; i < q; ++i) dst[(size_t)i * q] = src[(size_t)i * LD];
        }
    } else {
#pragma unroll 1
        for (int x = s.lane; x < q * q; x += 32) { const int i = x / q, c = x - i * q; dH[x] = s.W[(size_t)i * LD + c]; }
    }
#pragma unroll 1
    for (int p = s.lane; p < q; p += 32) dI[p] = s.zrow[p];
#pragma unroll 1
    for (int j = s.lane; j < s.m; j += 32) dB[j] = s.sg[j];
}
// loads a q x q head into the working matrix
__device__ __forceinline__ void chain_load(WS &s, const double *dH, const int *dI, const signed char *dB, int q) {
    const int LD = s.LD;
    s.qd = q;
    if (q <= 32) {
        if (s.lane < q) {  // eight loads in flight per lane (the source is global memory)
            const double *src = dH + s.lane;
            double *dst = s.W + s.lane;
#pragma unroll 1
            for (int i = 0; i < q; i += 8) {
                double v[8];
#pragma unroll
                for (int e = 0; e < 8; ++e) v[e] = i + e < q ? src[(size_t)(i + e) * q] : 0.0;
#pragma unroll
                for (int e = 0; e < 8; ++e) if (i + e < q) dst[(size_t)(i + e) * LD] = v[e];
            }
        }
    } else {
#pragma unroll 1
        for (int x = s.lane; x < q * q; x += 32) { const int i = x / q, c = x - i * q; s.W[(size_t)i * LD + c] = dH[x]; }
    }
#pragma unroll 1
    for (int j = s.lane; j < s.m; j += 32) { s.sg[j] = dB[j]; s.pos[j] = 0; }
#pragma unroll 1
    for (int p = s.lane; p < q; p += 32) s.zrow[p] = dI[p];
    __syncwarp();
#pragma unroll 1
    for (int p = 1 + s.lane; p < q; p += 32) if (s.zrow[p] >= 0) s.pos[s.zrow[p]] = (unsigned char)p;
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
#pragma unroll 1
    for (int j = lane; j < P.m; j += 32) {
        unsigned long long z = 0;
        for (int i = 0; i < P.n; ++i) z += (unsigned long long)u[i] * (unsigned long long)P.lat[(size_t)i * P.m + j];
        const long long zz = (long long)z;
        if (P.con[2 * j] > zz || zz > P.con[2 * j + 1]) ok = false;
    }
    return __all_sync(FULLMASK, ok);
}

// ------------------------------------------------------------------------------------------------------------
// One task = the sibling loop below one explored node
// ------------------------------------------------------------------------------------------------------------

// per-warp private stack of pending records (global memory): [abot, acount) entries of the offset table are live
struct Arena {
    unsigned char *base;
    int *offs;       // offs[i] = byte offset of entry i; offs[acount] = top of the stack
    int acount, abot;
    size_t cap;
    int slots;
};

__device__ __forceinline__ void fail_search(const Params &P, int lane, int status) {
    if (lane == 0) { P.q->status = status; P.q->abort = 1; }
}
// BIG = the basis of this task is larger than the shared-memory matrices: H lives in the per-warp global buffers.
// The choice is a template parameter so that in the common case every access to H compiles to LDS/STS.
template <bool BIG>
__device__ __forceinline__ void use_side(WS &s, int side) {
    if (BIG) { s.W = side ? s.Wg_b : s.Wg_a; s.LD = s.LDG; } else { s.W = side ? s.Ws_b : s.Ws_a; s.LD = s.LDS; }
    s.zrow = side ? s.zrow_b : s.zrow_a; s.pos = side ? s.pos_b : s.pos_a; s.sg = side ? s.sg_b : s.sg_a;
}
// Four loads in flight per lane before the first store: a rolled load-store loop pays one L2 round trip per iteration.
__device__ __forceinline__ void copy_words(void *dst, const void *src, size_t bytes, int lane) {  // bytes % 8 == 0
    unsigned long long *d = (unsigned long long *)dst;
    const unsigned long long *s = (const unsigned long long *)src;
    const size_t nw = bytes / 8;
#pragma unroll 1
    for (size_t i = lane; i < nw; i += 128) {
        const unsigned long long v0 = s[i];
        const unsigned long long v1 = i + 32 < nw ? s[i + 32] : 0ull;
        const unsigned long long v2 = i + 64 < nw ? s[i + 64] : 0ull;
        const unsigned long long v3 = i + 96 < nw ? s[i + 96] : 0ull;
        d[i] = v0;
        if (i + 32 < nw) d[i + 32] = v1;
        if (i + 64 < nw) d[i + 64] = v2;
        if (i + 96 < nw) d[i + 96] = v3;
    }
}
// writes the record of an explored child (children level k-1) whose vertex is the current working state
// Takes an export ticket if the host has granted one (QCtl::export_want) and copies the record there; 1 if it did.
__device__ __noinline__ int export_oldest(QCtl *q, const int *exp_slots, int *exp_list, unsigned char *pool, size_t stride,
                                          const unsigned char *src, size_t bytes, int lane) {
    int ticket = -1, eslot = -1;
    if (lane == 0) {
        const int want = *(volatile int *)&q->export_want;
        int c = *(volatile int *)&q->export_claim;
        while (c < want) {
            const int old = atomicCAS(&q->export_claim, c, c + 1);
            if (old == c) { ticket = c; eslot = ((const volatile int *)exp_slots)[c]; break; }
            c = old;
        }
    }
    ticket = __shfl_sync(FULLMASK, ticket, 0);
    eslot = __shfl_sync(FULLMASK, eslot, 0);
    if (ticket < 0) return 0;
    copy_words(pool + (size_t)eslot * stride, src, bytes, lane);
    __syncwarp();
    __threadfence_system();
    if (lane == 0) ((volatile int *)exp_list)[ticket] = 1;
    return 1;
}

__device__ __forceinline__ void write_child(const Params &P, const RecLayout &RL, WS &s, unsigned char *C, int k, double xlast,
                                            double gn) {
    const int n = P.n, m = P.m, lane = s.lane, qd = s.qd;
    if (lane == 0) { ((int *)C)[0] = k - 1; ((int *)C)[1] = 0; ((double *)C)[1] = xlast; ((double *)C)[2] = gn; ((int *)C)[6] = 0; ((int *)C)[7] = 0; }
#pragma unroll 1
    for (int e = lane; e < n; e += 32) { ((long long *)(C + RL.off_u))[e] = s.u[e]; ((ord_t *)(C + RL.off_ord))[e] = s.ord[e]; }
#pragma unroll 1
    for (int j = lane; j < m; j += 32) ((double *)(C + RL.off_cpar))[j] = s.c[j];
#pragma unroll 1
    for (int e = lane; e < qd; e += 32) ((double *)(C + RL.off_chat))[e] = s.chat[e];
    if (k - 1 > 1) chain_store(s, (double *)(C + RL.off_H), (int *)(C + RL.off_zrow), (signed char *)(C + RL.off_sg));
    __syncwarp();
}

// Expands record R.  Returns the number of children left for this warp in `next` (0 or 1); others went to the arena,
// the queue or the static-split output.
template <bool BIG>
__device__ __forceinline__ int expand_impl(const Params &P, WS &s, const unsigned char *R, int pool_slot, unsigned char *next, Arena &A,
                      int gwarp, unsigned int *hist) {
    const int n = P.n, m = P.m, lane = s.lane;
    const RecLayout RL = rec_layout(n, m);
    const int k0 = ((const int *)R)[0];
    const int rflags = ((const int *)R)[1];
    double xc0 = ((const double *)R)[1];
    double gb = ((const double *)R)[2];
    const bool is_top = (rflags & REC_TOP) != 0, is_cont = (rflags & REC_CONT) != 0;
    const int emax = (rflags & REC_DIVE) ? 1 : P.eager_max;
    const int c_step = ((const int *)R)[6], c_state = ((const int *)R)[7];
    const int kk = k0 - 1, qd0 = m - kk, qpar = m - k0;

    if (BIG) { s.Qs = P.Qrev; s.LDQ = m; } else { s.Qs = s.Qsm; s.LDQ = s.LDQsm; }
    // ---- take what is needed out of the record; afterwards the record is dead ----
#pragma unroll 1
    for (int e = lane; e < n; e += 32) { s.u[e] = ((const long long *)(R + RL.off_u))[e]; s.ord[e] = ((const ord_t *)(R + RL.off_ord))[e]; }
#pragma unroll 1
    for (int j = lane; j < m; j += 32) s.cpar[j] = ((const double *)(R + RL.off_cpar))[j];
    if (qpar > 0 && k0 > 1) {
        copy_words(s.Wpar, R + RL.off_H, (size_t)qpar * qpar * 8, lane);
#pragma unroll 1
        for (int p = lane; p < qpar; p += 32) s.pzrow[p] = ((const int *)(R + RL.off_zrow))[p];
#pragma unroll 1
        for (int j = lane; j < m; j += 32) s.psg[j] = ((const signed char *)(R + RL.off_sg))[j];
#pragma unroll 1
        for (int e = lane; e < qpar; e += 32) s.chat[e] = ((const double *)(R + RL.off_chat))[e];
    }
    if (is_cont) {
        const ContLayout CL = cont_layout(RL, m, k0);
        const int st = c_state;
        for (int side = 0; side < 2; ++side) {
            if (!((st >> (1 + side)) & 1)) continue;
            use_side<BIG>(s, side);
            chain_load(s, (const double *)(R + (side ? CL.off_HB : CL.off_HA)), (const int *)(R + (side ? CL.off_zB : CL.off_zA)),
                       (const signed char *)(R + (side ? CL.off_sB : CL.off_sA)), qd0);
        }
    }
    __syncwarp();
    if (pool_slot >= 0) {
        __threadfence();
        if (lane == 0) ring_push(P.q, &P.q->ftail, P.fring, P.ring_mask, pool_slot);  // the record slot is free again
    }

    double d = 0.0, xlast = 0.0;
    // ---- constants of this level ----
#pragma unroll 1
    for (int j = lane; j < m; j += 32) s.brow[j] = P.B[(size_t)kk * m + j];
    const double Rkk = P.Rm[(size_t)kk * n + kk];
    double Rk1 = 1.0, Rk1k = 0.0, g2 = 0.0;
    if (kk >= 1) {
        const double *Rr = P.Rm + (size_t)(kk - 1) * n;
        Rk1 = Rr[kk - 1]; Rk1k = Rr[kk];
        double part = 0.0;
#pragma unroll 1
        for (int i = kk + 1 + lane; i < n; i += 32) part = fma(-(double)s.u[i], Rr[i], part);
        g2 = P.QTp[kk - 1] + butterfly(part);
    }
    long long ucur = (long long)round(xc0);
    int step = sgn_d(xc0 - (double)ucur);
    bool pruned = false;
    int inited = 0, nkept = 0, nexp = 0;
    if (is_cont) {  // resume a suspended loop: the next sibling, the zig-zag state and both side vertices were saved
        ucur = s.u[kk];
        step = c_step; pruned = (c_state & 1) != 0; inited = (c_state >> 1) & 3;
    } else if (lane == 0) { s.u[kk] = ucur; s.ord[kk] = 0; }
    __syncwarp();
    const bool forced = (n - k0) < P.static_depth;  // static split: every explored child of this record is published
    unsigned int wd = 0;

    for (;;) {
        {   // one lane reads the stop flag so that the whole warp takes the same branch
            int ab = 0;
            if (lane == 0) {
                ab = *(volatile int *)&P.q->abort;
                if (!ab && (++wd & 63u) == 0) ab = overdue(P.q, 1) ? 1 : 0;
            }
            if (__shfl_sync(FULLMASK, ab, 0)) return nkept;
        }
        // c = cpar - u * B[kk]
        double cmax = 0.0;
        {
            const double mu = -(double)ucur;
#pragma unroll 1
            for (int j = lane; j < m; j += 32) { const double cj = fma(mu, s.brow[j], s.cpar[j]); s.c[j] = cj; const double a = fabs(cj); cmax = a > cmax ? a : cmax; }
            cmax = warp_max(cmax);
            __syncwarp();
        }
        if (lane == 0) { s.ctr[C_NODES]++; hist[kk]++; }
        int status;
        double gn = 0.0;
        if (k0 == 1) {
            d = cmax;
            status = approxLE(d, P.radius) ? NODE_EXPLORE : NODE_PRUNE;
        } else {
            const int dir = sgn_d(xc0 - (double)ucur);
            const int side = dir > 0 ? 0 : 1;
            use_side<BIG>(s, side);
            const double clast = fma(-(double)ucur, Rkk, gb);
            if (lane == 0) s.chat[qd0 - 1] = clast;
            if (!((inited >> side) & 1)) {
                if (qd0 == 1) { if (lp_init1(s)) { fail_search(P, lane, BLTCU_STATUS_LP_FAIL); return nkept; } }
                else chain_load(s, s.Wpar, s.pzrow, s.psg, qpar);  // the parent's vertex, one coordinate short
                inited |= 1 << side;
            } else s.qd = qd0;
            __syncwarp();
            // One loop serves the warm start (at most one extension) and the last resort after a failed warm start:
            // rebuild the vertex coordinate by coordinate from nothing, every intermediate problem solved to optimality.
            bool cold = false;
            status = NODE_FAIL;
#pragma unroll 1
            for (;;) {
                int art = 0;
                double rad = P.radius_lp;
                if (s.qd < qd0) {
                    lp_extend(s);
                    if (cold) { lp_values<!BIG>(s); art = s.val[s.qd - 1] > 0.0 ? 1 : -1; if (s.qd < qd0) rad = INFINITY; }
                    else art = dir;
                }
                status = lp_solve<!BIG>(s, rad, cmax, d, art);
                if (status == NODE_FAIL) {
                    if (cold || lp_init1(s)) { status = NODE_FAIL; break; }
                    cold = true;
                    continue;
                }
                if (s.qd == qd0) break;
                if (status != NODE_EXPLORE) { status = NODE_FAIL; break; }
            }
            if (status == NODE_FAIL) {
                // Remember the failure and go on as if the node had been pruned: another warp may find a point inside
                // the box, and then the failure is irrelevant (the sequential order might never have come here).
                if (lane == 0 && atomicCAS(&P.q->fail_info, 0, s.failcode | (k0 << 8) | (qd0 << 16)) == 0) { for (int e = 0; e < n; ++e) P.q->fail_u[e] = s.u[e]; }
                inited &= ~(1 << side);  // this side's vertex is unusable: the next node of the side starts over
                status = NODE_PRUNE;
            }
            if (status == NODE_EXPLORE) {  // real centre of the next level: x_{kk-1} of the optimum
                if (P.radius_lp != P.radius && !approxLE(d, P.radius)) CTR(s, C_INBAND);  // kept by the sound margin alone
                gn = fma(-(double)ucur, Rk1k, g2);
                xlast = (gn - s.v[qd0]) / Rk1;
            }
        }
        if (status == NODE_PRUNE) {
            if (is_top) return nkept;  // closest.cc:280-286: first prune at the top level ends the search
            CTR(s, C_NPRUNE);
            if (pruned) return nkept;  // second prune: the level is left
            pruned = true;
        } else if (k0 > 1) {
            // explored inner node: its children are one more task
            CTR(s, C_NITER);
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
                else {  // private stack full: publish instead
                    if (lane == 0) slot = ring_pop(P.q, &P.q->fhead, &P.q->ftail, P.fring, P.ring_mask);
                    slot = __shfl_sync(FULLMASK, slot, 0);
                    if (slot < 0) { fail_search(P, lane, BLTCU_STATUS_ERROR); return nkept; }
                    C = P.pool + (size_t)slot * RL.stride;
                    dest = 2;
                }
            }
            write_child(P, RL, s, C, k0, xlast, gn);
            if ((rflags & REC_DIVE) && dest == 0 && lane == 0) ((int *)C)[1] = REC_DIVE;
            nexp++;
            if (dest == 0) nkept = 1;
            else if (dest == 1) { if (lane == 0) A.offs[A.acount + 1] = (int)(A.offs[A.acount] + need); A.acount++; __syncwarp(); }
            else {
                __threadfence();
                if (lane == 0) {
                    if (dest == 3) { const unsigned int o = atomicAdd(&P.q->out_count, 1u); P.out_list[o] = slot; }
                    else { atomicAdd(&P.q->pending, 1); ring_push(P.q, &P.q->tail, P.ring, P.ring_mask, slot); }
                }
                CTR(s, C_DONATED);
            }
        } else {
            // leaf: record, closest.cc:300-312
            {
                ord_t *bk = P.best_key + (size_t)gwarp * n;
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
#pragma unroll 1
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
#pragma unroll 1
                    for (int e = lane; e < n; e += 32) P.sol[e] = s.u[e];
                    __threadfence();
                }
                if (!P.no_early_exit) { if (lane == 0) P.q->abort = 1; return nkept; }
            }
            if (n == 1) return nkept;  // D4 of oracle/blt_oracle.c
            // explore(), closest.cc:99-104: the leaf level is abandoned and the parent level moves on (its loop has
            // already run here), so this task simply ends
            CTR(s, C_NITER);
            return nkept;
        }
        // next sibling, closest.cc:118-121
        ucur += step; step = -step - (step > 0 ? 1 : -1);
        if (lane == 0) { s.u[kk] = ucur; if (s.ord[kk] < ORD_MAX) s.ord[kk]++; }
        __syncwarp();
        // Wide level or little stack space left: park the rest of this loop and go down into the first child, the way
        // the depth-first order would.  The parked loop is a record like any other (it can even be donated).
        if (status == NODE_EXPLORE && !forced && k0 > 1 && (nexp >= emax || (size_t)A.offs[A.acount] * 2 > A.cap)) {
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
            if (lane == 0) {
                ((int *)C)[0] = k0; ((int *)C)[1] = (is_top ? REC_TOP : 0) | REC_CONT; ((double *)C)[1] = xc0; ((double *)C)[2] = gb;
                ((int *)C)[6] = step; ((int *)C)[7] = (pruned ? 1 : 0) | (inited << 1); ((int *)C)[8] = (int)CL.bytes;
            }
#pragma unroll 1
            for (int e = lane; e < n; e += 32) { ((long long *)(C + RL.off_u))[e] = s.u[e]; ((ord_t *)(C + RL.off_ord))[e] = s.ord[e]; }
#pragma unroll 1
            for (int j = lane; j < m; j += 32) { ((double *)(C + RL.off_cpar))[j] = s.cpar[j]; ((signed char *)(C + RL.off_sg))[j] = s.psg[j]; }
#pragma unroll 1
            for (int e = lane; e < qpar; e += 32) { ((double *)(C + RL.off_chat))[e] = s.chat[e]; ((int *)(C + RL.off_zrow))[e] = s.pzrow[e]; }
            if (qpar > 0) copy_words(C + RL.off_H, s.Wpar, (size_t)qpar * qpar * 8, lane);
            for (int side = 0; side < 2; ++side) {
                if (!((inited >> side) & 1)) continue;
                use_side<BIG>(s, side);
                s.qd = qd0;
                chain_store(s, (double *)(C + (side ? CL.off_HB : CL.off_HA)), (int *)(C + (side ? CL.off_zB : CL.off_zA)),
                            (signed char *)(C + (side ? CL.off_sB : CL.off_sA)));
            }
            __syncwarp();
            if (in_arena) { if (lane == 0) A.offs[A.acount + 1] = (int)(top + CL.bytes); A.acount++; __syncwarp(); }
            else {
                __threadfence();
                if (lane == 0) { atomicAdd(&P.q->pending, 1); ring_push(P.q, &P.q->tail, P.ring, P.ring_mask, slot); }
                CTR(s, C_DONATED);
            }
            return nkept;
        }
    }
}

__device__ __forceinline__ int expand(const Params &P, WS &s, const unsigned char *R, int pool_slot, unsigned char *next, Arena &A,
                                      int gwarp, unsigned int *hist) {
    const int k0 = ((const int *)R)[0];
    if (P.m - k0 + 1 <= QS_MAX) return expand_impl<false>(P, s, R, pool_slot, next, A, gwarp, hist);
    return expand_impl<true>(P, s, R, pool_slot, next, A, gwarp, hist);
}

// carves the per-warp shared-memory block; returns the node histogram
__device__ __forceinline__ unsigned int *carve_warp(WS &s, double *base, int n, int m, int QS, int LDS) {
    const int M1 = m + 1;
    s.Ws_a = base; base += (size_t)QS * LDS;
    s.Ws_b = base; base += (size_t)QS * LDS;
    s.chat = base; base += M1; s.val = base; base += M1; s.w = base; base += M1; s.a = base; base += M1; s.v = base; base += M1;
    s.r = base; base += m; s.c = base; base += m; s.cpar = base; base += m; s.brow = base; base += m;
    s.al = base; base += m; s.be = base; base += m;
    s.u = (long long *)base; base += n;
    s.ctr = (unsigned long long *)base; base += C_COUNT;
    int *ib = (int *)base;
    s.zrow_a = ib; ib += M1; s.zrow_b = ib; ib += M1; s.pzrow = ib; ib += M1;
    unsigned int *hist = (unsigned int *)ib; ib += n + 1;
    unsigned char *bb = (unsigned char *)ib;
    const int nb = ((n > m ? n : m) + 8) / 8 * 8;
    s.pos_a = bb; bb += nb; s.pos_b = bb; bb += nb;
    s.sg_a = (signed char *)bb; bb += nb; s.sg_b = (signed char *)bb; bb += nb;
    s.psg = (signed char *)bb; bb += nb;
    s.ord = (ord_t *)bb;
    s.chatpar = nullptr;
    return hist;
}
__host__ __device__ inline size_t warp_smem_bytes(int n, int m, int QS, int LDS) {
    const size_t M1 = (size_t)m + 1;
    const size_t d = 2 * (size_t)QS * LDS + 5 * M1 + 6 * (size_t)m + n + C_COUNT;
    const size_t i = 3 * M1 + (size_t)(n + 1);
    const size_t nb = (size_t)((n > m ? n : m) + 8) / 8 * 8;
    return (d * 8 + i * 4 + 7 * nb + 15) / 16 * 16;
}

// Compile-time layout of the same block for the search kernels.  MC is the size class of the problem (m <= MC): every
// array sits at a constant offset from the warp's base, so the compiler addresses all of them off ONE register with
// immediate offsets instead of keeping (or recomputing) thirty pointers.
template <int MC>
struct SL {
    static constexpr int V = MC + 2;             // doubles per vector (even, so everything stays 16-byte aligned)
    static constexpr int H = QS_MAX * LDS_C;     // doubles per matrix
    static constexpr int NB = MC + 8;            // bytes per byte array
    static constexpr int o_vec = 2 * H;          // chat val w a v r c cpar brow al be
    static constexpr int o_u = o_vec + 11 * V;
    static constexpr int o_ctr = o_u + MC;       // C_COUNT counters
    static constexpr int o_int = o_ctr + C_COUNT; // zrow_a zrow_b pzrow hist, V ints each = 2 V doubles
    static constexpr int o_byte = o_int + 2 * V; // pos_a pos_b sg_a sg_b psg ord
    static constexpr size_t bytes = ((size_t)o_byte * 8 + 7 * NB + 15) / 16 * 16;
};
template <int MC>
__device__ __forceinline__ unsigned int *carve_fixed(WS &s, double *base) {
    typedef SL<MC> L;
    s.Ws_a = base; s.Ws_b = base + L::H;
    double *v = base + L::o_vec;
    s.chat = v; s.val = v + L::V; s.w = v + 2 * L::V; s.a = v + 3 * L::V; s.v = v + 4 * L::V; s.r = v + 5 * L::V;
    s.c = v + 6 * L::V; s.cpar = v + 7 * L::V; s.brow = v + 8 * L::V; s.al = v + 9 * L::V; s.be = v + 10 * L::V;
    s.u = (long long *)(base + L::o_u);
    s.ctr = (unsigned long long *)(base + L::o_ctr);
    int *ib = (int *)(base + L::o_int);
    s.zrow_a = ib; s.zrow_b = ib + L::V; s.pzrow = ib + 2 * L::V;
    unsigned int *hist = (unsigned int *)(ib + 3 * L::V);
    unsigned char *bb = (unsigned char *)(base + L::o_byte);
    s.pos_a = bb; s.pos_b = bb + L::NB;
    s.sg_a = (signed char *)(bb + 2 * L::NB); s.sg_b = (signed char *)(bb + 3 * L::NB); s.psg = (signed char *)(bb + 4 * L::NB);
    s.ord = (ord_t *)(bb + 5 * L::NB);
    s.chatpar = nullptr;
    return hist;
}
__host__ inline int size_class(int m) { return m <= 32 ? 32 : (m <= 64 ? 64 : 96); }
__host__ inline size_t fixed_warp_bytes(int mc) { return mc == 32 ? SL<32>::bytes : (mc == 64 ? SL<64>::bytes : SL<96>::bytes); }

// ------------------------------------------------------------------------------------------------------------
// The persistent kernel
// ------------------------------------------------------------------------------------------------------------
template <int MC>
__global__ void __launch_bounds__(384, 1) bltcu_search_kernel(const Params P) {
    extern __shared__ __align__(16) double smem[];
    const int n = P.n, m = P.m;
    double *Qs = smem;
    for (int e = threadIdx.x; e < m * LDQ_C; e += blockDim.x) {
        const int j = e / LDQ_C, i = e - j * LDQ_C;
        Qs[e] = i < m ? P.Qrev[(size_t)j * m + i] : 0.0;
    }
    __syncthreads();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int gwarp = blockIdx.x * P.warps_per_cta + warp;
    const RecLayout RL = rec_layout(n, m);
    WS s;
    unsigned int *hist = carve_fixed<MC>(s, smem + (size_t)m * LDQ_C + (size_t)warp * (SL<MC>::bytes / 8));
    {
        double *wb = P.wbuf + (size_t)gwarp * P.wbuf_stride;
        s.Wpar = wb; wb += (size_t)m * m;
        s.Wg_a = wb; wb += (size_t)m * P.LDG;
        s.Wg_b = wb; wb += (size_t)m * P.LDG;
        s.tmpA = wb;
    }
    s.LDS = LDS_C; s.LDG = P.LDG; s.QS = QS_MAX;
    s.Qsm = Qs; s.LDQsm = LDQ_C;
    s.Qs = Qs; s.LDQ = LDQ_C; s.n = n; s.m = m; s.qd = 0; s.lane = lane;
    s.repairs = 0; s.failcode = 0;
    if (lane < C_COUNT) s.ctr[lane] = 0ull;
    use_side<false>(s, 0);
#pragma unroll 1
    for (int e = lane; e <= n; e += 32) hist[e] = 0;
    Arena A;
    A.base = P.arena + (size_t)gwarp * P.arena_stride;
    A.offs = (int *)(A.base + P.arena_bytes);
    A.cap = P.arena_bytes; A.slots = P.arena_slots; A.acount = 0; A.abot = 0;
    if (lane == 0) A.offs[0] = 0;
    __syncwarp();
    unsigned char *nextbuf = P.nextrec + (size_t)gwarp * 2 * RL.stride;
    int nextsel = 0, have_next = 0;
    QCtl *q = P.q;
    bool am_idle = false, have_token = false;
    unsigned int tasks = 0;
    unsigned long long t_mark = gtimer();
    for (;;) {
        const unsigned char *R = nullptr;
        int slot = -1;
        if (have_next) {
            R = nextbuf + (size_t)nextsel * RL.stride;
            nextsel ^= 1; have_next = 0;
            int ab = 0;
            if (lane == 0) ab = *(volatile int *)&q->abort;
            if (__shfl_sync(FULLMASK, ab, 0)) break;
        } else if (A.acount > A.abot) {
            A.acount--;
            R = A.base + A.offs[A.acount];
            if (A.acount == A.abot) { A.acount = 0; A.abot = 0; if (lane == 0) A.offs[0] = 0; __syncwarp(); }
            int ab = 0;
            if (lane == 0) ab = *(volatile int *)&q->abort;
            if (__shfl_sync(FULLMASK, ab, 0)) break;
        } else {
            int stop = 0;
            if (lane == 0) {
                if (have_token) { __threadfence(); atomicSub(&q->pending, 1); }
                if (*(volatile int *)&q->abort) stop = 1;
                else {
                    // While the first descent is under way the other warps stay out of its way: it decides how fast a
                    // problem with a point near the Babai point is answered, and it takes n node latencies at most.
                    const bool hold = P.dive_hold && *(volatile int *)&q->dive_state == 1;
                    if (hold) { __nanosleep(500); if (overdue(q, 2)) stop = 1; }
                    else slot = ring_pop(q, &q->head, &q->tail, P.ring, P.ring_mask);
                    if (slot >= 0 && (((const int *)(P.pool + (size_t)slot * RL.stride))[1] & REC_DIVE)) q->dive_state = 1;
                    if (!hold && slot < 0 && P.feed &&
                        (int)(*(volatile unsigned int *)&q->feed_tail - *(volatile unsigned int *)&q->feed_head) > 0) {  // a subtree from the host's feed
                        atomicAdd(&q->pending, 1);      // counted before it is taken, so that nobody sees an empty search
                        for (;;) {
                            const unsigned int fh = *(volatile unsigned int *)&q->feed_head, ft = *(volatile unsigned int *)&q->feed_tail;
                            if ((int)(ft - fh) <= 0) { atomicSub(&q->pending, 1); break; }
                            if (atomicCAS(&q->feed_head, fh, fh + 1) == fh) { slot = ((volatile int *)P.feed)[fh]; break; }
                        }
                    }
                    if (hold) {}
                    else if (slot < 0) {
                        const bool fed_out = !P.feed || (*(volatile int *)&q->feed_open == 0 &&
                                                         (int)(*(volatile unsigned int *)&q->feed_tail - *(volatile unsigned int *)&q->feed_head) <= 0);
                        if (*(volatile int *)&q->pending == 0 && fed_out) stop = 1;
                        else {
                            if (!am_idle) { atomicAdd(&q->idle, 1); am_idle = true; }
                            __nanosleep(500);
                            if (overdue(q, 2)) stop = 1;
                        }
                    } else if (am_idle) { atomicSub(&q->idle, 1); am_idle = false; }
                }
            }
            have_token = false;
            stop = __shfl_sync(FULLMASK, stop, 0);
            slot = __shfl_sync(FULLMASK, slot, 0);
            if (stop) break;
            if (slot < 0) continue;
            have_token = true;  // the record's count in `pending` now stands for this busy warp
            R = P.pool + (size_t)slot * RL.stride;
        }
        // a popped stack record sits where new children would be pushed: expand() copies what it needs first, and the
        // stack top was lowered above, so the children overwrite it only after it is dead
        { const unsigned long long t = gtimer(); if (lane == 0) s.ctr[C_IDLE_NS] += t - t_mark; t_mark = t; }
        const int was_dive = ((const int *)R)[1] & REC_DIVE;
        // the watchdog must reach warps that never run dry: one look at the clock every 16 tasks
        if ((++tasks & 15u) == 0) {
            int od = 0;
            if (lane == 0) od = overdue(q, 1) ? 1 : 0;
            if (__shfl_sync(FULLMASK, od, 0)) break;
        }
        have_next = expand(P, s, R, slot, nextbuf + (size_t)nextsel * RL.stride, A, gwarp, hist);
        __syncwarp();
        int burst = 0;  // the first descent just ended: its parked loops go to the waiting warps all at once
        if (was_dive && !(have_next && (((const int *)(nextbuf + (size_t)nextsel * RL.stride))[1] & REC_DIVE))) {
            if (lane == 0) { q->dive_end_ns = gtimer(); q->dive_state = 2; __threadfence(); }  // everybody in
            burst = P.dive_hold;
        }
        { const unsigned long long t = gtimer(); if (lane == 0) s.ctr[C_BUSY_NS] += t - t_mark; t_mark = t; }
        // another GPU of the search asked for work: the oldest record of the stack goes into the slot its ticket names
        // (a call, so that searches without a work-sharing block carry one uniform test here and nothing else)
        if (P.exp_list && A.acount > A.abot) {
            if (export_oldest(q, P.exp_slots, P.exp_list, P.pool, RL.stride, A.base + A.offs[A.abot],
                              (size_t)(A.offs[A.abot + 1] - A.offs[A.abot]), lane)) {
                A.abot++;
                CTR(s, C_DONATED);
                if (A.acount == A.abot) { A.acount = 0; A.abot = 0; if (lane == 0) A.offs[0] = 0; }
                __syncwarp();
            }
        }
        // hand the oldest pending record to an idle warp
#pragma unroll 1
        for (int rounds = burst ? A.acount - A.abot : 1; rounds > 0 && P.dynamic && A.acount > A.abot; --rounds) {
            int want = 0, dslot = -1;
            if (lane == 0) {
                const int idle = *(volatile int *)&q->idle;
                const int qlen = (int)(*(volatile unsigned int *)&q->tail - *(volatile unsigned int *)&q->head);
                want = (burst || (idle > 0 && qlen < idle + P.lowwater)) ? 1 : 0;
                if (want) dslot = ring_pop(q, &q->fhead, &q->ftail, P.fring, P.ring_mask);
            }
            dslot = __shfl_sync(FULLMASK, dslot, 0);
            if (dslot >= 0) {
                const unsigned char *src = A.base + A.offs[A.abot];
                const size_t bytes = (size_t)(A.offs[A.abot + 1] - A.offs[A.abot]);
                copy_words(P.pool + (size_t)dslot * RL.stride, src, bytes, lane);
                __syncwarp();
                __threadfence();
                if (lane == 0) { atomicAdd(&q->pending, 1); ring_push(q, &q->tail, P.ring, P.ring_mask, dslot); }
                A.abot++;
                CTR(s, C_DONATED);
                if (A.acount == A.abot) { A.acount = 0; A.abot = 0; if (lane == 0) A.offs[0] = 0; __syncwarp(); }
            }
        }
    }
    if (lane == 0) {
        if (have_token) { __threadfence(); atomicSub(&q->pending, 1); }
        if (am_idle) atomicSub(&q->idle, 1);
        atomicAdd(&q->niter, s.ctr[C_NITER]); atomicAdd(&q->nprune, s.ctr[C_NPRUNE]); atomicAdd(&q->nodes, s.ctr[C_NODES]);
        atomicAdd(&q->donated, s.ctr[C_DONATED]);
        atomicAdd(&q->pivots, s.ctr[C_PIVOTS]); atomicAdd(&q->quick, s.ctr[C_QUICK]); atomicAdd(&q->refactors, s.ctr[C_REFAC]);
        atomicAdd(&q->colds, s.ctr[C_COLDS]); atomicAdd(&q->inband, s.ctr[C_INBAND]);
        atomicAdd(&q->idle_ns, s.ctr[C_IDLE_NS] + (gtimer() - t_mark)); atomicAdd(&q->busy_ns, s.ctr[C_BUSY_NS]);
    }
    __syncwarp();
#pragma unroll 1
    for (int e = lane; e <= n; e += 32) if (hist[e]) atomicAdd(&q->hist[e], (unsigned long long)hist[e]);
}

// small helper kernels ---------------------------------------------------------------------------------------
// One warp: fp64 basis, centre, the geometry of the problem, then the initial solve LP(n) (closest.cc:206) whose
// optimum is the parent vertex of the top level.  Writes the top-level record into `rec`.  The basis of this one
// solve lives in global memory (Wg); it is a cold path.  smem: warp_smem_bytes(n, m, 0, 1) bytes of shared memory.
// cnt_out (global): pivots, cold starts, refactorisations, quick prunes of this solve.
__device__ __forceinline__ int setup_warp(const int64_t *lat, const int64_t *con, int n, int m, double *B, double *p, double *Q,
                                          double *Qrev, double *Rm, double *QTp, double *tmp, double *Wg, unsigned char *rec,
                                          double *smem, unsigned long long *cnt_out, int lane) {
#pragma unroll 1
    for (int e = lane; e < n * m; e += 32) B[e] = (double)lat[e];                                 // closest.cc:203
#pragma unroll 1
    for (int j = lane; j < m; j += 32) p[j] = (double)(con[2 * j + 1] + con[2 * j]) / 2.0;        // closest.cc:180-185
    __syncwarp();
    int r = geo_build_warp(B, p, n, m, Q, Qrev, Rm, QTp, tmp, tmp + m, lane);
    if (r == 0) {
        WS s;
        (void)carve_warp(s, smem, n, m, 0, 1);
        s.Wg_a = Wg; s.Wg_b = Wg; s.tmpA = Wg + (size_t)m * (m | 1); s.Wpar = nullptr;
        s.LDS = 1; s.LDG = m | 1; s.QS = 0;
        use_side<true>(s, 0);
        s.Qs = Qrev; s.LDQ = m; s.n = n; s.m = m; s.qd = 0; s.lane = lane;
        s.repairs = 0;
        if (lane < C_COUNT) s.ctr[lane] = 0ull;
        __syncwarp();
        const RecLayout RL = rec_layout(n, m);
        const int qpar = m - n;
        double cm = 0.0, d = 0.0, vq = 0.0;
#pragma unroll 1
        for (int j = lane; j < m; j += 32) { const double a = fabs(p[j]); cm = a > cm ? a : cm; }
        cm = warp_max(cm);
#pragma unroll 1
        for (size_t i = lane; i < RL.stride / 8; i += 32) ((unsigned long long *)rec)[i] = 0ull;
        __syncwarp();
        if (qpar > 0) {
#pragma unroll 1
            for (int i = lane; i < qpar; i += 32) s.chat[i] = QTp[m - 1 - i];
            __syncwarp();
            r = lp_init1(s);
            for (int i = 1; i < qpar && r == 0; ++i) {
                lp_extend(s);
                lp_values<false>(s);
                r = lp_drop_artificial(s, s.val[i] > 0.0 ? 1 : -1);
            }
            if (r == 0 && lp_solve<false>(s, INFINITY, cm, d, 0) != NODE_EXPLORE) r = -1;
            if (r == 0) {
                vq = s.v[qpar];
#pragma unroll 1
                for (int e = lane; e < qpar; e += 32) ((double *)(rec + RL.off_chat))[e] = s.chat[e];
                chain_store(s, (double *)(rec + RL.off_H), (int *)(rec + RL.off_zrow), (signed char *)(rec + RL.off_sg));
            }
        }
        if (r == 0) {
            const double gb = QTp[n - 1];
            const double xlast = (gb - vq) / Rm[(size_t)(n - 1) * n + (n - 1)];
            if (lane == 0) { ((int *)rec)[0] = n; ((int *)rec)[1] = REC_TOP | REC_DIVE; ((double *)rec)[1] = xlast; ((double *)rec)[2] = gb; }
#pragma unroll 1
            for (int j = lane; j < m; j += 32) ((double *)(rec + RL.off_cpar))[j] = p[j];
        }
        __syncwarp();
        if (lane == 0) { cnt_out[0] = s.ctr[C_PIVOTS]; cnt_out[1] = s.ctr[C_COLDS]; cnt_out[2] = s.ctr[C_REFAC]; cnt_out[3] = s.ctr[C_QUICK]; }
    }
    __syncwarp();
    return r;
}
__global__ void bltcu_setup_kernel(const int64_t *lat, const int64_t *con, int n, int m, double *B, double *p, double *Q,
                                   double *Qrev, double *Rm, double *QTp, double *tmp, unsigned long long *now, int *rc,
                                   double *Wg, unsigned char *rec) {
    extern __shared__ __align__(16) double smem[];
    const int lane = threadIdx.x;
    if (lane == 0) *now = gtimer();
    const int r = setup_warp(lat, con, n, m, B, p, Q, Qrev, Rm, QTp, tmp, Wg, rec, smem, now + 2, lane);
    if (lane == 0) *rc = r;
}
// ------------------------------------------------------------------------------------------------------------
// Batch form: many small independent problems of one shape, one warp per problem
// ------------------------------------------------------------------------------------------------------------
// Every warp takes the next problem off a counter, builds that problem's geometry in its own buffers, and searches it
// alone with its private stack.  Sibling loops are parked after their first explored child (eager_max = 1), which is
// exactly the reference's depth-first order: status, found, niter, nprune and the returned vector are those of
// closest() on that problem alone, for SAT problems too.
struct BatchParams {
    Params P;                  // layout fields shared by all problems; the per-problem pointers are set by each warp
    const int64_t *lat, *con;  // count x n*m, count x 2m
    int count;
    int *next;                 // problem counter
    unsigned char *geo;        // per-warp: B, p, Q, Qrev, Rm, QTp, tmp (doubles), then 4 counters
    size_t geo_stride;
    QCtl *ctl;                 // per-warp control blocks
    int64_t *sol;              // count x n
    int *status, *found;       // count
    unsigned long long *niter, *nprune;  // count
    unsigned long long *totals;  // nodes, pivots, quick prunes, refactorisations, cold starts, then hist[MAXD + 1]
    unsigned long long timeout_ns;
    size_t warp_doubles;       // shared memory per warp: Q (m x LDQ) + the search block
    // helping: warps that ran out of problems take subtrees of the problems still running
    int allow_help;
    QCtl *gq;                  // global block: help ring, free ring, idle helpers, problems not finished
    int *hring, *hfring;       // help ring / free-slot ring (record indices)
    unsigned int hmask;
    unsigned char *hpool;      // records handed over
    int64_t *sbest_u;          // per owner warp: the problem's shared best candidate, its key and validity
    ord_t *sbest_key;
    int *sbest_valid;
    int *helped;               // count: 1 if the problem was searched by more than one warp
};
// what a warp of the batch kernel owns
struct BWarp {
    double *Qs, *wsm, *wb;       // shared-memory copy of Q, shared-memory search block, global per-warp buffers
    unsigned char *nextbuf;
    int gwarp, lane;
};
// loads Q of the problem whose geometry P points to and readies the warp's search state; returns the node histogram
template <int MC>
__device__ __forceinline__ unsigned int *batch_prepare(const Params &P, WS &s, Arena &A, const BWarp &w) {
    const int n = P.n, m = P.m, lane = w.lane;
#pragma unroll 1
    for (int e = lane; e < m * LDQ_C; e += 32) {
        const int j = e / LDQ_C, i = e - j * LDQ_C;
        w.Qs[e] = i < m ? P.Qrev[(size_t)j * m + i] : 0.0;
    }
    unsigned int *hist = carve_fixed<MC>(s, w.wsm);
    s.Wpar = w.wb; s.Wg_a = w.wb + (size_t)m * m; s.Wg_b = s.Wg_a + (size_t)m * P.LDG; s.tmpA = s.Wg_b + (size_t)m * P.LDG;
    s.LDS = LDS_C; s.LDG = P.LDG; s.QS = QS_MAX;
    s.Qsm = w.Qs; s.LDQsm = LDQ_C;
    s.Qs = w.Qs; s.LDQ = LDQ_C; s.n = n; s.m = m; s.qd = 0; s.lane = lane;
    s.repairs = 0; s.failcode = 0;
    if (lane < C_COUNT) s.ctr[lane] = 0ull;
    use_side<false>(s, 0);
#pragma unroll 1
    for (int e = lane; e <= n; e += 32) hist[e] = 0;
    A.base = P.arena + (size_t)w.gwarp * P.arena_stride;
    A.offs = (int *)(A.base + P.arena_bytes);
    A.cap = P.arena_bytes; A.slots = P.arena_slots; A.acount = 0; A.abot = 0;
    if (lane == 0) { A.offs[0] = 0; P.best_valid[w.gwarp] = 0; }
    __syncwarp();
    return hist;
}
// folds this warp's last recorded leaf into the problem's shared candidate (later in depth-first order wins)
__device__ __forceinline__ void batch_merge_best(const BatchParams &BP, const Params &P, QCtl *oq, int owner, int gwarp, int lane) {
    const int n = P.n;
    int bv = 0;
    if (lane == 0) bv = P.best_valid[gwarp];
    if (!__shfl_sync(FULLMASK, bv, 0)) return;
    const ord_t *mk = P.best_key + (size_t)gwarp * n;
    ord_t *sk = BP.sbest_key + (size_t)owner * n;
    int newer = 1;
    if (lane == 0) {
        while (atomicCAS(&oq->lock, 0, 1) != 0) __nanosleep(50);
        __threadfence();
        if (*(volatile int *)&BP.sbest_valid[owner]) {
            int cmp = 0;
            for (int e = n - 1; e >= 0 && cmp == 0; --e) cmp = (int)mk[e] - (int)((volatile ord_t *)sk)[e];
            newer = cmp >= 0;
        }
    }
    newer = __shfl_sync(FULLMASK, newer, 0);
    if (newer) {
#pragma unroll 1
        for (int e = lane; e < n; e += 32) { sk[e] = mk[e]; BP.sbest_u[(size_t)owner * n + e] = P.best_u[(size_t)gwarp * n + e]; }
    }
    __syncwarp();
    __threadfence();
    if (lane == 0) { if (newer) BP.sbest_valid[owner] = 1; __threadfence(); atomicExch(&oq->lock, 0); }
    __syncwarp();
}
// Runs the warp's own stack to exhaustion.  have_next: a record waits in the first `next` buffer.  While other warps
// are idle, the OLDEST record of the stack (the largest subtree left) is handed over after every task.
template <int MC>
__device__ __forceinline__ void batch_run(const BatchParams &BP, const Params &P, WS &s, Arena &A, const BWarp &w,
                                          const RecLayout &RL, int have_next, int owner, unsigned int *hist) {
    const int lane = w.lane;
    QCtl *oq = P.q;
    int nextsel = 0;
    unsigned int tasks = 0;
    for (;;) {
        const unsigned char *R;
        if (have_next) { R = w.nextbuf + (size_t)nextsel * RL.stride; nextsel ^= 1; have_next = 0; }
        else if (A.acount > A.abot) {
            A.acount--;
            R = A.base + A.offs[A.acount];
            if (A.acount == A.abot) { A.acount = 0; A.abot = 0; if (lane == 0) A.offs[0] = 0; __syncwarp(); }
        } else break;
        int ab = 0;
        if (lane == 0) ab = *(volatile int *)&oq->abort;
        if (__shfl_sync(FULLMASK, ab, 0)) break;
        if ((++tasks & 15u) == 0) {  // the watchdog must reach warps that never run dry
            int od = 0;
            if (lane == 0) od = overdue(oq, 1) ? 1 : 0;
            if (__shfl_sync(FULLMASK, od, 0)) break;
        }
        have_next = expand(P, s, R, -1, w.nextbuf + (size_t)nextsel * RL.stride, A, w.gwarp, hist);
        __syncwarp();
        if (BP.allow_help && A.acount > A.abot) {
            int dslot = -1;
            if (lane == 0) {
                const int idle = *(volatile int *)&BP.gq->idle;
                // a record near the leaves is less work than handing it over
                if (idle > 0 && ((const int *)(A.base + A.offs[A.abot]))[0] >= P.donate_min_level) {
                    const int qlen = (int)(*(volatile unsigned int *)&BP.gq->tail - *(volatile unsigned int *)&BP.gq->head);
                    if (qlen < idle) dslot = ring_pop(BP.gq, &BP.gq->fhead, &BP.gq->ftail, BP.hfring, BP.hmask);
                }
            }
            dslot = __shfl_sync(FULLMASK, dslot, 0);
            if (dslot >= 0) {
                unsigned char *dst = BP.hpool + (size_t)dslot * RL.stride;
                const unsigned char *src = A.base + A.offs[A.abot];
                copy_words(dst, src, (size_t)(A.offs[A.abot + 1] - A.offs[A.abot]), lane);
                __syncwarp();
                if (lane == 0) ((int *)dst)[9] = owner;
                __threadfence();
                if (lane == 0) { oq->helped = 1; atomicAdd(&oq->pending, 1); ring_push(BP.gq, &BP.gq->tail, BP.hring, BP.hmask, dslot); }
                A.abot++;
                if (A.acount == A.abot) { A.acount = 0; A.abot = 0; if (lane == 0) A.offs[0] = 0; }
                __syncwarp();
            }
        }
    }
}
// The kernel is a loop over work units with ONE call site of the search loop: a unit is either the warp's own share of
// a new problem (it becomes the problem's owner) or a subtree handed over by another owner.  Every unit folds its
// counters and its last recorded leaf into the owner's control block; the owner writes the problem's results when
// its own share and all handed-over subtrees are in (pending == 0).
template <int MC>
__global__ void __launch_bounds__(384, 1) bltcu_batch_kernel(const BatchParams BP) {
    extern __shared__ __align__(16) double smem[];
    Params P = BP.P;
    const int n = P.n, m = P.m;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int gwarp = blockIdx.x * P.warps_per_cta + warp;
    const RecLayout RL = rec_layout(n, m);
    BWarp w;
    w.Qs = smem + (size_t)warp * BP.warp_doubles;
    w.wsm = w.Qs + (size_t)m * LDQ_C;
    w.wb = P.wbuf + (size_t)gwarp * P.wbuf_stride;
    w.nextbuf = P.nextrec + (size_t)gwarp * 2 * RL.stride;
    w.gwarp = gwarp; w.lane = lane;
    const size_t geo_off_p = (size_t)n * m, geo_off_Q = geo_off_p + m, geo_off_Qrev = geo_off_Q + (size_t)m * m;
    const size_t geo_off_Rm = geo_off_Qrev + (size_t)m * m, geo_off_QTp = geo_off_Rm + (size_t)n * n, geo_off_tmp = geo_off_QTp + m;
    double *gown = (double *)(BP.geo + (size_t)gwarp * BP.geo_stride);
    unsigned long long *sc = (unsigned long long *)(gown + geo_off_tmp + 2 * (size_t)m);
    QCtl *q = BP.ctl + gwarp;
    P.static_depth = 0; P.dynamic = 0; P.eager_max = 1;
    unsigned long long deadline = 0;
    if (BP.timeout_ns) deadline = gtimer() + BP.timeout_ns;
    deadline = __shfl_sync(FULLMASK, deadline, 0);
    unsigned long long t_nodes = 0, t_piv = 0, t_quick = 0, t_ref = 0, t_cold = 0;
    int owning = -1;   // problem this warp owns and has not written the results of yet
    int more = 1;      // the problem counter is not exhausted
    int am_idle = 0;
    for (;;) {
        int owner = -1, have_next = 0, am_idle_unit = 0;  // am_idle_unit: this unit came from the help ring
        // ---- 1. results of the own problem, once everything handed over has come back ----
        if (owning >= 0) {
            int pend = 0;
            if (lane == 0) pend = *(volatile int *)&q->pending;
            if (__shfl_sync(FULLMASK, pend, 0) == 0) {
                __threadfence();
                int st = 0, fnd = 0, fi = 0, helped = 0;
                unsigned long long h[7] = {0, 0, 0, 0, 0, 0, 0};
                if (lane == 0) {
                    st = *(volatile int *)&q->status; fnd = *(volatile int *)&q->found; fi = *(volatile int *)&q->fail_info;
                    helped = *(volatile int *)&q->helped;
                    h[0] = *(volatile unsigned long long *)&q->niter; h[1] = *(volatile unsigned long long *)&q->nprune;
                    h[2] = *(volatile unsigned long long *)&q->nodes; h[3] = *(volatile unsigned long long *)&q->pivots;
                    h[4] = *(volatile unsigned long long *)&q->colds; h[5] = *(volatile unsigned long long *)&q->refactors;
                    h[6] = *(volatile unsigned long long *)&q->quick;
                }
                st = __shfl_sync(FULLMASK, st, 0); fnd = __shfl_sync(FULLMASK, fnd, 0); fi = __shfl_sync(FULLMASK, fi, 0);
                if (!fnd && fi && st == BLTCU_STATUS_OK) st = BLTCU_STATUS_LP_FAIL;
                int64_t *sol = BP.sol + (size_t)owning * n;
                if (!fnd) {  // the last recorded candidate (uhat), as the reference returns it
                    int bv = 0;
                    if (lane == 0) bv = *(volatile int *)&BP.sbest_valid[gwarp];
                    if (__shfl_sync(FULLMASK, bv, 0)) {
#pragma unroll 1
                        for (int e = lane; e < n; e += 32) sol[e] = BP.sbest_u[(size_t)gwarp * n + e];
                    }
                }
                if (lane == 0) {
                    t_nodes += h[2] + 1; t_piv += h[3] + sc[0]; t_cold += h[4] + sc[1]; t_ref += h[5] + sc[2]; t_quick += h[6] + sc[3];
                    BP.status[owning] = st; BP.found[owning] = fnd; BP.niter[owning] = h[0]; BP.nprune[owning] = h[1];
                    BP.helped[owning] = helped;
                    atomicAdd(&BP.totals[5 + n], 1ull);
                    __threadfence();
                    atomicSub(&BP.gq->active, 1);
                }
                __syncwarp();
                owning = -1;
            }
        }
        // ---- 2. a new problem of its own ----
        if (owning < 0 && more) {
            int inst = 0;
            if (lane == 0) inst = atomicAdd(BP.next, 1);
            inst = __shfl_sync(FULLMASK, inst, 0);
            if (inst >= BP.count) more = 0;
            else {
                const int64_t *lat = BP.lat + (size_t)inst * n * m, *con = BP.con + (size_t)inst * 2 * m;
                int64_t *sol = BP.sol + (size_t)inst * n;
#pragma unroll 1
                for (int e = lane; e < (int)(sizeof(QCtl) / 8); e += 32) ((unsigned long long *)q)[e] = 0ull;
#pragma unroll 1
                for (int e = lane; e < n; e += 32) sol[e] = 0;
                __syncwarp();
                const double radius = (double)(con[1] - con[0]) / 2.0;  // closest.cc:193-194: row 0 only
                if (lane == 0) {
                    sc[0] = sc[1] = sc[2] = sc[3] = 0; BP.sbest_valid[gwarp] = 0;
                    q->deadline_ns = deadline; q->owner_inst = inst; q->radius_b = radius; q->pending = 1;  // 1 = the own share
                }
                __syncwarp();
                __threadfence();
                const int r = setup_warp(lat, con, n, m, gown, gown + geo_off_p, gown + geo_off_Q, gown + geo_off_Qrev, gown + geo_off_Rm,
                                         gown + geo_off_QTp, gown + geo_off_tmp, w.wb, w.nextbuf, w.wsm, sc, lane);
                if (r != 0) {  // dependent basis rows or a failed initial solve
                    if (lane == 0) {
                        BP.status[inst] = BLTCU_STATUS_LP_FAIL; BP.found[inst] = 0; BP.niter[inst] = 0; BP.nprune[inst] = 0; BP.helped[inst] = 0;
                        __threadfence();
                        atomicSub(&BP.gq->active, 1);
                    }
                    __syncwarp();
                    continue;
                }
                owning = inst; owner = gwarp; have_next = 1;  // setup_warp left the top-level record in the first buffer
            }
        }
        // ---- 3. otherwise a subtree somebody handed over ----
        if (owner < 0) {
            int slot = -1, stop = 0;
            if (lane == 0) {
                if (BP.allow_help) slot = ring_pop(BP.gq, &BP.gq->head, &BP.gq->tail, BP.hring, BP.hmask);
                if (slot < 0) {
                    if (owning < 0 && !more && (!BP.allow_help || *(volatile int *)&BP.gq->active == 0)) stop = 1;
                    else {
                        if (!am_idle && BP.allow_help) { atomicAdd(&BP.gq->idle, 1); am_idle = 1; }
                        __nanosleep(1000);
                        if (deadline && gtimer() > deadline) stop = 1;
                    }
                } else if (am_idle) { atomicSub(&BP.gq->idle, 1); am_idle = 0; }
            }
            stop = __shfl_sync(FULLMASK, stop, 0);
            slot = __shfl_sync(FULLMASK, slot, 0);
            if (stop) break;
            if (slot < 0) continue;
            // the record moves into this warp's first buffer, the pool slot is free again
            const unsigned char *R = BP.hpool + (size_t)slot * RL.stride;
            owner = ((const int *)R)[9];
            const int k0 = ((const int *)R)[0], fl = ((const int *)R)[1];
            const size_t bytes = (fl & REC_CONT) ? (size_t)((const int *)R)[8] : rec_bytes(RL, m, k0);
            copy_words(w.nextbuf, R, (bytes + 7) / 8 * 8, lane);
            __syncwarp();
            __threadfence();
            if (lane == 0) ring_push(BP.gq, &BP.gq->ftail, BP.hfring, BP.hmask, slot);
            have_next = 1; am_idle_unit = 1;
        } else if (lane == 0 && am_idle) {
            atomicSub(&BP.gq->idle, 1);
            am_idle = 0;
        }
        // ---- 4. the search itself: geometry, bounds and control block are the owner's ----
        QCtl *oq = BP.ctl + owner;
        {
            double *g = (double *)(BP.geo + (size_t)owner * BP.geo_stride);
            P.B = g; P.p = g + geo_off_p; P.Qrev = g + geo_off_Qrev; P.Rm = g + geo_off_Rm; P.QTp = g + geo_off_QTp;
            int inst = 0;
            double radius = 0.0;
            if (lane == 0) { inst = *(volatile int *)&oq->owner_inst; radius = *(volatile double *)&oq->radius_b; }
            inst = __shfl_sync(FULLMASK, inst, 0);
            radius = __shfl_sync(FULLMASK, radius, 0);
            P.lat = BP.lat + (size_t)inst * n * m; P.con = BP.con + (size_t)inst * 2 * m; P.sol = BP.sol + (size_t)inst * n;
            P.radius = radius; P.radius_lp = radius;
            P.q = oq;
            // a handed-over subtree belongs to a shared problem: its order no longer matters (see the second pass on the
            // host), so it is expanded eagerly, which produces the records other idle warps are waiting for
            P.eager_max = (owner == gwarp && owning >= 0 && have_next && !am_idle_unit) ? 1 : EAGER_MAX;
        }
        WS s;
        Arena A;
        unsigned int *hist = batch_prepare<MC>(P, s, A, w);
        batch_run<MC>(BP, P, s, A, w, RL, have_next, owner, hist);
        __syncwarp();
        // ---- 5. fold this unit into the owner's block ----
        if (lane == 0) {
            atomicAdd(&oq->niter, s.ctr[C_NITER]); atomicAdd(&oq->nprune, s.ctr[C_NPRUNE]); atomicAdd(&oq->nodes, s.ctr[C_NODES]);
            atomicAdd(&oq->pivots, s.ctr[C_PIVOTS]); atomicAdd(&oq->quick, s.ctr[C_QUICK]); atomicAdd(&oq->refactors, s.ctr[C_REFAC]);
            atomicAdd(&oq->colds, s.ctr[C_COLDS]);
        }
#pragma unroll 1
        for (int e = lane; e <= n; e += 32) if (hist[e]) atomicAdd(&BP.totals[5 + e], (unsigned long long)hist[e]);
        batch_merge_best(BP, P, oq, owner, gwarp, lane);
        __threadfence();
        if (lane == 0) atomicSub(&oq->pending, 1);
        __syncwarp();
    }
    if (lane == 0) {
        if (am_idle) atomicSub(&BP.gq->idle, 1);
        atomicAdd(&BP.totals[0], t_nodes); atomicAdd(&BP.totals[1], t_piv); atomicAdd(&BP.totals[2], t_quick);
        atomicAdd(&BP.totals[3], t_ref); atomicAdd(&BP.totals[4], t_cold);
    }
}

// static split: collects the path keys of the published subtrees into one contiguous buffer (one copy to the host)
__global__ void bltcu_gather_keys_kernel(const unsigned char *pool, size_t stride, size_t off_ord, const int *out_list,
                                         int count, int n, ord_t *keys) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count * n) return;
    const int r = i / n, e = i - r * n;
    keys[i] = ((const ord_t *)(pool + (size_t)out_list[r] * stride + off_ord))[n - 1 - e];  // top level first
}
__global__ void bltcu_check_kernel(const int64_t *lat, const int64_t *con, int n, int m, const int64_t *u, int *ok) {
    Params P; P.n = n; P.m = m; P.lat = lat; P.con = con;
    const bool r = check_soln_dev(P, (const long long *)u, threadIdx.x & 31);
    if (threadIdx.x == 0) *ok = r ? 1 : 0;
}
// one cold layer-distance solve (test seam): c computed as glpk_wrapper.cc:156-166 does.  Everything but the small
// vectors lives in global memory here (QS = 0): this kernel is a test seam, not a hot path.
__global__ void bltcu_infdist_kernel(const double *B, int n, int m, const int64_t *u, int k, const double *p, double *x,
                                     double *dist, int *rc, double *Q, double *Qrev, double *Rm, double *QTp, double *Wg,
                                     double *tmp) {
    extern __shared__ __align__(16) double smem[];
    const int lane = threadIdx.x;
    WS s;
    (void)carve_warp(s, smem, n, m, 0, 1);
    s.Wg_a = Wg; s.Wg_b = Wg; s.tmpA = Wg + (size_t)m * (m | 1); s.Wpar = nullptr;
    s.LDS = 1; s.LDG = m | 1; s.QS = 0;
    use_side<true>(s, 0);
    s.Qs = Qrev; s.LDQ = m; s.n = n; s.m = m; s.qd = 0; s.lane = lane;
    s.repairs = 0;
    if (lane < C_COUNT) s.ctr[lane] = 0ull;
    __syncwarp();
    int r = geo_build_warp(B, p, n, m, Q, Qrev, Rm, QTp, tmp, tmp + m, lane);
    double cm = 0.0;
#pragma unroll 1
    for (int i = lane; i < m; i += 32) {
        double rhs1 = 0.0;
        for (int j = k; j < n; ++j) rhs1 += (double)(-u[j]) * B[(size_t)j * m + i];
        rhs1 += p[i];
        s.c[i] = rhs1; s.r[i] = 0.0;
        cm = fabs(rhs1) > cm ? fabs(rhs1) : cm;
    }
    cm = warp_max(cm);
    __syncwarp();
    double d = 0.0;
    const int qd = m - k;
    if (k == 0) d = cm;
    else if (r == 0) {
        if (qd > 0) {
#pragma unroll 1
            for (int i = lane; i < qd; i += 32) {
                const double *Qi = Qrev + i, *c = s.c;
                s.chat[i] = dot4(m, [&](int e) { return Qi[(size_t)e * m]; }, [&](int e) { return c[e]; });
            }
            __syncwarp();
            r = lp_init1(s);
            for (int i = 1; i < qd && r == 0; ++i) {
                lp_extend(s);
                lp_values<false>(s);
                r = lp_drop_artificial(s, s.val[i] > 0.0 ? 1 : -1);
            }
            if (r == 0 && lp_solve<false>(s, INFINITY, cm, d, 0) != NODE_EXPLORE) r = -1;
            if (r == 0) {
                const double t = s.val[0];
#pragma unroll 1
                for (int j = lane; j < m; j += 32) s.r[j] = s.pos[j] ? s.val[s.pos[j]] : (double)s.sg[j] * t;
                __syncwarp();
            }
        }
        if (r == 0 && lane == 0) {  // x = R_k^-1 Q_k^T (c - r), one lane, the oracle's loops
            for (int i = 0; i < k; ++i) {
                double acc = 0.0;
                for (int j = 0; j < m; ++j) acc = fma(Q[(size_t)i * m + j], s.c[j] - s.r[j], acc);
                s.w[i] = acc;
            }
            for (int i = k - 1; i >= 0; --i) {
                double acc = s.w[i];
                for (int t = i + 1; t < k; ++t) acc = fma(-Rm[(size_t)i * n + t], x[t], acc);
                x[i] = acc / Rm[(size_t)i * n + i];
            }
        }
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
    cudaStream_t stream = nullptr, stream2 = nullptr;  // stream2: small copies beside a running search kernel
    int *pin = nullptr;                                // page-locked staging for them
    unsigned char *rec_pin = nullptr;                  // page-locked staging for two records (work sharing between GPUs)
    size_t rec_cap = 0;
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    std::mutex mu;  // one call at a time per device; different devices run concurrently (opts.gpus)
};
std::mutex g_mu;
std::vector<DevWs *> g_ws;

int get_ws(int device, DevWs **out) {
    std::lock_guard<std::mutex> lk(g_mu);  // guards the list only
    for (auto *w : g_ws) if (w->device == device) { CK(cudaSetDevice(device)); *out = w; return 0; }
    CK(cudaSetDevice(device));
    DevWs *w = new DevWs();
    w->device = device;
    int v = 0;
    CK(cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, device)); w->sms = v;
    CK(cudaDeviceGetAttribute(&v, cudaDevAttrMaxSharedMemoryPerBlockOptin, device)); w->smem_optin = (size_t)v;
    CK(cudaStreamCreateWithFlags(&w->stream, cudaStreamNonBlocking));
    CK(cudaStreamCreateWithFlags(&w->stream2, cudaStreamNonBlocking));
    CK(cudaHostAlloc((void **)&w->pin, 4096, cudaHostAllocDefault));
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

// Layout of one search's global workspace, shared by the cached-workspace path and the fixed-array path.
struct Layout {
    int KQS, LDS, LDG, LDQ, wpc, ctas, nwarps, pool_cap, arena_slots;
    size_t warp_bytes, smem_bytes, wbuf_stride, arena_bytes, arena_stride;
    size_t o_exp;  // export tickets: EXP_TICKETS slots, then EXP_TICKETS ready marks
    size_t o_q, o_in, o_B, o_p, o_Q, o_Qrev, o_Rm, o_QTp, o_tmp, o_rc, o_ring, o_fring, o_out, o_sol, o_bu, o_bk, o_bv, o_pool,
        o_wbuf, o_next, o_arena, total;
    RecLayout RL;
};
// Bytes of private stack with which a warp never has to publish a record for lack of room, whatever the tree: a
// loop parks after every explored child once the stack is half full, so from there on each level adds one suspended
// loop at most (batch_launch sizes its stacks the same way).
size_t arena_worst_case(const RecLayout &RL, int n, int m) {
    size_t b = RL.stride;
    for (int k0 = 2; k0 <= n; ++k0) b += 2 * cont_layout(RL, m, k0).bytes;
    return b;
}
// arena_recs: size of each warp's private stack in units of the largest record (0: the worst-case size above)
int make_layout(Layout &L, int n, int m, size_t smem_budget, int sms, int want_wpc, int want_ctas, int max_wpc, int pool_cap,
                int arena_recs) {
    L.KQS = QS_MAX; L.LDS = LDS_C; L.LDG = m | 1; L.LDQ = LDQ_C;
    L.warp_bytes = fixed_warp_bytes(size_class(m));
    const size_t q_bytes = (size_t)m * L.LDQ * 8;
    int wpc = (int)((smem_budget > q_bytes ? smem_budget - q_bytes : 0) / L.warp_bytes);
    if (wpc > max_wpc) wpc = max_wpc;
    if (want_wpc > 0 && want_wpc < wpc) wpc = want_wpc;
    if (wpc < 1) return -1;
    L.wpc = wpc;
    L.ctas = want_ctas > 0 ? want_ctas : sms;
    L.nwarps = L.ctas * wpc;
    L.smem_bytes = q_bytes + (size_t)wpc * L.warp_bytes;
    L.RL = rec_layout(n, m);
    L.pool_cap = pool_cap;
    while ((size_t)L.pool_cap * L.RL.stride > ((size_t)3 << 30) && L.pool_cap > 256) L.pool_cap >>= 1;
    L.wbuf_stride = align_up((size_t)2 * m * m + (size_t)2 * m * L.LDG, 32);
    L.arena_slots = 1024;
    L.arena_bytes = align_up(arena_recs > 0 ? (size_t)arena_recs * L.RL.stride : arena_worst_case(L.RL, n, m), 256);
    L.arena_stride = align_up(L.arena_bytes + (size_t)L.arena_slots * 4, 256);
    size_t off = 0;
    auto carve = [&](size_t bytes) { size_t r = off; off = align_up(off + bytes, 256); return r; };
    L.o_q = carve(sizeof(QCtl));
    L.o_in = carve(((size_t)n * m + 2 * (size_t)m) * 8);
    L.o_B = carve((size_t)n * m * 8);
    L.o_p = carve((size_t)m * 8);
    L.o_Q = carve((size_t)m * m * 8);
    L.o_Qrev = carve((size_t)m * m * 8);
    L.o_Rm = carve((size_t)n * n * 8);
    L.o_QTp = carve((size_t)m * 8);
    L.o_tmp = carve((size_t)2 * m * 8);
    L.o_rc = carve(64);
    L.o_ring = carve((size_t)L.pool_cap * 4);
    L.o_fring = carve((size_t)L.pool_cap * 4);
    L.o_out = carve((size_t)L.pool_cap * 4);
    L.o_exp = carve((size_t)2 * EXP_TICKETS * 4);
    L.o_sol = carve((size_t)n * 8 + 8);
    L.o_bu = carve((size_t)L.nwarps * n * 8);
    L.o_bk = carve((size_t)L.nwarps * n * sizeof(ord_t));
    L.o_bv = carve((size_t)L.nwarps * 4);
    L.o_pool = carve((size_t)L.pool_cap * L.RL.stride);
    L.o_wbuf = carve((size_t)L.nwarps * L.wbuf_stride * 8);
    L.o_next = carve((size_t)L.nwarps * 2 * L.RL.stride);
    L.o_arena = carve((size_t)L.nwarps * L.arena_stride);
    L.total = off;
    return 0;
}
void fill_params(Params &P, const Layout &L, unsigned char *base, int n, int m) {
    memset(&P, 0, sizeof P);
    P.n = n; P.m = m; P.LDQ = L.LDQ; P.QS = L.KQS; P.LDS = L.LDS; P.LDG = L.LDG;
    P.B = (double *)(base + L.o_B); P.p = (double *)(base + L.o_p);
    P.Qrev = (double *)(base + L.o_Qrev); P.Rm = (double *)(base + L.o_Rm); P.QTp = (double *)(base + L.o_QTp);
    P.q = (QCtl *)(base + L.o_q);
    P.ring = (int *)(base + L.o_ring); P.fring = (int *)(base + L.o_fring); P.ring_mask = (unsigned int)L.pool_cap - 1;
    P.pool = base + L.o_pool; P.rec_stride = L.RL.stride; P.pool_cap = L.pool_cap;
    P.wbuf = (double *)(base + L.o_wbuf); P.wbuf_stride = L.wbuf_stride;
    P.nextrec = base + L.o_next;
    P.arena = base + L.o_arena; P.arena_stride = L.arena_stride; P.arena_bytes = L.arena_bytes; P.arena_slots = L.arena_slots;
    P.sol = (int64_t *)(base + L.o_sol);
    P.best_u = (int64_t *)(base + L.o_bu); P.best_key = (ord_t *)(base + L.o_bk); P.best_valid = (int *)(base + L.o_bv);
    P.out_list = (int *)(base + L.o_out);
    P.warps_per_cta = L.wpc;
    P.warp_smem_doubles = L.warp_bytes / 8;
    P.lowwater = std::min(L.nwarps, L.pool_cap / 4);
    P.eager_max = EAGER_MAX;
    P.dive_hold = getenv("BLTCU_NO_DIVE_HOLD") ? 0 : 1;
}
}  // namespace

// the kernels are compiled once per size class of m (32, 64, 96): see SL<MC>
static cudaError_t launch_search(int m, int ctas, int threads, size_t smem, cudaStream_t st, const Params &P) {
    cudaError_t e;
    switch (size_class(m)) {
    case 32:
        if ((e = cudaFuncSetAttribute(bltcu_search_kernel<32>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)) != cudaSuccess) return e;
        bltcu_search_kernel<32><<<ctas, threads, smem, st>>>(P);
        break;
    case 64:
        if ((e = cudaFuncSetAttribute(bltcu_search_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)) != cudaSuccess) return e;
        bltcu_search_kernel<64><<<ctas, threads, smem, st>>>(P);
        break;
    default:
        if ((e = cudaFuncSetAttribute(bltcu_search_kernel<96>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)) != cudaSuccess) return e;
        bltcu_search_kernel<96><<<ctas, threads, smem, st>>>(P);
        break;
    }
    return cudaGetLastError();
}
static cudaError_t launch_batch(int m, int ctas, int threads, size_t smem, cudaStream_t st, const BatchParams &BP) {
    cudaError_t e;
    switch (size_class(m)) {
    case 32:
        if ((e = cudaFuncSetAttribute(bltcu_batch_kernel<32>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)) != cudaSuccess) return e;
        bltcu_batch_kernel<32><<<ctas, threads, smem, st>>>(BP);
        break;
    case 64:
        if ((e = cudaFuncSetAttribute(bltcu_batch_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)) != cudaSuccess) return e;
        bltcu_batch_kernel<64><<<ctas, threads, smem, st>>>(BP);
        break;
    default:
        if ((e = cudaFuncSetAttribute(bltcu_batch_kernel<96>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)) != cudaSuccess) return e;
        bltcu_batch_kernel<96><<<ctas, threads, smem, st>>>(BP);
        break;
    }
    return cudaGetLastError();
}


// ------------------------------------------------------------------------------------------------------------
// Work sharing between the GPUs of one search (bltcu_opts.stop_flag pointing at a block larger than the 16 legacy
// ints, its size in bytes in word 2).  All traffic is host-relayed: a GPU that runs dry asks (need[r]); the host
// thread of a GPU whose warps are all busy grants export tickets to its device, whose warps put the oldest record of
// their stacks into the named pool slots; the host copies the record out, through the destination's mailbox in the
// block, and the destination's host thread copies it into a reserved pool slot of its own device and appends it to
// the feed the idle warps pop.  Records are self-contained (every rank has built the same geometry), so they can
// be expanded on any GPU.  Termination: every rank idle (device empty, feed empty, no claimed ticket unprocessed,
// mailbox empty) and exported == imported, read twice around the scan of the idle flags.
// Block layout in ints: [0] stop word, [1] on-demand counter, [2] bytes of the block, [4] exported, [5] imported,
// [6] done, [16+r] idle, [32+r] need, [48+r] mailbox head, [64+r] mailbox tail, [80+r] left; mailboxes from byte 4096:
// per rank `slots` entries of 64 bytes of header (int seq) + one record.
struct Share {
    volatile int *w = nullptr;
    int world = 0, rank = 0, slots = 0;
    size_t stride = 0, entry = 0;
    unsigned char *mail = nullptr;
    bool on() const { return slots > 0; }
    static constexpr int MAXW = 16;
    void attach(volatile int *block, int world_, int rank_, size_t rec_stride) {
        w = block; world = world_; rank = rank_; stride = rec_stride; slots = 0;
        // opt-in (BLTCU_SHARE=1): correct at 2 and 8 GPUs, but on the bench tree it moves ~1,000-node records at the
        // ~25 records/ms a host thread can relay and does not shorten the search (profiles/r2_feed_policy.md, section 6)
        if (!block || world_ < 2 || world_ > MAXW || !getenv("BLTCU_SHARE") || atoi(getenv("BLTCU_SHARE")) == 0) return;
        const long long bytes = (long long)(unsigned int)__atomic_load_n((int *)(block + 2), __ATOMIC_ACQUIRE);
        entry = 64 + ((rec_stride + 63) / 64) * 64;
        const long long avail = (bytes - 4096) / (long long)world_ / (long long)entry;
        if (bytes < 4096 || avail < 2) return;
        slots = (int)std::min<long long>(8, avail);
        mail = (unsigned char *)block + 4096;
    }
    int *exported() const { return (int *)(w + 4); }
    int *imported() const { return (int *)(w + 5); }
    int *done() const { return (int *)(w + 6); }
    int *idle(int r) const { return (int *)(w + 16 + r); }
    int *need(int r) const { return (int *)(w + 32 + r); }
    int *mhead(int r) const { return (int *)(w + 48 + r); }
    int *mtail(int r) const { return (int *)(w + 64 + r); }
    int *left(int r) const { return (int *)(w + 80 + r); }
    unsigned char *slot(int r, unsigned int ticket) const { return mail + ((size_t)r * slots + ticket % (unsigned int)slots) * entry; }
    void leave() {  // this rank takes no further part: it counts as idle, asks for nothing
        if (!on()) return;
        __atomic_store_n(need(rank), 0, __ATOMIC_SEQ_CST);
        __atomic_store_n(idle(rank), 1, __ATOMIC_SEQ_CST);
        __atomic_store_n(left(rank), 1, __ATOMIC_SEQ_CST);
    }
    bool all_idle() const {
        for (int r = 0; r < world; ++r) if (!__atomic_load_n(idle(r), __ATOMIC_SEQ_CST)) return false;
        return true;
    }
};
struct ShareGuard {  // whatever way search_impl returns, the other ranks must not wait for this one
    Share *s;
    ~ShareGuard() { if (s) s->leave(); }
};

static int search_impl(DevWs *w, const int64_t *d_lat, const int64_t *d_con, int n, int m, const bltcu_opts *opts,
                       int64_t *sol, bltcu_stats *stats) {
    const auto t_start = std::chrono::steady_clock::now();
    if (stats) memset(stats, 0, sizeof *stats);  // every early return leaves zeros, not the caller's stack
    bltcu_opts o;
    memset(&o, 0, sizeof o);
    if (opts) o = *opts;
    const int world = o.world > 1 ? o.world : 1;
    const int rank = world > 1 ? o.rank : 0;
    if (world > 1 && o.split_depth <= 0) o.split_depth = std::min(6, std::max(1, n - 3));
    if (o.split_depth > n - 2) o.split_depth = std::max(0, n - 2);

    Layout L;
    if (make_layout(L, n, m, w->smem_optin, w->sms, o.warps_per_cta, o.ctas, 12, 16384, 64)) {
        fprintf(stderr, "bltcu: problem %d x %d does not fit the shared-memory budget (%zu bytes)\n", n, m, w->smem_optin);
        return BLTCU_STATUS_ERROR;
    }
    if (ws_reserve(w, L.total)) return BLTCU_STATUS_ERROR;
    unsigned char *base = (unsigned char *)w->buf;
    cudaStream_t st = w->stream;
    const int ctas = L.ctas, wpc = L.wpc, nwarps = L.nwarps, pool_cap = L.pool_cap;
    const unsigned int ring_cap = (unsigned int)pool_cap;
    const RecLayout RL = L.RL;
    const size_t smem_bytes = L.smem_bytes;

    Params P;
    fill_params(P, L, base, n, m);
    P.lat = d_lat; P.con = d_con;
    P.donate_min_level = o.donate_min_level > 0 ? std::max(2, o.donate_min_level) : std::min(6, std::max(2, n / 3));
    P.no_early_exit = o.no_early_exit;
    if (o.eager > 0) P.eager_max = o.eager;
    // The stop word of a multi-rank search is plain host memory shared by the ranks' host threads (or processes).  GPUs
    // never touch it: while its kernel runs, each rank's host thread looks at its own device's `found` through a second
    // stream, raises the word when it is set, and writes `abort` into its device when somebody else raised the word.
    // Word 1 of the same block is a counter from which the ranks take the last eighth of the static split's subtrees one
    // at a time, whoever has idle warps first; the host thread writes the record index into the device's feed (`dyn`).
    volatile int *shared = (world > 1 && o.stop_flag) ? (volatile int *)o.stop_flag : nullptr;
    volatile int *host_stop = (shared && !o.no_early_exit) ? shared : nullptr;
    std::vector<int> dyn;   // record slots of the subtrees dealt on demand, in depth-first order
    bool feeding = false;
    Share share;            // work sharing between the GPUs (a block larger than the legacy 16 ints)
    share.attach(shared, world, rank, RL.stride);
    ShareGuard share_guard{&share};
    if (share.on() && w->rec_cap < 2 * RL.stride) {
        if (w->rec_pin) cudaFreeHost(w->rec_pin);
        w->rec_pin = nullptr; w->rec_cap = 0;
        CK(cudaHostAlloc((void **)&w->rec_pin, 2 * RL.stride, cudaHostAllocDefault));
        w->rec_cap = 2 * RL.stride;
    }
    std::vector<int> resv;  // pool slots kept out of the free ring: imports first, then SHARE_EXPORT_SLOTS for exports
    int want_resv = 0;

    int launches = 0;
    const size_t setup_smem = warp_smem_bytes(n, m, 0, 1);
    CK(cudaFuncSetAttribute(bltcu_setup_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)setup_smem));
    bltcu_setup_kernel<<<1, 32, setup_smem, st>>>(d_lat, d_con, n, m, (double *)P.B, (double *)P.p, (double *)(base + L.o_Q),
                                                 (double *)P.Qrev, (double *)P.Rm, (double *)P.QTp, (double *)(base + L.o_tmp),
                                                 (unsigned long long *)(base + L.o_rc), (int *)(base + L.o_rc + 8), P.wbuf,
                                                 P.pool);
    CK(cudaGetLastError());
    launches++;
    // radius: closest.cc:193-194 uses row 0 only; device clock for the watchdog deadline; status of the geometry
    int64_t con0[2];
    unsigned long long now_rc[6] = {0, 0, 0, 0, 0, 0};
    unsigned long long deadline_ns = 0;
    CK(cudaMemcpyAsync(con0, d_con, 16, cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(now_rc, base + L.o_rc, 48, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    if ((int)(now_rc[1] & 0xffffffffu) != 0) return BLTCU_STATUS_LP_FAIL;  // dependent basis rows
    P.radius = (double)(con0[1] - con0[0]) / 2.0;
    P.radius_lp = o.sound ? P.radius * (1.0 + SOUND_MARGIN) : P.radius;
    {
        long long tmo = o.timeout_ms;
        if (tmo <= 0) { const char *e = getenv("BLTCU_TIMEOUT_MS"); if (e) tmo = atoll(e); }
        if (tmo > 0) deadline_ns = now_rc[0] + (unsigned long long)tmo * 1000000ull;
    }

    // host images of the control block and rings
    std::vector<int> ring(ring_cap, -1), fring(ring_cap);
    QCtl hq;
    auto reset_ctl = [&](const std::vector<int> &initial) -> int {
        memset(&hq, 0, sizeof hq);
        std::fill(ring.begin(), ring.end(), -1);
        for (size_t i = 0; i < initial.size(); ++i) ring[i] = initial[i];
        hq.tail = (unsigned int)initial.size();
        hq.pending = (int)initial.size();
        std::fill(fring.begin(), fring.end(), -1);
        unsigned int nf = 0;
        std::vector<char> used(pool_cap, 0);
        for (int r : initial) used[r] = 1;
        for (int r : dyn) used[r] = 1;  // records waiting to be fed stay where they are
        resv.clear();
        for (int i = pool_cap - 1; i >= 0 && (int)resv.size() < want_resv; --i) if (!used[i]) { used[i] = 1; resv.push_back(i); }
        for (int i = 0; i < pool_cap; ++i) if (!used[i]) fring[nf++] = i;
        hq.ftail = nf;
        hq.deadline_ns = deadline_ns;
        CK(cudaMemcpyAsync(P.q, &hq, sizeof hq, cudaMemcpyHostToDevice, st));
        CK(cudaMemcpyAsync(P.ring, ring.data(), ring_cap * 4, cudaMemcpyHostToDevice, st));
        CK(cudaMemcpyAsync(P.fring, fring.data(), ring_cap * 4, cudaMemcpyHostToDevice, st));
        return 0;
    };
    CK(cudaMemsetAsync(P.best_valid, 0, (size_t)nwarps * 4, st));
    CK(cudaMemsetAsync(P.sol, 0, (size_t)n * 8, st));

    float kernel_ms = 0.f;
    unsigned int fed_total = 0;
    bool lp_failed = false;
    const bool verbose = getenv("BLTCU_VERBOSE") != nullptr;
    QCtl acc;  // totals
    memset(&acc, 0, sizeof acc);
    auto add_counts = [&](const QCtl &x) {
        acc.niter += x.niter; acc.nprune += x.nprune; acc.nodes += x.nodes; acc.pivots += x.pivots; acc.quick += x.quick;
        acc.refactors += x.refactors; acc.colds += x.colds; acc.donated += x.donated; acc.inband += x.inband; acc.idle_ns += x.idle_ns; acc.busy_ns += x.busy_ns;
        for (int i = 0; i <= n; ++i) acc.hist[i] += x.hist[i];
    };
    auto launch = [&]() -> int {
        CK(cudaEventRecord(w->e0, st));
        CK(launch_search(m, ctas, wpc * 32, smem_bytes, st, P));
        CK(cudaEventRecord(w->e1, st));
        if (host_stop || feeding) {
            bool told = false, closed = !feeding;
            unsigned int fed = 0;
            const auto t_launch = std::chrono::steady_clock::now();
            double closed_ms = -1.0, first_hungry_ms = -1.0;
            struct Sample { double t; int idle, pending, qlen, fed; };
            std::vector<Sample> samples;
            const bool trace = getenv("BLTCU_TRACE") != nullptr;  // developer aid: the relay's looks at the device, binned
            int hungry_polls = 0, feed_polls = 1, feed_idle_div = 8;
            if (const char *e = getenv("BLTCU_FEED_POLLS")) feed_polls = std::max(1, atoi(e));
            if (const char *e = getenv("BLTCU_FEED_IDLE_DIV")) feed_idle_div = std::max(1, atoi(e));
            bool ramped = getenv("BLTCU_FEED_RAMP_GUARD") == nullptr, seen_idle = false;  // the guard is off unless asked for
            bool fed_state = getenv("BLTCU_FEED_EAGER") != nullptr;  // (eager: top up from the start, for measurements)
            int feed_per_idle = 0, feed_hold_us = 150, feed_backlog = 0;
            if (const char *e = getenv("BLTCU_FEED_BACKLOG")) { feed_backlog = std::max(0, atoi(e)); if (feed_backlog > 0) { ramped = false; feed_hold_us = 0; } }
            if (const char *e = getenv("BLTCU_FEED_PER_IDLE")) feed_per_idle = std::max(0, atoi(e));
            if (const char *e = getenv("BLTCU_FEED_HOLD_US")) feed_hold_us = std::max(0, atoi(e));
            // pin[0..11]: pending idle abort status found out_count timed_out fail_info dive_state feed_head feed_tail
            // feed_open, as laid out in QCtl; [12..13] head, tail of the ring (trace); [20..22] values to send; [32..] a chunk
            int *pin = w->pin;
            static_assert(offsetof(QCtl, feed_head) - offsetof(QCtl, pending) == 36 && offsetof(QCtl, feed_open) - offsetof(QCtl, pending) == 44, "relay snapshot layout");
            static_assert(offsetof(QCtl, export_claim) - offsetof(QCtl, pending) == 52, "relay snapshot layout");
            // ---- work sharing between the GPUs (see struct Share) ----
            const bool sharing = share.on() && P.exp_list != nullptr;
            std::vector<int> imp_slots, exp_free, ticket_dst, ticket_slot;
            size_t imp_next = 0;
            int granted = 0, processed = 0, n_exported = 0, n_imported = 0, n_requests = 0;
            unsigned int my_mhead = 0;
            auto t_last_req = t_launch;
            if (sharing) {
                const int ne = std::min<int>(SHARE_EXPORT_SLOTS, (int)resv.size() / 2);
                exp_free.assign(resv.end() - ne, resv.end());
                imp_slots.assign(resv.begin(), resv.end() - ne);
                ticket_dst.assign(EXP_TICKETS, -1); ticket_slot.assign(EXP_TICKETS, -1);
                my_mhead = (unsigned int)__atomic_load_n(share.mhead(rank), __ATOMIC_ACQUIRE);
                ramped = false;
            }
            const int epoch = sharing ? __atomic_load_n((int *)(share.w + 7), __ATOMIC_ACQUIRE) : 0;
            auto seq_of = [&](unsigned int ticket) { return (int)((unsigned int)epoch * 0x10000u + ticket + 1u); };
            // a record from this rank's mailbox into a reserved slot of the device's pool, then into the feed
            auto import_step = [&]() -> int {
                for (;;) {
                    if ((int)((unsigned int)__atomic_load_n(share.mtail(rank), __ATOMIC_ACQUIRE) - my_mhead) <= 0) break;
                    unsigned char *e = share.slot(rank, my_mhead);
                    if (__atomic_load_n((int *)e, __ATOMIC_ACQUIRE) != seq_of(my_mhead)) break;  // claimed, not written yet
                    if (imp_next >= imp_slots.size()) break;  // no slot left: the record waits (nothing more is asked for)
                    __atomic_store_n(share.idle(rank), 0, __ATOMIC_SEQ_CST);
                    unsigned char *stage = w->rec_pin + RL.stride;
                    memcpy(stage, e + 64, RL.stride);
                    my_mhead++;
                    __atomic_store_n(share.mhead(rank), (int)my_mhead, __ATOMIC_RELEASE);
                    const int slot = imp_slots[imp_next++];
                    pin[24] = slot; pin[21] = (int)fed + 1;
                    CK(cudaMemcpyAsync(P.pool + (size_t)slot * RL.stride, stage, RL.stride, cudaMemcpyHostToDevice, w->stream2));
                    CK(cudaMemcpyAsync(P.feed + fed, pin + 24, 4, cudaMemcpyHostToDevice, w->stream2));
                    CK(cudaMemcpyAsync(&P.q->feed_tail, pin + 21, 4, cudaMemcpyHostToDevice, w->stream2));
                    CK(cudaStreamSynchronize(w->stream2));
                    fed++; n_imported++;
                    __atomic_fetch_add(share.imported(), 1, __ATOMIC_SEQ_CST);
                }
                return 0;
            };
            // the record of export ticket t out of the device and into the mailbox of the rank that asked for it
            auto deliver = [&](int t) -> int {
                const int slot = ticket_slot[t], dst = ticket_dst[t];
                CK(cudaMemcpyAsync(w->rec_pin, P.pool + (size_t)slot * RL.stride, RL.stride, cudaMemcpyDeviceToHost, w->stream2));
                CK(cudaStreamSynchronize(w->stream2));
                exp_free.push_back(slot);
                if ((host_stop && *host_stop) || __atomic_load_n(share.left(dst), __ATOMIC_SEQ_CST)) return 0;  // the search is over
                __atomic_fetch_add(share.exported(), 1, __ATOMIC_SEQ_CST);
                const unsigned int ticket = (unsigned int)__atomic_fetch_add(share.mtail(dst), 1, __ATOMIC_SEQ_CST);
                // room in the ring of `slots` entries; while waiting this rank keeps serving its own mailbox (no cycles)
                while ((int)(ticket - (unsigned int)__atomic_load_n(share.mhead(dst), __ATOMIC_ACQUIRE)) >= share.slots) {
                    if ((host_stop && *host_stop) || __atomic_load_n(share.left(dst), __ATOMIC_SEQ_CST)) return 0;
                    if (import_step()) return 1;
                    std::this_thread::sleep_for(std::chrono::microseconds(20));
                }
                unsigned char *e = share.slot(dst, ticket);
                memcpy(e + 64, w->rec_pin, RL.stride);
                __atomic_store_n((int *)e, seq_of(ticket), __ATOMIC_RELEASE);
                n_exported++;
                return 0;
            };
            auto share_poll = [&]() -> int {
                const int pending = pin[0], idle = pin[1], claim = pin[13];
                // exports the device has completed (the list is read after the snapshot, on the same stream)
                while (processed < claim) {
                    const int cnt = std::min(claim - processed, 960);
                    CK(cudaMemcpyAsync(pin + 32, P.exp_list + processed, (size_t)cnt * 4, cudaMemcpyDeviceToHost, w->stream2));
                    CK(cudaStreamSynchronize(w->stream2));
                    int k = 0;
                    while (k < cnt && pin[32 + k]) { if (deliver(processed)) return 1; processed++; k++; }
                    if (k < cnt) break;  // a warp has its ticket but is still copying
                }
                if (import_step()) return 1;
                if (told) return 0;
                const auto now = std::chrono::steady_clock::now();
                if (pin[1] > 0) seen_idle = true;
                if (!ramped && ((seen_idle && idle * 8 < nwarps) || std::chrono::duration<double, std::milli>(now - t_launch).count() > 5.0)) ramped = true;
                const bool mailbox_empty = (unsigned int)__atomic_load_n(share.mtail(rank), __ATOMIC_ACQUIRE) == my_mhead;
                const bool is_done = __atomic_load_n(share.done(), __ATOMIC_SEQ_CST) != 0;
                // ask: an eighth of the warps wait and nothing is on its way
                const bool hungry = ramped && idle * 8 > nwarps && !is_done;
                const int asked = __atomic_load_n(share.need(rank), __ATOMIC_SEQ_CST);
                if (hungry && asked == 0 && mailbox_empty && imp_slots.size() - imp_next >= 32 &&
                    std::chrono::duration<double, std::micro>(now - t_last_req).count() > 200.0) {
                    __atomic_store_n(share.need(rank), std::min(16, std::max(2, idle / 64)), __ATOMIC_SEQ_CST);
                    t_last_req = now; n_requests++;
                } else if (!hungry && asked > 0) __atomic_store_n(share.need(rank), 0, __ATOMIC_SEQ_CST);
                // give: every warp busy, records queued beyond them, few tickets still unclaimed
                if (idle * 16 < nwarps && pending >= nwarps && granted - claim < 4 && !is_done) {
                    for (int k = 1; k < world; ++k) {
                        const int d = (rank + k) % world;
                        int nd = __atomic_load_n(share.need(d), __ATOMIC_SEQ_CST);
                        if (nd <= 0 || __atomic_load_n(share.left(d), __ATOMIC_SEQ_CST)) continue;
                        const int c = std::min(std::min(nd, 2), std::min((int)exp_free.size(), EXP_TICKETS - granted));
                        if (c <= 0) break;
                        if (!__atomic_compare_exchange_n(share.need(d), &nd, nd - c, false, __ATOMIC_SEQ_CST, __ATOMIC_SEQ_CST)) continue;
                        for (int i = 0; i < c; ++i) {
                            const int slot = exp_free.back(); exp_free.pop_back();
                            ticket_dst[granted + i] = d; ticket_slot[granted + i] = slot; pin[24 + i] = slot;
                        }
                        pin[23] = granted + c;
                        CK(cudaMemcpyAsync(P.exp_slots + granted, pin + 24, (size_t)c * 4, cudaMemcpyHostToDevice, w->stream2));
                        CK(cudaMemcpyAsync(&P.q->export_want, pin + 23, 4, cudaMemcpyHostToDevice, w->stream2));
                        CK(cudaStreamSynchronize(w->stream2));
                        granted += c;
                        break;
                    }
                }
                // idle flag and the end of the search
                const bool my_idle = pending == 0 && fed == (unsigned int)pin[9] && processed == claim && mailbox_empty;
                __atomic_store_n(share.idle(rank), my_idle ? 1 : 0, __ATOMIC_SEQ_CST);
                if (my_idle && !is_done) {
                    const int e1 = __atomic_load_n(share.exported(), __ATOMIC_SEQ_CST);
                    const int i1 = __atomic_load_n(share.imported(), __ATOMIC_SEQ_CST);
                    if (e1 == i1 && share.all_idle() && __atomic_load_n(share.exported(), __ATOMIC_SEQ_CST) == e1)
                        __atomic_store_n(share.done(), 1, __ATOMIC_SEQ_CST);
                }
                if (__atomic_load_n(share.done(), __ATOMIC_SEQ_CST) && !closed) {
                    pin[22] = 0;
                    CK(cudaMemcpyAsync(&P.q->feed_open, pin + 22, 4, cudaMemcpyHostToDevice, w->stream2));
                    CK(cudaStreamSynchronize(w->stream2));
                    closed = true;
                    closed_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_launch).count();
                }
                return 0;
            };
            CK(cudaStreamWaitEvent(w->stream2, w->e0, 0));  // the snapshots below must see this launch's control block
            while (cudaEventQuery(w->e1) == cudaErrorNotReady) {
                CK(cudaMemcpyAsync(pin, &P.q->pending, 56, cudaMemcpyDeviceToHost, w->stream2));
                if (trace) CK(cudaMemcpyAsync(pin + 14, &P.q->head, 8, cudaMemcpyDeviceToHost, w->stream2));
                CK(cudaStreamSynchronize(w->stream2));
                if (trace) {
                    const double t = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_launch).count();
                    samples.push_back({t, pin[1], pin[0], (int)((unsigned int)pin[15] - (unsigned int)pin[14]), (int)fed});
                }
                if (host_stop && pin[4]) *host_stop = 1;
                if (host_stop && *host_stop && !told) {
                    pin[20] = 1;
                    CK(cudaMemcpyAsync(&P.q->abort, pin + 20, 4, cudaMemcpyHostToDevice, w->stream2));
                    CK(cudaStreamSynchronize(w->stream2));
                    told = true;
                }
                if (sharing) {
                    if (share_poll()) return BLTCU_STATUS_ERROR;
                    std::this_thread::sleep_for(std::chrono::microseconds(40));
                    continue;
                }
                // hungry: an eighth of the warps wait for work (on BLTCU_FEED_POLLS looks in a row, default 1).  Measured
                // (profiles/r2_feed_policy.md): stricter tests, larger chunks and larger on-demand parts all lose, because
                // a subtree that starts late cannot finish before its own critical path has run (tens of ms here).
                // BLTCU_FEED_RAMP_GUARD: while a search ramps up every warp looks idle; with the guard nothing is handed out
                // before the warps have been busy once (or 5 ms have passed).  Measured equal at 8 GPUs, and it would delay
                // small searches, so it is off by default.
                if (pin[1] > 0) seen_idle = true;  // the kernel has started
                if (!ramped && ((seen_idle && pin[1] * 8 < nwarps) ||
                                std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_launch).count() > 5.0)) ramped = true;
                hungry_polls = (ramped && pin[1] * feed_idle_div > nwarps) ? hungry_polls + 1 : 0;
                // backlog policy (BLTCU_FEED_BACKLOG=k): keep k subtrees waiting in the feed, whether or not anybody is idle:
                // a warp that runs dry with the ring empty takes one at once instead of waiting for the next look
                const int backlog = (int)(fed - (unsigned int)pin[9]);
                // ... from the moment this rank's own share first leaves an eighth of its warps without work
                if (ramped && pin[1] * feed_idle_div > nwarps) fed_state = true;
                const bool top_up = feed_backlog > 0 && fed_state && backlog < feed_backlog;
                if (!closed && !told && (feed_backlog > 0 ? top_up : hungry_polls >= feed_polls)) {  // one more chunk
                    if (verbose && first_hungry_ms < 0.0 && fed > 0) first_hungry_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_launch).count();
                    // a chunk is 1/16 of a rank's fair share of the on-demand part (at most what the staging page holds)
                    int chunk = std::min(960, std::max(1, (int)(dyn.size() / (16 * (size_t)world))));
                    if (feed_backlog > 0) {  // fill up, but never more than half of a rank's fair share of what is left
                        const int left = std::max(0, (int)dyn.size() - __atomic_load_n((int *)(shared + 1), __ATOMIC_RELAXED));
                        chunk = std::min(960, std::max(1, std::min(feed_backlog - backlog, left / (2 * world))));
                    } else if (feed_per_idle > 0) {  // one subtree per so many waiting warps, less as the on-demand part runs out
                        const int left = std::max(0, (int)dyn.size() - __atomic_load_n((int *)(shared + 1), __ATOMIC_RELAXED));
                        chunk = std::max(chunk, std::min(pin[1] / feed_per_idle, std::max(1, left / (2 * world))));
                        chunk = std::min(960, chunk);
                    }
                    const int i0 = __atomic_fetch_add((int *)(shared + 1), chunk, __ATOMIC_RELAXED);
                    if (i0 < (int)dyn.size()) {
                        const int cnt = std::min(chunk, (int)dyn.size() - i0);
                        for (int e = 0; e < cnt; ++e) pin[32 + e] = dyn[(size_t)(i0 + e)];
                        pin[21] = (int)fed + cnt;
                        CK(cudaMemcpyAsync(P.feed + fed, pin + 32, (size_t)cnt * 4, cudaMemcpyHostToDevice, w->stream2));
                        CK(cudaMemcpyAsync(&P.q->feed_tail, pin + 21, 4, cudaMemcpyHostToDevice, w->stream2));
                        fed += (unsigned int)cnt;
                    } else {
                        pin[22] = 0;
                        CK(cudaMemcpyAsync(&P.q->feed_open, pin + 22, 4, cudaMemcpyHostToDevice, w->stream2));
                        closed = true;
                        closed_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_launch).count();
                    }
                    CK(cudaStreamSynchronize(w->stream2));
                    // one chunk per 150 us at most: while a search ramps up every warp looks idle, and a rank must not
                    // swallow the whole on-demand half in those first milliseconds
                    std::this_thread::sleep_for(std::chrono::microseconds(feed_hold_us));
                    continue;
                }
                std::this_thread::sleep_for(std::chrono::microseconds(50));
            }
            (void)cudaGetLastError();
            fed_total += fed;
            if (sharing) {
                share.leave();
                if (verbose) fprintf(stderr, "bltcu: rank %d asked %d times, took in %d records, handed out %d (%d tickets granted), feed closed after %.2f ms\n",
                                     rank, n_requests, n_imported, n_exported, granted, closed_ms);
            }
            if (trace && !samples.empty()) {
                const double bin = std::max(1.0, samples.back().t / 40.0);
                size_t i = 0;
                for (double t0 = 0.0; i < samples.size(); t0 += bin) {
                    long cnt = 0; double idle = 0, pend = 0, ql = 0; int fedv = 0;
                    for (; i < samples.size() && samples[i].t < t0 + bin; ++i) { cnt++; idle += samples[i].idle; pend += samples[i].pending; ql += samples[i].qlen; fedv = samples[i].fed; }
                    if (cnt) fprintf(stderr, "bltcu: trace rank %d t %7.1f ms idle %6.0f of %d warps, pending %7.0f, ring %6.0f, fed %d\n", rank, t0, idle / cnt, nwarps, pend / cnt, ql / cnt, fedv);
                }
            }
            if (verbose && feeding && !sharing)
                fprintf(stderr, "bltcu: rank %d took %u of %zu on-demand subtrees, second request after %.2f ms, none left after %.2f ms\n", rank, fed,
                        dyn.size(), first_hungry_ms, closed_ms);
        }
        CK(cudaMemcpyAsync(&hq, P.q, sizeof hq, cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));
        if (host_stop && hq.found) *host_stop = 1;
        float ms = 0.f;
        CK(cudaEventElapsedTime(&ms, w->e0, w->e1));
        kernel_ms += ms;
        launches++;
        if (verbose) fprintf(stderr, "bltcu: rank %d search launch %d (static depth %d): %.3f ms, %llu nodes, busy %.3f, %u subtrees published, first descent over after %.3f ms\n", rank, launches, P.static_depth, ms, (unsigned long long)hq.nodes, (double)hq.busy_ns / (double)std::max<unsigned long long>(1, hq.busy_ns + hq.idle_ns), hq.out_count, hq.dive_end_ns ? (double)(hq.dive_end_ns - now_rc[0]) * 1e-6 : -1.0);
        if (hq.fail_info) lp_failed = true;
        if (hq.fail_info && verbose) {
            fprintf(stderr, "bltcu: LP failure code %d at level k=%d (basis size %d), u =", hq.fail_info & 255, (hq.fail_info >> 8) & 255, hq.fail_info >> 16);
            for (int e = ((hq.fail_info >> 8) & 255) - 1; e < n; ++e) fprintf(stderr, " %lld", hq.fail_u[e]);
            fprintf(stderr, "\n");
        }
        if (hq.timed_out)
            fprintf(stderr, "bltcu: watchdog fired (where=%d) pending=%d idle=%d head=%u tail=%u fhead=%u ftail=%u found=%d\n",
                    hq.timed_out, hq.pending, hq.idle, hq.head, hq.tail, hq.fhead, hq.ftail, hq.found);
        return 0;
    };

    int status = BLTCU_STATUS_OK, found = 0;
    std::vector<int> initial;
    if (rank == 0) {  // the initial solve LP(n) done by the setup kernel
        acc.nodes += 1; acc.hist[n] += 1;
        acc.pivots += now_rc[2]; acc.colds += now_rc[3]; acc.refactors += now_rc[4]; acc.quick += now_rc[5];
    }
    if (o.split_depth > 0) {
        // phase A: deterministic static split of the top levels (every rank computes the same list)
        // The depth is lowered until the published subtrees fill at most half of the record pool: the count is a property
        // of the tree, so every rank settles on the same depth (a run can only run out of records above that count).
        for (int depth = o.split_depth;; --depth) {
            // the top-level record written by the setup kernel sits in slot 0; a failed attempt has consumed it
            if (depth != o.split_depth) {
                bltcu_setup_kernel<<<1, 32, setup_smem, st>>>(d_lat, d_con, n, m, (double *)P.B, (double *)P.p, (double *)(base + L.o_Q),
                                                             (double *)P.Qrev, (double *)P.Rm, (double *)P.QTp, (double *)(base + L.o_tmp),
                                                             (unsigned long long *)(base + L.o_rc), (int *)(base + L.o_rc + 8), P.wbuf,
                                                             P.pool);
                CK(cudaGetLastError());
                launches++;
                CK(cudaMemsetAsync(P.best_valid, 0, (size_t)nwarps * 4, st));
            }
            initial.assign(1, 0);
            if (reset_ctl(initial)) return BLTCU_STATUS_ERROR;
            P.static_depth = depth; P.dynamic = 0;
            lp_failed = false;
            if (launch()) return BLTCU_STATUS_ERROR;
            const bool overflow = (hq.status == BLTCU_STATUS_ERROR && !hq.timed_out) || hq.out_count > (unsigned int)pool_cap / 2;
            if (overflow && depth > 1 && !(hq.found && !o.no_early_exit)) {
                if (verbose) fprintf(stderr, "bltcu: static split at depth %d publishes too many subtrees (%u), trying %d\n", depth, hq.out_count, depth - 1);
                continue;
            }
            break;
        }
        status = hq.status; found = hq.found;
        if (rank == 0) add_counts(hq);
        std::vector<int> out(hq.out_count);
        if (hq.out_count) CK(cudaMemcpy(out.data(), P.out_list, (size_t)hq.out_count * 4, cudaMemcpyDeviceToHost));
        // order the collected subtrees by their path (DFS order) so that every rank deals the same hands
        std::vector<std::vector<ord_t>> keys(out.size(), std::vector<ord_t>(n));
        if (!out.empty()) {
            // the per-warp buffers are idle between the two phases: borrow them for the gathered keys
            ord_t *dkeys = (ord_t *)P.wbuf;
            if ((size_t)out.size() * n * sizeof(ord_t) > (size_t)nwarps * P.wbuf_stride * 8) return BLTCU_STATUS_ERROR;
            const int total = (int)out.size() * n;
            bltcu_gather_keys_kernel<<<(total + 255) / 256, 256, 0, st>>>(P.pool, RL.stride, RL.off_ord, P.out_list, (int)out.size(), n, dkeys);
            CK(cudaGetLastError());
            std::vector<ord_t> flat((size_t)total);
            CK(cudaMemcpyAsync(flat.data(), dkeys, (size_t)total * sizeof(ord_t), cudaMemcpyDeviceToHost, st));
            CK(cudaStreamSynchronize(st));
            launches++;
            for (size_t i = 0; i < out.size(); ++i) keys[i].assign(flat.begin() + i * n, flat.begin() + (i + 1) * n);
        }
        std::vector<int> idx(out.size());
        for (size_t i = 0; i < idx.size(); ++i) idx[i] = (int)i;
        std::sort(idx.begin(), idx.end(), [&](int a, int b) { return keys[a] < keys[b]; });
        initial.clear();
        // Seven eighths of the subtrees are dealt round-robin now; with a shared block the last eighth is taken on demand
        // (measured: an on-demand HALF starves the GPUs between refills).
        // Developer knobs: BLTCU_FEED_DIV=d takes 1/d of the subtrees on demand (default 8), the LAST 1/d of the
        // depth-first order unless BLTCU_FEED_STRIDED is set (then every d-th subtree: measured slower).
        int feed_div = 8;
        if (const char *e = getenv("BLTCU_FEED_DIV")) feed_div = std::min(64, std::max(2, atoi(e)));
        const bool feed_tail = getenv("BLTCU_FEED_STRIDED") == nullptr;
        size_t nstat = shared ? idx.size() - idx.size() / (size_t)feed_div : idx.size();
        if (const char *e = getenv("BLTCU_FEED_PCT")) if (shared) nstat = idx.size() - idx.size() * (size_t)std::min(95, std::max(1, atoi(e))) / 100;
        size_t dealt = 0;
        for (size_t i = 0; i < idx.size(); ++i) {
            // (with a work-sharing block every subtree is dealt now and records move between the GPUs later)
            const bool on_demand = shared && !share.on() && (feed_tail ? i >= nstat : (i % (size_t)feed_div) == (size_t)feed_div - 1);
            if (on_demand) dyn.push_back(out[idx[i]]);
            else if ((int)(dealt++ % (size_t)world) == rank) {
                initial.push_back(out[idx[i]]);
                if (getenv("BLTCU_DUMP_KEYS")) {  // developer aid: the path (sibling ordinals from the top level down) of this rank's subtrees
                    fprintf(stderr, "bltcu: subtree %zu of %zu key", i, idx.size());
                    for (int e = 0; e < n && e < 10; ++e) fprintf(stderr, " %d", (int)keys[idx[i]][e]);  // top level first
                    fprintf(stderr, "\n");
                }
            }
        }
        feeding = !dyn.empty() || share.on();
        if (share.on()) {
            want_resv = std::min(2048 + SHARE_EXPORT_SLOTS, std::max(0, (pool_cap - (int)initial.size()) / 2));
            P.exp_slots = (int *)(base + L.o_exp); P.exp_list = P.exp_slots + EXP_TICKETS;
            CK(cudaMemsetAsync(P.exp_slots, 0, (size_t)2 * EXP_TICKETS * 4, st));
        }
        const bool stop = (status != BLTCU_STATUS_OK) || (found && !o.no_early_exit && (hq.abort != 0)) || (host_stop && *host_stop);
        if (!stop && (!initial.empty() || feeding)) {
            if (reset_ctl(initial)) return BLTCU_STATUS_ERROR;
            if (feeding) { hq.feed_open = 1; P.feed = P.out_list; CK(cudaMemcpyAsync(P.q, &hq, sizeof hq, cudaMemcpyHostToDevice, st)); }
            if (found) { hq.found = 1; CK(cudaMemcpyAsync(P.q, &hq, sizeof hq, cudaMemcpyHostToDevice, st)); }
            P.static_depth = 0; P.dynamic = 1;
            if (launch()) return BLTCU_STATUS_ERROR;
            if (hq.status != BLTCU_STATUS_OK) status = hq.status;
            found = found || hq.found;
            add_counts(hq);
        }
    } else {
        initial.assign(1, 0);
        if (reset_ctl(initial)) return BLTCU_STATUS_ERROR;
        P.static_depth = 0; P.dynamic = 1;
        if (launch()) return BLTCU_STATUS_ERROR;
        status = hq.status; found = hq.found;
        add_counts(hq);
    }

    if (!found && lp_failed && status == BLTCU_STATUS_OK) status = BLTCU_STATUS_LP_FAIL;
    // ---- result: the point found, else the last recorded candidate in DFS order (uhat) ----
    std::vector<int64_t> hsol(n, 0);
    bool have_key = false;
    std::vector<ord_t> best_path;  // top level first
    if (found) {
        CK(cudaMemcpy(hsol.data(), P.sol, (size_t)n * 8, cudaMemcpyDeviceToHost));
    } else {
        std::vector<int> bv(nwarps);
        CK(cudaMemcpy(bv.data(), P.best_valid, (size_t)nwarps * 4, cudaMemcpyDeviceToHost));
        int bestw = -1;
        std::vector<ord_t> bkey(n), key(n), keys;
        bool any = false;
        for (int i = 0; i < nwarps; ++i) any = any || bv[i];
        if (any) {  // all keys in one copy
            keys.resize((size_t)nwarps * n);
            CK(cudaMemcpy(keys.data(), P.best_key, (size_t)nwarps * n * sizeof(ord_t), cudaMemcpyDeviceToHost));
        }
        for (int i = 0; i < nwarps; ++i) {
            if (!bv[i]) continue;
            key.assign(keys.begin() + (size_t)i * n, keys.begin() + (size_t)(i + 1) * n);
            std::reverse(key.begin(), key.end());
            if (bestw < 0 || key >= bkey) { bestw = i; bkey = key; }
        }
        if (bestw >= 0) CK(cudaMemcpy(hsol.data(), P.best_u + (size_t)bestw * n, (size_t)n * 8, cudaMemcpyDeviceToHost));
        if (bestw >= 0) { have_key = true; best_path = bkey; }
    }
    if (sol) memcpy(sol, hsol.data(), (size_t)n * 8);
    if (stats) {
        memset(stats, 0, sizeof *stats);
        stats->niter = acc.niter; stats->nprune = acc.nprune; stats->nodes = acc.nodes; stats->pivots = acc.pivots;
        stats->quick_prunes = acc.quick; stats->refactors = acc.refactors; stats->cold_starts = acc.colds;
        stats->donated = acc.donated + fed_total;
        stats->busy_frac = (acc.idle_ns + acc.busy_ns) ? (double)acc.busy_ns / (double)(acc.idle_ns + acc.busy_ns) : 0.0;
        for (int i = 0; i <= n; ++i) stats->hist[i] = acc.hist[i];
        stats->kernel_ms = kernel_ms;
        stats->found = found;
        stats->launches = launches;
        stats->warps = nwarps; stats->ctas = ctas;
        stats->scratch_bytes = L.total;
        stats->inband = acc.inband;
        stats->stopped = (!found && host_stop && *host_stop) ? 1 : 0;
        stats->gpus = 1;
        stats->best_valid = have_key ? 1 : 0;
        if (have_key) for (int i = 0; i < n; ++i) stats->best_key[i] = best_path[i];
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
    DevWs *w = nullptr;
    if (get_ws(opts ? opts->device : 0, &w)) return BLTCU_STATUS_ERROR;
    std::lock_guard<std::mutex> lk(w->mu);
    return search_impl(w, d_lat, d_con, n, m, opts, sol, stats);
}

static int closest_one_device(const int64_t *lat, const int64_t *con, int n, int m, const bltcu_opts *opts, int64_t *sol,
                              bltcu_stats *stats) {
    const auto t0 = std::chrono::steady_clock::now();
    DevWs *w = nullptr;
    if (get_ws(opts ? opts->device : 0, &w)) return BLTCU_STATUS_ERROR;
    std::lock_guard<std::mutex> lk(w->mu);
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

// One search on several GPUs of this process (opts.gpus / BLT_GPUS): a host thread per device, each with the share
// i % gpus of the deterministic static split (the same code path as one process per GPU with opts.rank / opts.world),
// one stop word in host memory for the first point found (see search_impl), results merged here.  No collective: the ranks
// exchange nothing but that word.
static int closest_multi_device(const int64_t *lat, const int64_t *con, int n, int m, const bltcu_opts &base, int gpus, int64_t *sol,
                                bltcu_stats *stats) {
    const auto t0 = std::chrono::steady_clock::now();
    // the ranks' shared block (bltcu_opts.stop_flag): stop word, counter of the subtrees dealt on demand, and the
    // mailboxes through which records move between the GPUs (struct Share, used when BLTCU_SHARE=1)
    const size_t share_entry = 64 + ((rec_layout(n, m).stride + 63) / 64) * 64;
    const char *share_env = getenv("BLTCU_SHARE");
    const size_t block_bytes = 4096 + ((share_env && atoi(share_env)) ? (size_t)gpus * 8 * share_entry : 0);
    std::vector<unsigned long long> block((block_bytes + 7) / 8, 0ull);
    int *flag = (int *)block.data();
    flag[2] = (block_bytes > 4096 && gpus <= Share::MAXW && block_bytes < (1ull << 31)) ? (int)block_bytes : 0;
    std::vector<bltcu_stats> st((size_t)gpus);
    std::vector<std::vector<int64_t>> sols((size_t)gpus, std::vector<int64_t>((size_t)n, 0));
    std::vector<int> rcs((size_t)gpus, BLTCU_STATUS_ERROR);
    std::vector<std::thread> th;
    for (int r = 0; r < gpus; ++r)
        th.emplace_back([&, r]() {
            bltcu_opts o = base;
            o.device = base.device + r; o.rank = r; o.world = gpus; o.gpus = 1; o.stop_flag = flag;
            rcs[r] = closest_one_device(lat, con, n, m, &o, sols[r].data(), &st[r]);
        });
    for (auto &t : th) t.join();
    int rc = BLTCU_STATUS_OK, finder = -1, last = -1;
    bool lp_fail = false;
    for (int r = 0; r < gpus; ++r) {
        if (rcs[r] == BLTCU_STATUS_ERROR || rcs[r] == BLTCU_STATUS_INPUT_ERROR) rc = rcs[r];
        if (rcs[r] == BLTCU_STATUS_LP_FAIL) lp_fail = true;
        if (st[r].found && finder < 0) finder = r;
        // the reference returns the last leaf it recorded in depth-first order: the lexicographically last path key
        if (st[r].best_valid && (last < 0 || !std::lexicographical_compare(st[r].best_key, st[r].best_key + n, st[last].best_key, st[last].best_key + n))) last = r;
    }
    if (rc == BLTCU_STATUS_OK && finder < 0 && lp_fail) rc = BLTCU_STATUS_LP_FAIL;
    const int src = finder >= 0 ? finder : last;
    if (sol) for (int e = 0; e < n; ++e) sol[e] = src >= 0 ? sols[src][e] : 0;
    if (stats) {
        memset(stats, 0, sizeof *stats);
        double busy = 0.0;
        for (int r = 0; r < gpus; ++r) {
            stats->niter += st[r].niter; stats->nprune += st[r].nprune; stats->nodes += st[r].nodes; stats->pivots += st[r].pivots;
            stats->quick_prunes += st[r].quick_prunes; stats->refactors += st[r].refactors; stats->cold_starts += st[r].cold_starts;
            stats->donated += st[r].donated; stats->inband += st[r].inband; stats->launches += st[r].launches; stats->warps += st[r].warps; stats->ctas += st[r].ctas;
            stats->scratch_bytes += st[r].scratch_bytes;
            for (int i = 0; i <= n; ++i) stats->hist[i] += st[r].hist[i];
            stats->kernel_ms = std::max(stats->kernel_ms, st[r].kernel_ms);
            busy += st[r].busy_frac;
        }
        stats->busy_frac = busy / gpus;
        stats->found = finder >= 0 ? 1 : 0;
        stats->gpus = gpus;
        if (finder < 0 && last >= 0) { stats->best_valid = 1; memcpy(stats->best_key, st[last].best_key, sizeof stats->best_key); }
        stats->total_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    }
    return rc;
}

extern "C" int bltcu_closest_ex(const int64_t *lat, const int64_t *con, int n, int m, const bltcu_opts *opts,
                                int64_t *sol, bltcu_stats *stats) {
    const int v = validate(n, m);
    if (v != BLTCU_STATUS_OK) return v;
    if (!lat || !con) return BLTCU_STATUS_INPUT_ERROR;
    bltcu_opts o;
    memset(&o, 0, sizeof o);
    if (opts) o = *opts;
    int gpus = o.gpus;
    if (gpus <= 0) { const char *e = getenv("BLT_GPUS"); gpus = e ? atoi(e) : 1; }
    if (o.world > 1) gpus = 1;  // the caller already runs one rank per GPU
    if (gpus > 1) {
        int have = 0;
        CK(cudaGetDeviceCount(&have));
        gpus = std::min(gpus, have - o.device);
    }
    if (gpus > 1 && n >= 4) return closest_multi_device(lat, con, n, m, o, gpus, sol, stats);
    return closest_one_device(lat, con, n, m, &o, sol, stats);
}

extern "C" int bltcu_closest(const int64_t *lat, const int64_t *con, int n, int m, int64_t *sol, uint64_t *niter,
                             uint64_t *nprune) {
    bltcu_stats st = {};
    const int rc = bltcu_closest_ex(lat, con, n, m, nullptr, sol, &st);
    if (rc == BLTCU_STATUS_OK || rc == BLTCU_STATUS_LP_FAIL) {
        if (niter) *niter = st.niter;
        if (nprune) *nprune = st.nprune;
    }
    return rc;
}

// One launch of the batch kernel over `mine` problems (host pointers already offset to them).  allow_help: warps
// that run out of problems take subtrees of the problems still running; helped[i] then says whether problem i was
// searched by more than one warp.  *stats is accumulated into.
static int batch_launch(DevWs *w, const int64_t *lat, const int64_t *con, int mine, int n, int m, const bltcu_opts &o, int64_t *sol,
                        int *status, uint64_t *niter, uint64_t *nprune, int *found, int *helped, int allow_help, bltcu_stats *stats) {
    const int lo = 0;

    const int KQS = QS_MAX, LDS = LDS_C, LDG = m | 1, LDQ = LDQ_C;
    const size_t warp_bytes = fixed_warp_bytes(size_class(m));
    const size_t per_warp = align_up((size_t)m * LDQ * 8 + warp_bytes, 16);
    int wpc = (int)std::min<size_t>(12, w->smem_optin / per_warp);
    if (o.warps_per_cta > 0 && o.warps_per_cta < wpc) wpc = o.warps_per_cta;
    if (wpc < 1) {
        fprintf(stderr, "bltcu: problem %d x %d does not fit the shared-memory budget (%zu bytes)\n", n, m, w->smem_optin);
        return BLTCU_STATUS_ERROR;
    }
    int ctas = o.ctas > 0 ? o.ctas : w->sms;
    if (!allow_help) ctas = std::max(1, std::min(ctas, (mine + wpc - 1) / wpc));  // with helping, warps without a problem of their own are useful
    const int nwarps = ctas * wpc;
    const RecLayout RL = rec_layout(n, m);
    size_t arena_bytes = RL.stride;
    for (int k0 = 2; k0 <= n; ++k0) arena_bytes += 2 * cont_layout(RL, m, k0).bytes;  // parks when half full
    // handed-over subtrees are expanded eagerly (several children per loop): a stack like the one-problem search's
    if (allow_help) arena_bytes = std::max(arena_bytes, (size_t)48 * RL.stride);
    arena_bytes = align_up(arena_bytes, 256);
    const int arena_slots = allow_help ? 1024 : 2 * n + 16;
    const size_t arena_stride = align_up(arena_bytes + (size_t)arena_slots * 4, 256);
    const size_t wbuf_stride = align_up((size_t)2 * m * m + (size_t)2 * m * LDG, 32);
    const size_t geo_stride = align_up(((size_t)n * m + 4 * (size_t)m + 2 * (size_t)m * m + (size_t)n * n) * 8 + 32, 256);

    size_t off = 0;
    auto carve = [&](size_t bytes) { size_t r = off; off = align_up(off + bytes, 256); return r; };
    const size_t o_ctl = carve((size_t)nwarps * sizeof(QCtl)), o_next = carve(64), o_tot = carve((5 + MAXD + 1) * 8);
    const size_t o_geo = carve((size_t)nwarps * geo_stride), o_bu = carve((size_t)nwarps * n * 8), o_bk = carve((size_t)nwarps * n * sizeof(ord_t));
    const size_t o_bv = carve((size_t)nwarps * 4), o_wbuf = carve((size_t)nwarps * wbuf_stride * 8);
    const size_t o_nrec = carve((size_t)nwarps * 2 * RL.stride), o_arena = carve((size_t)nwarps * arena_stride);
    const size_t o_lat = carve((size_t)mine * n * m * 8), o_con = carve((size_t)mine * 2 * m * 8), o_sol = carve((size_t)mine * n * 8);
    const size_t o_st = carve((size_t)mine * 4), o_fd = carve((size_t)mine * 4), o_ni = carve((size_t)mine * 8), o_np = carve((size_t)mine * 8);
    int hpool_cap = 2048;  // handed-over records in flight (power of two)
    while ((size_t)hpool_cap * RL.stride > ((size_t)1 << 30) && hpool_cap > 64) hpool_cap >>= 1;
    const size_t o_gq = carve(sizeof(QCtl)), o_hring = carve((size_t)hpool_cap * 4), o_hfring = carve((size_t)hpool_cap * 4);
    const size_t o_hpool = carve((size_t)hpool_cap * RL.stride), o_sbu = carve((size_t)nwarps * n * 8), o_sbk = carve((size_t)nwarps * n * sizeof(ord_t));
    const size_t o_sbv = carve((size_t)nwarps * 4), o_hl = carve((size_t)mine * 4);
    if (ws_reserve(w, off)) return BLTCU_STATUS_ERROR;
    unsigned char *base = (unsigned char *)w->buf;
    cudaStream_t st = w->stream;

    BatchParams BP;
    memset(&BP, 0, sizeof BP);
    Params &P = BP.P;
    P.n = n; P.m = m; P.LDQ = LDQ; P.QS = KQS; P.LDS = LDS; P.LDG = LDG;
    P.ring_mask = 0; P.rec_stride = RL.stride;
    P.wbuf = (double *)(base + o_wbuf); P.wbuf_stride = wbuf_stride;
    P.nextrec = base + o_nrec;
    P.arena = base + o_arena; P.arena_stride = arena_stride; P.arena_bytes = arena_bytes; P.arena_slots = arena_slots;
    P.best_u = (int64_t *)(base + o_bu); P.best_key = (ord_t *)(base + o_bk); P.best_valid = (int *)(base + o_bv);
    P.no_early_exit = o.no_early_exit;
    P.warps_per_cta = wpc;
    P.warp_smem_doubles = warp_bytes / 8;
    P.eager_max = 1;
    P.donate_min_level = std::min(6, std::max(2, n / 3));
    BP.lat = (const int64_t *)(base + o_lat); BP.con = (const int64_t *)(base + o_con);
    BP.count = mine;
    BP.next = (int *)(base + o_next);
    BP.geo = base + o_geo; BP.geo_stride = geo_stride;
    BP.ctl = (QCtl *)(base + o_ctl);
    BP.sol = (int64_t *)(base + o_sol); BP.status = (int *)(base + o_st); BP.found = (int *)(base + o_fd);
    BP.niter = (unsigned long long *)(base + o_ni); BP.nprune = (unsigned long long *)(base + o_np);
    BP.totals = (unsigned long long *)(base + o_tot);
    {
        long long tmo = o.timeout_ms;
        if (tmo <= 0) { const char *e = getenv("BLTCU_TIMEOUT_MS"); if (e) tmo = atoll(e); }
        BP.timeout_ns = tmo > 0 ? (unsigned long long)tmo * 1000000ull : 0ull;
    }
    BP.warp_doubles = per_warp / 8;
    const size_t smem_bytes = (size_t)wpc * per_warp;
    BP.allow_help = allow_help && nwarps > 1;
    BP.gq = (QCtl *)(base + o_gq); BP.hring = (int *)(base + o_hring); BP.hfring = (int *)(base + o_hfring);
    BP.hmask = (unsigned int)hpool_cap - 1; BP.hpool = base + o_hpool;
    BP.sbest_u = (int64_t *)(base + o_sbu); BP.sbest_key = (ord_t *)(base + o_sbk); BP.sbest_valid = (int *)(base + o_sbv);
    BP.helped = (int *)(base + o_hl);
    {
        QCtl hg;
        memset(&hg, 0, sizeof hg);
        hg.ftail = (unsigned int)hpool_cap;  // every slot is free
        hg.active = mine;
        std::vector<int> ring((size_t)hpool_cap, -1), fring((size_t)hpool_cap);
        for (int i = 0; i < hpool_cap; ++i) fring[i] = i;
        CK(cudaMemcpyAsync(BP.gq, &hg, sizeof hg, cudaMemcpyHostToDevice, st));
        CK(cudaMemcpyAsync(BP.hring, ring.data(), (size_t)hpool_cap * 4, cudaMemcpyHostToDevice, st));
        CK(cudaMemcpyAsync(BP.hfring, fring.data(), (size_t)hpool_cap * 4, cudaMemcpyHostToDevice, st));
        CK(cudaMemsetAsync(BP.helped, 0, (size_t)mine * 4, st));
        CK(cudaStreamSynchronize(st));  // the host images above go out of scope
    }

    CK(cudaMemcpyAsync(base + o_lat, lat + (size_t)lo * n * m, (size_t)mine * n * m * 8, cudaMemcpyHostToDevice, st));
    CK(cudaMemcpyAsync(base + o_con, con + (size_t)lo * 2 * m, (size_t)mine * 2 * m * 8, cudaMemcpyHostToDevice, st));
    CK(cudaMemsetAsync(base + o_next, 0, 64, st));
    CK(cudaMemsetAsync(base + o_tot, 0, (5 + MAXD + 1) * 8, st));
    // a problem whose owner stops on the watchdog never writes its results: they must not be the last batch's
    {
        std::vector<int> err((size_t)mine, BLTCU_STATUS_ERROR);
        CK(cudaMemcpyAsync(base + o_st, err.data(), (size_t)mine * 4, cudaMemcpyHostToDevice, st));
        CK(cudaMemsetAsync(base + o_fd, 0, (size_t)mine * 4, st));
        CK(cudaMemsetAsync(base + o_ni, 0, (size_t)mine * 8, st));
        CK(cudaMemsetAsync(base + o_np, 0, (size_t)mine * 8, st));
        CK(cudaMemsetAsync(base + o_sol, 0, (size_t)mine * n * 8, st));
        CK(cudaStreamSynchronize(st));
    }
    CK(cudaEventRecord(w->e0, st));
    CK(launch_batch(m, ctas, wpc * 32, smem_bytes, st, BP));
    CK(cudaEventRecord(w->e1, st));
    std::vector<unsigned long long> tot(5 + MAXD + 1), hni(mine), hnp(mine);
    std::vector<int> hfd(mine), hhl(mine);
    CK(cudaMemcpyAsync(hhl.data(), base + o_hl, (size_t)mine * 4, cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(sol + (size_t)lo * n, base + o_sol, (size_t)mine * n * 8, cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(status + lo, base + o_st, (size_t)mine * 4, cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(hfd.data(), base + o_fd, (size_t)mine * 4, cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(hni.data(), base + o_ni, (size_t)mine * 8, cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(hnp.data(), base + o_np, (size_t)mine * 8, cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(tot.data(), base + o_tot, tot.size() * 8, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    float ms = 0.f;
    CK(cudaEventElapsedTime(&ms, w->e0, w->e1));
    if (getenv("BLTCU_VERBOSE")) fprintf(stderr, "bltcu: batch launch of %d problems (help %d): %.3f ms, %d warps\n", mine, BP.allow_help, ms, nwarps);
    unsigned long long sni = 0, snp = 0;
    int nfound = 0;
    for (int i = 0; i < mine; ++i) {
        if (niter) niter[lo + i] = hni[i];
        if (nprune) nprune[lo + i] = hnp[i];
        if (found) found[lo + i] = hfd[i];
        if (helped) helped[i] = hhl[i];
        sni += hni[i]; snp += hnp[i]; nfound += hfd[i] ? 1 : 0;
    }
    if (stats) {
        stats->niter += sni; stats->nprune += snp; stats->nodes += tot[0]; stats->pivots += tot[1]; stats->quick_prunes += tot[2];
        stats->refactors += tot[3]; stats->cold_starts += tot[4];
        for (int i = 0; i <= n; ++i) stats->hist[i] += tot[5 + i];
        stats->kernel_ms += ms; stats->found += nfound; stats->launches += 1; stats->warps = nwarps; stats->ctas = ctas;
        stats->scratch_bytes = off;
    }
    return BLTCU_STATUS_OK;
}

// Batch form of the search seam ("arrays of problems, one status per problem").
extern "C" int bltcu_closest_batch(const int64_t *lat, const int64_t *con, int count, int n, int m, const bltcu_opts *opts,
                                   int64_t *sol, int *status, uint64_t *niter, uint64_t *nprune, int *found,
                                   bltcu_stats *stats) {
    const auto t0 = std::chrono::steady_clock::now();
    if (count < 0) return BLTCU_STATUS_INPUT_ERROR;
    const int v = validate(n, m);
    if (v != BLTCU_STATUS_OK) return v;
    if (stats) memset(stats, 0, sizeof *stats);
    if (count == 0) return BLTCU_STATUS_OK;
    if (!lat || !con || !sol || !status) return BLTCU_STATUS_INPUT_ERROR;
    bltcu_opts o;
    memset(&o, 0, sizeof o);
    if (opts) o = *opts;
    const int world = o.world > 1 ? o.world : 1, rank = world > 1 ? o.rank : 0;
    if (rank < 0 || rank >= world) return BLTCU_STATUS_INPUT_ERROR;
    const int lo = (int)((long long)count * rank / world), hi = (int)((long long)count * (rank + 1) / world);
    const int mine = hi - lo;
    if (mine == 0) return BLTCU_STATUS_OK;
    DevWs *w = nullptr;
    if (get_ws(o.device, &w)) return BLTCU_STATUS_ERROR;
    std::lock_guard<std::mutex> lk(w->mu);
    const size_t nm = (size_t)n * m;
    std::vector<uint64_t> ni(mine), np(mine);
    std::vector<int> fd(mine), hl(mine);
    const int allow_help = getenv("BLTCU_BATCH_NO_HELP") ? 0 : 1;
    int rc = batch_launch(w, lat + lo * nm, con + (size_t)lo * 2 * m, mine, n, m, o, sol + (size_t)lo * n, status + lo, ni.data(),
                          np.data(), fd.data(), hl.data(), allow_help, stats);
    if (rc != BLTCU_STATUS_OK) return rc;
    // A problem that was shared between warps AND has a point inside the box is searched again by one warp alone:
    // the point and the counters are then those of the depth-first order, as for every other problem of the batch.
    std::vector<int> again;
    for (int i = 0; i < mine; ++i) if (hl[i] && fd[i]) again.push_back(i);
    if (!again.empty()) {
        const int k = (int)again.size();
        std::vector<int64_t> l2(k * nm), c2((size_t)k * 2 * m), s2((size_t)k * n);
        std::vector<int> st2(k), f2(k);
        std::vector<uint64_t> ni2(k), np2(k);
        for (int j = 0; j < k; ++j) {
            memcpy(&l2[j * nm], lat + (lo + again[j]) * nm, nm * 8);
            memcpy(&c2[(size_t)j * 2 * m], con + (size_t)(lo + again[j]) * 2 * m, (size_t)2 * m * 8);
        }
        if (stats) {  // the first attempt at these problems does not count
            for (int j = 0; j < k; ++j) { stats->niter -= ni[again[j]]; stats->nprune -= np[again[j]]; stats->found -= 1; }
        }
        rc = batch_launch(w, l2.data(), c2.data(), k, n, m, o, s2.data(), st2.data(), ni2.data(), np2.data(), f2.data(), nullptr, 0, stats);
        if (rc != BLTCU_STATUS_OK) return rc;
        for (int j = 0; j < k; ++j) {
            const int i = again[j];
            memcpy(sol + (size_t)(lo + i) * n, &s2[(size_t)j * n], (size_t)n * 8);
            status[lo + i] = st2[j]; ni[i] = ni2[j]; np[i] = np2[j]; fd[i] = f2[j];
        }
    }
    for (int i = 0; i < mine; ++i) {
        if (niter) niter[lo + i] = ni[i];
        if (nprune) nprune[lo + i] = np[i];
        if (found) found[lo + i] = fd[i];
    }
    if (stats) {
        stats->donated = 0;
        for (int i = 0; i < mine; ++i) stats->donated += hl[i] ? 1 : 0;  // problems searched by more than one warp
        stats->total_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    }
    return BLTCU_STATUS_OK;
}

extern "C" int bltcu_check_soln(const int64_t *lat, const int64_t *con, int n, int m, const int64_t *u, int *ok) {
    if (n <= 0 || m <= 0 || !lat || !con || !u || !ok) return BLTCU_STATUS_INPUT_ERROR;
    DevWs *w = nullptr;
    if (get_ws(0, &w)) return BLTCU_STATUS_ERROR;
    std::lock_guard<std::mutex> lk(w->mu);
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
    DevWs *w = nullptr;
    if (get_ws(0, &w)) return BLTCU_STATUS_ERROR;
    std::lock_guard<std::mutex> lk(w->mu);
    const size_t smem = warp_smem_bytes(n, m, 0, 1);
    const size_t nb = (size_t)n * m, mm = (size_t)m * m;
    // doubles: B, p, x, dist+rc, Q, Qrev, Rm, QTp, Wg (m*LDG + m*m), tmp (2m); then u
    const size_t nd = nb + m + n + 2 + mm + mm + (size_t)n * n + m + ((size_t)m * (m | 1) + mm) + 2 * (size_t)m;
    unsigned char *d = nullptr;
    CK(cudaMalloc((void **)&d, nd * 8 + (size_t)n * 8));
    double *dB = (double *)d, *dp = dB + nb, *dx = dp + m, *dd = dx + n, *dQ = dd + 2, *dQrev = dQ + mm, *dRm = dQrev + mm;
    double *dQTp = dRm + (size_t)n * n, *dWg = dQTp + m, *dtmp = dWg + ((size_t)m * (m | 1) + mm);
    int64_t *du = (int64_t *)(dtmp + 2 * (size_t)m);
    int *drc = (int *)(dd + 1);
    CK(cudaMemcpy(dB, B, nb * 8, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(du, u, (size_t)n * 8, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(dp, p, (size_t)m * 8, cudaMemcpyHostToDevice));
    CK(cudaFuncSetAttribute(bltcu_infdist_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    bltcu_infdist_kernel<<<1, 32, smem, w->stream>>>(dB, n, m, du, k, dp, dx, dd, drc, dQ, dQrev, dRm, dQTp, dWg, dtmp);
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
        if (w->stream2) cudaStreamDestroy(w->stream2);
        if (w->pin) cudaFreeHost(w->pin);
        if (w->rec_pin) cudaFreeHost(w->rec_pin);
        delete w;
    }
    g_ws.clear();
}

extern "C" const char *bltcu_version(void) { return "blt-b200 0.2 (sm_100a)"; }

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
#define BLTCU_NOALLOC_WPC 1       // warps per CTA of this variant (fixed so that the per-warp stacks have a fixed size)
#define BLTCU_NOALLOC_CTAS 148    // at most this many CTAs
#define BLTCU_NOALLOC_POOL 256    // subtree records (power of two)
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
    for (int i = tid; i < P.n; i += nth) P.sol[i] = 0;
    for (int i = tid; i < nwarps; i += nth) P.best_valid[i] = 0;
    __threadfence();
    __syncthreads();
    if (blockIdx.x == 0) {
        if (threadIdx.x == 0) {
            P.q->tail = 1; P.q->pending = 1; P.q->ftail = P.ring_mask; P.q->deadline_ns = deadline_ns;
        }
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
    // same shared-memory and workspace layout as search_impl(), but inside the fixed array
    // Every warp's private stack has the worst-case size of a depth-first path (one suspended loop per level), so a
    // search never depends on the small shared pool for room; the number of CTAs is what gives way: as many as fit.
    Layout L;
    int ctas_fit = std::min(g_noalloc_sms[device], BLTCU_NOALLOC_CTAS);
    for (;; --ctas_fit) {
        if (ctas_fit < 1 || make_layout(L, n, m, g_noalloc_smem[device], ctas_fit, BLTCU_NOALLOC_WPC, ctas_fit, BLTCU_NOALLOC_WPC,
                                        BLTCU_NOALLOC_POOL, 0))
            return BLTCU_STATUS_ERROR;
        if (L.total <= (size_t)BLTCU_NOALLOC_BYTES) break;
    }
    if (L.total > (size_t)BLTCU_NOALLOC_BYTES) {
        fprintf(stderr, "bltcu_closest_noalloc: internal layout (%zu bytes) exceeds the scratch array\n", L.total);
        return BLTCU_STATUS_ERROR;
    }
    const int wpc = L.wpc, ctas = L.ctas, nwarps = L.nwarps;
    const size_t smem_bytes = L.smem_bytes;
    int64_t *d_lat = (int64_t *)(base + L.o_in), *d_con = d_lat + (size_t)n * m;
    Params P;
    fill_params(P, L, base, n, m);
    P.lat = d_lat; P.con = d_con;
    P.static_depth = 0; P.dynamic = 1;
    P.donate_min_level = std::min(6, std::max(2, n / 3));
    P.no_early_exit = 0;
    P.radius = (double)(con[1] - con[0]) / 2.0;  // closest.cc:193-194: row 0 only
    P.radius_lp = P.radius;

    cudaStream_t st = 0;  // legacy default stream: creating a stream would be an allocation
    CK(cudaMemcpyAsync(d_lat, lat, (size_t)n * m * 8, cudaMemcpyHostToDevice, st));
    CK(cudaMemcpyAsync(d_con, con, (size_t)2 * m * 8, cudaMemcpyHostToDevice, st));
    // also writes the top-level record into pool slot 0 (the root record of the search)
    bltcu_setup_kernel<<<1, 32, warp_smem_bytes(n, m, 0, 1), st>>>(
        d_lat, d_con, n, m, (double *)P.B, (double *)P.p, (double *)(base + L.o_Q), (double *)P.Qrev, (double *)P.Rm,
        (double *)P.QTp, (double *)(base + L.o_tmp), (unsigned long long *)(base + L.o_rc), (int *)(base + L.o_rc + 8),
        P.wbuf, P.pool);
    CK(cudaGetLastError());
    unsigned long long now_rc[2] = {0, 0};
    CK(cudaMemcpyAsync(now_rc, base + L.o_rc, 16, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    if ((int)(now_rc[1] & 0xffffffffu) != 0) return BLTCU_STATUS_LP_FAIL;
    unsigned long long deadline_ns = 0;
    if (const char *e = getenv("BLTCU_TIMEOUT_MS")) {
        const long long tmo = atoll(e);
        if (tmo > 0) deadline_ns = now_rc[0] + (unsigned long long)tmo * 1000000ull;
    }
    bltcu_noalloc_init_kernel<<<1, 256, 0, st>>>(P, deadline_ns, nwarps);  // one block: its barrier orders the clears
    CK(cudaGetLastError());
    CK(launch_search(m, ctas, wpc * 32, smem_bytes, st, P));
    // host copies live in static storage: no heap use on the host either.  The result slot, the per-warp candidates,
    // their keys and flags are adjacent in the layout: one copy brings them all.
    static QCtl hq;
    static unsigned char h_res[BLTCU_NOALLOC_CTAS * BLTCU_NOALLOC_WPC * (BLTCU_NOALLOC_MAX_DIM * (8 + sizeof(ord_t)) + 4) + 4096];
    const size_t res_bytes = L.o_bv + (size_t)nwarps * 4 - L.o_sol;
    if (res_bytes > sizeof h_res) return BLTCU_STATUS_ERROR;
    CK(cudaMemcpyAsync(&hq, P.q, sizeof hq, cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(h_res, base + L.o_sol, res_bytes, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    if (hq.timed_out) fprintf(stderr, "bltcu_closest_noalloc: watchdog fired (where=%d)\n", hq.timed_out);
    const int64_t *h_sol = (const int64_t *)h_res, *h_bu = (const int64_t *)(h_res + (L.o_bu - L.o_sol));
    const ord_t *h_bk = (const ord_t *)(h_res + (L.o_bk - L.o_sol));
    const int *h_bv = (const int *)(h_res + (L.o_bv - L.o_sol));
    int64_t hsol[BLTCU_NOALLOC_MAX_DIM] = {0};
    if (hq.found) {
        for (int e = 0; e < n; ++e) hsol[e] = h_sol[e];
    } else {  // last recorded candidate in depth-first order, as the reference returns on UNSAT
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
    return (hq.status == BLTCU_STATUS_OK && !hq.found && hq.fail_info) ? BLTCU_STATUS_LP_FAIL : hq.status;
}
