// lll_batch.cu -- batched exact LLL on the GPU: bltcu_lll_batch().
//
// for many small problems (config 5: 10 000 instances of 20 x 20; config 2: 40 x 20 of rank 20)
// the time-to-decision is the reduction, not the search.  One warp reduces one problem with the fixed-capacity exact
// arithmetic of lll_fixed.h (512-bit integers): the same operations in the same order as bltcu_lll() on the host, so the
// same reduced basis and the same U, bit for bit.  The algorithm is a chain of data-dependent steps, so problems do not
// share a warp (32 of them would execute their different branches one after the other); inside a problem the row
// operations of the size reduction and of the swaps, which are the bulk of the work, are dealt to the 32 lanes.  A problem whose numbers outgrow 512 bits is
// reduced by the arbitrary-precision host implementation instead (reported in stats->found).
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <chrono>
#include <vector>

#include "../../include/bltcu.h"
#include "lll_fixed.h"
#include "lll_fp.h"

#include <algorithm>
#include <atomic>
#include <mutex>
#include <thread>

#define LLL_W 16  // limbs of 32 bits

__global__ void __launch_bounds__(128) bltcu_lll_batch_kernel(int64_t *B, int64_t *U, int count, int r, int c, int *rank, int *status,
                                                              unsigned char *work, size_t work_stride) {
    const int w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (w >= count) return;  // whole warps leave together
    const int rk = lllfix::reduce_one<LLL_W, lllfix::WarpLanes>(B + (size_t)w * r * c, U ? U + (size_t)w * r * r : nullptr, r, c,
                                                               work + (size_t)w * work_stride);
    if ((threadIdx.x & 31) == 0) {
        rank[w] = rk < 0 ? 0 : rk;
        status[w] = rk < 0 ? 1 : 0;
    }
}

#define CKL(call)                                                                                          \
    do {                                                                                                   \
        cudaError_t e_ = (call);                                                                           \
        if (e_ != cudaSuccess) {                                                                           \
            fprintf(stderr, "bltcu: CUDA error %s at %s:%d: %s\n", cudaGetErrorName(e_), __FILE__, __LINE__, \
                    cudaGetErrorString(e_));                                                               \
            if (dbuf) cudaFree(dbuf);                                                                      \
            return BLTCU_STATUS_ERROR;                                                                     \
        }                                                                                                  \
    } while (0)

// The multiprecision kernel over a batch (every problem, or the ones the floating-point pass deferred).
static int lll_batch_exact(int64_t *B, int64_t *U, int count, int r, int c, int *rank, int device, bltcu_stats *stats) {
    const auto t0 = std::chrono::steady_clock::now();
    unsigned char *dbuf = nullptr;
    CKL(cudaSetDevice(device));
    const size_t nb = (size_t)count * r * c * 8, nu = (size_t)count * r * r * 8;
    const size_t stride = ((size_t)lllfix::bytes_needed<LLL_W>(r, c) + 255) / 256 * 256;
    auto al = [](size_t x) { return (x + 255) / 256 * 256; };
    const size_t o_b = 0, o_u = al(nb), o_rank = o_u + al(nu), o_st = o_rank + al((size_t)count * 4), o_work = o_st + al((size_t)count * 4);
    const size_t total = o_work + (size_t)count * stride;
    CKL(cudaMalloc((void **)&dbuf, total));
    cudaStream_t st = nullptr;
    CKL(cudaMemcpyAsync(dbuf + o_b, B, nb, cudaMemcpyHostToDevice, st));
    cudaEvent_t e0, e1;
    CKL(cudaEventCreate(&e0));
    CKL(cudaEventCreate(&e1));
    CKL(cudaEventRecord(e0, st));
    const int threads = 128, wpb = threads / 32;
    bltcu_lll_batch_kernel<<<(count + wpb - 1) / wpb, threads, 0, st>>>((int64_t *)(dbuf + o_b), (int64_t *)(dbuf + o_u), count, r, c,
                                                                         (int *)(dbuf + o_rank), (int *)(dbuf + o_st), dbuf + o_work, stride);
    CKL(cudaGetLastError());
    CKL(cudaEventRecord(e1, st));
    std::vector<int> status(count);
    std::vector<int64_t> hb((size_t)count * r * c), hu;
    CKL(cudaMemcpyAsync(hb.data(), dbuf + o_b, nb, cudaMemcpyDeviceToHost, st));
    if (U) { hu.resize((size_t)count * r * r); CKL(cudaMemcpyAsync(hu.data(), dbuf + o_u, nu, cudaMemcpyDeviceToHost, st)); }
    CKL(cudaMemcpyAsync(rank, dbuf + o_rank, (size_t)count * 4, cudaMemcpyDeviceToHost, st));
    CKL(cudaMemcpyAsync(status.data(), dbuf + o_st, (size_t)count * 4, cudaMemcpyDeviceToHost, st));
    CKL(cudaStreamSynchronize(st));
    float ms = 0.f;
    CKL(cudaEventElapsedTime(&ms, e0, e1));
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(dbuf);
    dbuf = nullptr;
    int on_host = 0, rc = BLTCU_STATUS_OK;
    for (int i = 0; i < count; ++i) {
        int64_t *bi = B + (size_t)i * r * c, *ui = U ? U + (size_t)i * r * r : nullptr;
        if (status[i] == 0) {
            memcpy(bi, hb.data() + (size_t)i * r * c, (size_t)r * c * 8);
            if (ui) memcpy(ui, hu.data() + (size_t)i * r * r, (size_t)r * r * 8);
        } else {  // numbers beyond 512 bits: the arbitrary-precision implementation (B still holds the input)
            ++on_host;
            const int s = bltcu_lll(bi, ui, r, c, &rank[i]);
            if (s != BLTCU_STATUS_OK) rc = s;
        }
    }
    if (stats) {
        stats->kernel_ms = ms;
        stats->launches = 1;
        stats->found = on_host;
        stats->warps = count;
        stats->scratch_bytes = total;
        stats->total_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    }
    return rc;
}


// ---- floating-point decisions, exact int64 row operations (lll_fp.h): the fast pass ---------------------------------
__global__ void __launch_bounds__(128) bltcu_lll_fp_kernel(int64_t *B, int64_t *U, int count, int r, int c, int *status, double *work,
                                                           size_t work_stride) {
    const int w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (w >= count) return;  // whole warps leave together
    const int rk = lllfp::reduce_one(B + (size_t)w * r * c, U + (size_t)w * r * r, r, c, work + (size_t)w * work_stride, lllfp::WarpLanes());
    if ((threadIdx.x & 31) == 0) status[w] = rk == r ? 0 : 1;
}

namespace {
struct FpWs {  // cached per process, grow-only: no allocation, event or stream creation on the steady-state path
    std::mutex mu;
    int device = -1;
    unsigned char *buf = nullptr;
    size_t cap = 0;
    cudaStream_t st = nullptr;
    cudaEvent_t e0 = nullptr, e1 = nullptr;
} g_fp;
}  // namespace

#define CKF(call)                                                                                          \
    do {                                                                                                   \
        cudaError_t e_ = (call);                                                                           \
        if (e_ != cudaSuccess) {                                                                           \
            fprintf(stderr, "bltcu: CUDA error %s at %s:%d: %s\n", cudaGetErrorName(e_), __FILE__, __LINE__, \
                    cudaGetErrorString(e_));                                                               \
            return BLTCU_STATUS_ERROR;                                                                     \
        }                                                                                                  \
    } while (0)

// Reduces what it can; status[i] = 0 where B/U of problem i were written, 1 where the problem was deferred (untouched).
static int lll_batch_fp(int64_t *B, int64_t *U, int count, int r, int c, int device, std::vector<int> &status, float *kernel_ms,
                        size_t *scratch) {
    std::lock_guard<std::mutex> lk(g_fp.mu);
    CKF(cudaSetDevice(device));
    if (g_fp.device != device) {
        if (g_fp.buf) { cudaFree(g_fp.buf); g_fp.buf = nullptr; g_fp.cap = 0; }
        if (!g_fp.st) { CKF(cudaStreamCreateWithFlags(&g_fp.st, cudaStreamNonBlocking)); CKF(cudaEventCreate(&g_fp.e0)); CKF(cudaEventCreate(&g_fp.e1)); }
        g_fp.device = device;
    }
    auto al = [](size_t x) { return (x + 255) / 256 * 256; };
    const size_t nb = (size_t)count * r * c * 8, nu = (size_t)count * r * r * 8;
    const size_t wstride = (lllfp::scratch_doubles(r) + 31) / 32 * 32;
    const size_t o_b = 0, o_u = al(nb), o_st = o_u + al(nu), o_work = o_st + al((size_t)count * 4);
    const size_t total = o_work + (size_t)count * wstride * 8;
    if (total > g_fp.cap) {
        if (g_fp.buf) CKF(cudaFree(g_fp.buf));
        g_fp.buf = nullptr; g_fp.cap = 0;
        CKF(cudaMalloc((void **)&g_fp.buf, total));
        g_fp.cap = total;
    }
    unsigned char *d = g_fp.buf;
    CKF(cudaMemcpyAsync(d + o_b, B, nb, cudaMemcpyHostToDevice, g_fp.st));
    CKF(cudaEventRecord(g_fp.e0, g_fp.st));
    const int threads = 128, wpb = threads / 32;
    bltcu_lll_fp_kernel<<<(count + wpb - 1) / wpb, threads, 0, g_fp.st>>>((int64_t *)(d + o_b), (int64_t *)(d + o_u), count, r, c,
                                                                          (int *)(d + o_st), (double *)(d + o_work), wstride);
    CKF(cudaGetLastError());
    CKF(cudaEventRecord(g_fp.e1, g_fp.st));
    status.assign((size_t)count, 1);
    std::vector<int64_t> hb((size_t)count * r * c), hu((size_t)count * r * r);
    CKF(cudaMemcpyAsync(status.data(), d + o_st, (size_t)count * 4, cudaMemcpyDeviceToHost, g_fp.st));
    CKF(cudaMemcpyAsync(hb.data(), d + o_b, nb, cudaMemcpyDeviceToHost, g_fp.st));
    CKF(cudaMemcpyAsync(hu.data(), d + o_u, nu, cudaMemcpyDeviceToHost, g_fp.st));
    CKF(cudaStreamSynchronize(g_fp.st));
    CKF(cudaEventElapsedTime(kernel_ms, g_fp.e0, g_fp.e1));
    for (int i = 0; i < count; ++i) {
        if (status[i] != 0) continue;
        memcpy(B + (size_t)i * r * c, hb.data() + (size_t)i * r * c, (size_t)r * c * 8);
        if (U) memcpy(U + (size_t)i * r * r, hu.data() + (size_t)i * r * r, (size_t)r * r * 8);
    }
    *scratch = total;
    return BLTCU_STATUS_OK;
}

extern "C" int bltcu_lll_batch(int64_t *B, int64_t *U, int count, int r, int c, int *rank, int device, bltcu_stats *stats) {
    const auto t0 = std::chrono::steady_clock::now();
    if (stats) memset(stats, 0, sizeof *stats);
    if (count < 0 || r <= 0 || c <= 0) return BLTCU_STATUS_INPUT_ERROR;
    if (count == 0) return BLTCU_STATUS_OK;
    if (!B || !rank) return BLTCU_STATUS_INPUT_ERROR;
    // Pass 1 (rows no more than columns, i.e. possibly independent): decisions in fp64 with margins, row operations in
    // exact int64.  Pass 2: whatever pass 1 deferred, with 512-bit integers; pass 3 inside it: the host routine.
    std::vector<int> status((size_t)count, 1);
    float fp_ms = 0.f;
    size_t fp_scratch = 0;
    if (r <= c && !getenv("BLTCU_LLL_EXACT_ONLY")) {
        const int rc = lll_batch_fp(B, U, count, r, c, device, status, &fp_ms, &fp_scratch);
        if (rc != BLTCU_STATUS_OK) return rc;
    }
    std::vector<int> rest;
    for (int i = 0; i < count; ++i) { if (status[i] == 0) rank[i] = r; else rest.push_back(i); }
    int rc = BLTCU_STATUS_OK;
    bltcu_stats st2;
    memset(&st2, 0, sizeof st2);
    int on_host_threads = 0;
    const unsigned hc = std::max(1u, std::thread::hardware_concurrency());
    if (!rest.empty() && status.size() > rest.size() && rest.size() <= 4 * (size_t)hc) {
        // A handful of stragglers of the fast pass (typically half a percent: a decision on a tie): one 512-bit warp per
        // problem would take ~100 ms for them however few they are; the host routine takes ~3 ms each, on all cores.
        std::atomic<size_t> next(0);
        std::atomic<int> bad(BLTCU_STATUS_OK);
        std::vector<std::thread> th;
        const unsigned nt = (unsigned)std::min<size_t>(hc, rest.size());
        for (unsigned t = 0; t < nt; ++t)
            th.emplace_back([&]() {
                for (size_t j = next++; j < rest.size(); j = next++) {
                    const int i = rest[j];
                    const int s = bltcu_lll(B + (size_t)i * r * c, U ? U + (size_t)i * r * r : nullptr, r, c, &rank[i]);
                    if (s != BLTCU_STATUS_OK) bad = s;
                }
            });
        for (auto &t : th) t.join();
        rc = bad.load();
        on_host_threads = (int)rest.size();
    } else if (!rest.empty()) {
        const size_t k = rest.size();
        std::vector<int64_t> b2(k * r * c), u2(k * r * r);
        std::vector<int> rk2(k, 0);
        for (size_t j = 0; j < k; ++j) memcpy(&b2[j * r * c], B + (size_t)rest[j] * r * c, (size_t)r * c * 8);
        rc = lll_batch_exact(b2.data(), u2.data(), (int)k, r, c, rk2.data(), device, &st2);
        for (size_t j = 0; j < k; ++j) {
            memcpy(B + (size_t)rest[j] * r * c, &b2[j * r * c], (size_t)r * c * 8);
            if (U) memcpy(U + (size_t)rest[j] * r * r, &u2[j * r * r], (size_t)r * r * 8);
            rank[rest[j]] = rk2[j];
        }
    }
    if (stats) {
        stats->kernel_ms = fp_ms + st2.kernel_ms;
        stats->launches = (fp_scratch ? 1 : 0) + st2.launches;
        stats->found = st2.found + on_host_threads;   // problems reduced by the host routine
        stats->donated = rest.size();             // problems the floating-point pass deferred to the 512-bit kernel
        stats->warps = count;
        stats->scratch_bytes = fp_scratch + st2.scratch_bytes;
        stats->total_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    }
    return rc;
}

// Test seam: the floating-point pass on the host (same source, one lane).  Returns 0 when it completed (B, U, rank
// written), 1 when it deferred (B, U unchanged).
extern "C" int bltcu_lll_fp_host(int64_t *B, int64_t *U, int r, int c, int *rank) {
    if (!B || !U || !rank || r <= 0 || c <= 0) return BLTCU_STATUS_INPUT_ERROR;
    std::vector<int64_t> b(B, B + (size_t)r * c), u((size_t)r * r, 0);
    std::vector<double> work(lllfp::scratch_doubles(r));
    const int rk = r <= c ? lllfp::reduce_one(b.data(), u.data(), r, c, work.data(), lllfp::HostLanes()) : -1;
    if (rk != r) return 1;
    memcpy(B, b.data(), (size_t)r * c * 8);
    memcpy(U, u.data(), (size_t)r * r * 8);
    *rank = r;
    return 0;
}
