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
#include <string.h>

#include <chrono>
#include <vector>

#include "../../include/bltcu.h"
#include "lll_fixed.h"

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

extern "C" int bltcu_lll_batch(int64_t *B, int64_t *U, int count, int r, int c, int *rank, int device, bltcu_stats *stats) {
    const auto t0 = std::chrono::steady_clock::now();
    if (stats) memset(stats, 0, sizeof *stats);
    if (count < 0 || r <= 0 || c <= 0) return BLTCU_STATUS_INPUT_ERROR;
    if (count == 0) return BLTCU_STATUS_OK;
    if (!B || !rank) return BLTCU_STATUS_INPUT_ERROR;
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
