// peaks.cu -- measures the FP64 FMA throughput of the device the search runs on.
// MEASURED_PEAKS.json carries HBM and bf16 numbers only; the search kernel is bound by the FP64 FMA pipe, so the
// roofline denominator is measured here with dependent-chain-free DFMA streams at the box's own clocks.
#include <cuda_runtime.h>
#include <stdio.h>

#include "../../include/bltcu.h"

__global__ void __launch_bounds__(256) bltcu_fp64_peak_kernel(double *out, int iters, double a, double b) {
    double x0 = threadIdx.x, x1 = x0 + 1, x2 = x0 + 2, x3 = x0 + 3, x4 = x0 + 4, x5 = x0 + 5, x6 = x0 + 6, x7 = x0 + 7;
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int r = 0; r < 16; ++r) {
            x0 = fma(x0, a, b); x1 = fma(x1, a, b); x2 = fma(x2, a, b); x3 = fma(x3, a, b);
            x4 = fma(x4, a, b); x5 = fma(x5, a, b); x6 = fma(x6, a, b); x7 = fma(x7, a, b);
        }
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = ((x0 + x1) + (x2 + x3)) + ((x4 + x5) + (x6 + x7));
}

/* Measured FP64 FMA peak in TFLOP/s (2 flops per FMA) on `device`; best of `reps` launches of >= 20 ms each. */
extern "C" int bltcu_fp64_peak(int device, int reps, double *tflops) {
    if (!tflops) return BLTCU_STATUS_INPUT_ERROR;
    if (cudaSetDevice(device) != cudaSuccess) return BLTCU_STATUS_ERROR;
    int sms = 0;
    if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device) != cudaSuccess) return BLTCU_STATUS_ERROR;
    const int blocks = sms * 8, threads = 256;
    double *d = nullptr;
    if (cudaMalloc((void **)&d, (size_t)blocks * threads * 8) != cudaSuccess) return BLTCU_STATUS_ERROR;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    int iters = 2000;
    double best = 0.0;
    for (int r = 0; r < reps + 1; ++r) {
        cudaEventRecord(e0);
        bltcu_fp64_peak_kernel<<<blocks, threads>>>(d, iters, 0.999999, 1e-9);
        cudaEventRecord(e1);
        if (cudaEventSynchronize(e1) != cudaSuccess) { cudaFree(d); return BLTCU_STATUS_ERROR; }
        float ms = 0.f;
        cudaEventElapsedTime(&ms, e0, e1);
        const double flops = 2.0 * 8 * 16 * (double)iters * blocks * threads;
        const double tf = flops / (ms * 1e-3) / 1e12;
        if (r == 0) { if (ms < 20.f) iters = (int)(iters * 25.0 / (ms > 0.01f ? ms : 0.01f)); continue; }  // warm-up sizes the run
        if (tf > best) best = tf;
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    cudaFree(d);
    *tflops = best;
    return BLTCU_STATUS_OK;
}

// The same for the INT32 multiply-add pipe (north_star quotes the search against "FP64/INT"): independent IMAD streams.
__global__ void __launch_bounds__(256) bltcu_int32_peak_kernel(int *out, int iters, int a, int b) {
    int x0 = threadIdx.x, x1 = x0 + 1, x2 = x0 + 2, x3 = x0 + 3, x4 = x0 + 4, x5 = x0 + 5, x6 = x0 + 6, x7 = x0 + 7;
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int r = 0; r < 16; ++r) {
            x0 = x0 * a + b; x1 = x1 * a + b; x2 = x2 * a + b; x3 = x3 * a + b;
            x4 = x4 * a + b; x5 = x5 * a + b; x6 = x6 * a + b; x7 = x7 * a + b;
        }
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = ((x0 + x1) + (x2 + x3)) + ((x4 + x5) + (x6 + x7));
}

/* Measured INT32 multiply-add peak in Tera-operations/s (2 operations per IMAD) on `device`; best of `reps` launches. */
extern "C" int bltcu_int32_peak(int device, int reps, double *tiops) {
    if (!tiops) return BLTCU_STATUS_INPUT_ERROR;
    if (cudaSetDevice(device) != cudaSuccess) return BLTCU_STATUS_ERROR;
    int sms = 0;
    if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device) != cudaSuccess) return BLTCU_STATUS_ERROR;
    const int blocks = sms * 8, threads = 256;
    int *d = nullptr;
    if (cudaMalloc((void **)&d, (size_t)blocks * threads * 4) != cudaSuccess) return BLTCU_STATUS_ERROR;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    int iters = 2000;
    double best = 0.0;
    for (int r = 0; r < reps + 1; ++r) {
        cudaEventRecord(e0);
        bltcu_int32_peak_kernel<<<blocks, threads>>>(d, iters, 1664525 + r, 1013904223);
        cudaEventRecord(e1);
        if (cudaEventSynchronize(e1) != cudaSuccess) { cudaFree(d); return BLTCU_STATUS_ERROR; }
        float ms = 0.f;
        cudaEventElapsedTime(&ms, e0, e1);
        const double ops = 2.0 * 8 * 16 * (double)iters * blocks * threads;
        const double t = ops / (ms * 1e-3) / 1e12;
        if (r == 0) { if (ms < 20.f) iters = (int)(iters * 25.0 / (ms > 0.01f ? ms : 0.01f)); continue; }
        if (t > best) best = t;
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    cudaFree(d);
    *tiops = best;
    return BLTCU_STATUS_OK;
}
