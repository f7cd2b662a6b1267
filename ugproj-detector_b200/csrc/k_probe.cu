// Measured denominator of k_scan's roofline: what a conflict-free stream of 32-bit shared-memory loads sustains on
// this part.  k_scan is gather work out of a shared-memory tile, so its roof is the LSU/shared pipe, nominally
// 148 SMs x 128 B/clk; this probe times the real thing (one kernel, every SM, full occupancy, ld.shared.b32 with the
// 32 lanes of a warp on 32 distinct banks, 16 independent loads in flight per thread) with CUDA events.
#include <cuda_runtime.h>

#include "../../include/ugdet.h"

namespace {

constexpr int kProbeThreads = 256, kProbeUnroll = 16;

__global__ void __launch_bounds__(kProbeThreads) k_lds_probe(int iters, unsigned* sink) {
    __shared__ unsigned tile[4096];                          // 16 KB
    for (int i = threadIdx.x; i < 4096; i += kProbeThreads) tile[i] = i * 2654435761u;
    __syncthreads();
    // lane l of every warp reads bank l (word index = lane + 32 * row); the 16 loads of an iteration use immediate
    // offsets from one base register, and one LOP3 folds two of them, so the loop issues ~1.6 instructions per LDS:
    // the shared-memory pipe, not the issue slots, is what it measures
    const unsigned base = (unsigned)__cvta_generic_to_shared(tile) + 4u * (threadIdx.x & 31) + 128u * (threadIdx.x >> 5);
    unsigned acc[kProbeUnroll / 2];
#pragma unroll
    for (int u = 0; u < kProbeUnroll / 2; ++u) acc[u] = 0;
#define UGD_LDS2(u)                                                                                        \
    {                                                                                                      \
        unsigned v0, v1;                                                                                   \
        asm volatile("ld.volatile.shared.b32 %0, [%1+" #u "*2048];" : "=r"(v0) : "r"(base) : "memory");                       \
        asm volatile("ld.volatile.shared.b32 %0, [%1+" #u "*2048+1024];" : "=r"(v1) : "r"(base) : "memory");                  \
        acc[u] ^= v0 ^ v1;                                                                                 \
    }
#pragma unroll 1
    for (int it = 0; it < iters; ++it) {
        UGD_LDS2(0) UGD_LDS2(1) UGD_LDS2(2) UGD_LDS2(3) UGD_LDS2(4) UGD_LDS2(5) UGD_LDS2(6) UGD_LDS2(7)
    }
#undef UGD_LDS2
    unsigned x = 0;
#pragma unroll
    for (int u = 0; u < kProbeUnroll / 2; ++u) x ^= acc[u];
    if (x == 0x12345u) sink[0] = x;                          // keeps the loads alive without a store per thread
}

}  // namespace

extern "C" int ugd_measure_lds_peak(int cuda_device, double* gbytes_per_s, double* bytes_per_clk_per_sm) {
    if (!gbytes_per_s) return UGD_ERR_ARG;
    if (cudaSetDevice(cuda_device) != cudaSuccess) return UGD_ERR_CUDA;
    int sms = 0, khz = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, cuda_device);
    cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, cuda_device);
    unsigned* sink = nullptr;
    if (cudaMalloc(&sink, 4) != cudaSuccess) return UGD_ERR_CUDA;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int blocks = sms * 8, iters = 2048;                // 8 x 256 threads = 64 warps per SM: full occupancy
    double best = 0.0;
    for (int rep = 0; rep < 6; ++rep) {                      // first repetitions warm the clocks up
        cudaEventRecord(e0);
        k_lds_probe<<<blocks, kProbeThreads>>>(iters, sink);
        cudaEventRecord(e1);
        if (cudaEventSynchronize(e1) != cudaSuccess) { cudaFree(sink); return UGD_ERR_CUDA; }
        float ms = 0;
        cudaEventElapsedTime(&ms, e0, e1);
        const double bytes = (double)blocks * kProbeThreads * (double)iters * kProbeUnroll * 4.0;
        const double gbs = bytes / (ms * 1e-3) / 1e9;
        if (rep >= 2 && gbs > best) best = gbs;
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    cudaFree(sink);
    *gbytes_per_s = best;
    if (bytes_per_clk_per_sm) *bytes_per_clk_per_sm = khz > 0 ? best * 1e9 / ((double)sms * khz * 1e3) : 0.0;
    return UGD_OK;
}
