// k_noalloc.cu -- ugd_integral_noalloc: the integral pass (S and SQ planes of one 8-bit image) for callers that
// cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).
//
// It has no counterpart in the reference.  It is a variant of
// the engine's three-phase integral (k_image.cu: strip column sums -> scan of the strip sums -> S / SQ planes), with
// the strip sums -- the scratch of that pass -- in a file-scope __device__ array instead of detector-owned memory.
//
//   * g_na_scratch is 4 MiB; every GPU that loads libugdet.so holds it until the process exits;
//   * limit: w <= UGD_NOALLOC_MAX_W (4096) and h <= UGD_NOALLOC_MAX_H (4096); larger images are rejected;
//   * the entry point calls no cudaMalloc / cudaFree / allocating library routine: three kernel launches on the
//     caller's stream, nothing else.  Its one-time set-up per device is cudaGetSymbolAddress (no allocation);
//   * every call shares the array and the array keeps the last call's contents.  Nothing in it has to start at zero:
//     phase 1 overwrites every entry that phases 2 and 3 of the same call read (2 planes x nstrips x w), so no clear
//     is needed;  CALLS THAT OVERLAP ON THE GPU (two streams, two host threads, two processes' worth of graphs
//     replayed concurrently on one device) RACE on it -- serialise them on one stream or with events.
//
// Arithmetic: uint32 wrap-around, as the engine's own integral pass and cv::integral's CV_32S planes do
// (SURVEY A.4): sums are exact modulo 2^32 whatever the summation order.
#include <cstdint>
#include <cstdio>
#include <mutex>

#include <cuda_runtime.h>

#include "../../include/ugdet.h"

namespace {

constexpr int kStripRows = 32;
constexpr int kMaxStrips = UGD_NOALLOC_MAX_H / kStripRows;
constexpr int kThreads = 256;
constexpr int kColsPerThread = UGD_NOALLOC_MAX_W / kThreads;   // 16

// [plane: S, SQ][strip][column], pitch = the call's w
static __device__ __align__(256) uint32_t g_na_scratch[2 * kMaxStrips * UGD_NOALLOC_MAX_W];

// phase 1: column sums of v and v*v over each strip of 32 rows
__global__ void __launch_bounds__(kThreads) k_na_strip_sums(const uint8_t* __restrict__ img, int w, int h, size_t stride,
                                                            int nstrips, uint32_t* __restrict__ scratch) {
    const int x = blockIdx.x * kThreads + threadIdx.x, s = blockIdx.y;
    if (x >= w) return;
    const int y0 = s * kStripRows, y1 = min(h, y0 + kStripRows);
    uint32_t a = 0, q = 0;
    for (int y = y0; y < y1; ++y) {
        const uint32_t v = img[(size_t)y * stride + x];
        a += v;
        q += v * v;
    }
    scratch[(size_t)s * w + x] = a;
    scratch[(size_t)(nstrips + s) * w + x] = q;
}

// phase 2: exclusive scan of the strip sums down each column, in place
__global__ void __launch_bounds__(kThreads) k_na_strip_scan(int w, int nstrips, uint32_t* __restrict__ scratch) {
    const int x = blockIdx.x * kThreads + threadIdx.x;
    if (x >= w) return;
    uint32_t* p = scratch + (size_t)blockIdx.y * nstrips * w + x;
    uint32_t run = 0;
    for (int s = 0; s < nstrips; ++s) {
        const uint32_t v = p[(size_t)s * w];
        p[(size_t)s * w] = run;
        run += v;
    }
}

// phase 3: one block per strip.  Thread t owns columns [t * E, t * E + E): it keeps the running column sums (carry of
// the strips above + the strip's rows so far) in registers, and every row is one block-wide inclusive scan of them.
__global__ void __launch_bounds__(kThreads) k_na_planes(const uint8_t* __restrict__ img, int w, int h, size_t stride,
                                                        int nstrips, const uint32_t* __restrict__ scratch,
                                                        uint32_t* __restrict__ S, uint32_t* __restrict__ SQ) {
    __shared__ uint32_t s_warp[2][kThreads / 32];
    const int s = blockIdx.x, t = threadIdx.x, lane = t & 31, wid = t >> 5;
    const int E = (w + kThreads - 1) / kThreads;
    const int x0 = t * E;
    const size_t P = (size_t)w + 1;
    uint32_t ca[kColsPerThread], cq[kColsPerThread];
#pragma unroll
    for (int j = 0; j < kColsPerThread; ++j) {
        const bool in = j < E && x0 + j < w;
        ca[j] = in ? scratch[(size_t)s * w + x0 + j] : 0u;
        cq[j] = in ? scratch[(size_t)(nstrips + s) * w + x0 + j] : 0u;
    }
    if (s == 0)
        for (int x = t; x <= w; x += kThreads) { S[x] = 0u; SQ[x] = 0u; }
    const int y0 = s * kStripRows, y1 = min(h, y0 + kStripRows);
    for (int y = y0; y < y1; ++y) {
        const uint8_t* row = img + (size_t)y * stride;
        uint32_t ta = 0, tq = 0;
#pragma unroll
        for (int j = 0; j < kColsPerThread; ++j)
            if (j < E && x0 + j < w) {
                const uint32_t v = row[x0 + j];
                ca[j] += v;
                cq[j] += v * v;
                ta += ca[j];
                tq += cq[j];
            }
        // exclusive prefix of the per-thread totals over the block
        uint32_t ia = ta, iq = tq;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const uint32_t na = __shfl_up_sync(0xffffffffu, ia, d), nq = __shfl_up_sync(0xffffffffu, iq, d);
            if (lane >= d) { ia += na; iq += nq; }
        }
        if (lane == 31) { s_warp[0][wid] = ia; s_warp[1][wid] = iq; }
        __syncthreads();
        uint32_t ba = ia - ta, bq = iq - tq;
        for (int k = 0; k < wid; ++k) { ba += s_warp[0][k]; bq += s_warp[1][k]; }
        __syncthreads();
        uint32_t* so = S + (size_t)(y + 1) * P;
        uint32_t* qo = SQ + (size_t)(y + 1) * P;
        if (t == 0) { so[0] = 0u; qo[0] = 0u; }
#pragma unroll
        for (int j = 0; j < kColsPerThread; ++j)
            if (j < E && x0 + j < w) {
                ba += ca[j];
                bq += cq[j];
                so[x0 + j + 1] = ba;
                qo[x0 + j + 1] = bq;
            }
    }
}

// Device address of g_na_scratch, looked up once per device with that device current.  In host code the name
// g_na_scratch is only a stand-in for the symbol: a kernel must be given the address cudaGetSymbolAddress returns.
std::mutex g_na_mu;
uint32_t* g_na_ptr[64];

}  // namespace

extern "C" int ugd_integral_noalloc(int cuda_device, const uint8_t* d_img, int w, int h, size_t stride_bytes,
                                    int32_t* d_S, int32_t* d_SQ, void* cuda_stream) {
    if (!d_img || !d_S || !d_SQ || w < 1 || h < 1 || stride_bytes < (size_t)w || cuda_device < 0 || cuda_device >= 64)
        return UGD_ERR_ARG;
    if (w > UGD_NOALLOC_MAX_W || h > UGD_NOALLOC_MAX_H) {
        fprintf(stderr, "ugdet: ugd_integral_noalloc: image %d x %d exceeds the static scratch limit of %d x %d "
                        "(UGD_NOALLOC_MAX_W x UGD_NOALLOC_MAX_H); use a detector for larger frames\n",
                w, h, UGD_NOALLOC_MAX_W, UGD_NOALLOC_MAX_H);
        return UGD_ERR_CAPACITY;
    }
    cudaError_t e = cudaSetDevice(cuda_device);
    if (e != cudaSuccess) {
        fprintf(stderr, "ugdet: ugd_integral_noalloc: cudaSetDevice(%d): %s\n", cuda_device, cudaGetErrorString(e));
        return UGD_ERR_CUDA;
    }
    uint32_t* scratch;
    {
        std::lock_guard<std::mutex> lk(g_na_mu);
        if (!g_na_ptr[cuda_device]) {
            void* p = nullptr;
            e = cudaGetSymbolAddress(&p, g_na_scratch);
            if (e != cudaSuccess || !p) {
                fprintf(stderr, "ugdet: ugd_integral_noalloc: cudaGetSymbolAddress on device %d: %s\n", cuda_device,
                        cudaGetErrorString(e));
                return UGD_ERR_CUDA;
            }
            g_na_ptr[cuda_device] = static_cast<uint32_t*>(p);
        }
        scratch = g_na_ptr[cuda_device];
    }
    cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
    const int nstrips = (h + kStripRows - 1) / kStripRows, xb = (w + kThreads - 1) / kThreads;
    k_na_strip_sums<<<dim3(xb, nstrips), kThreads, 0, st>>>(d_img, w, h, stride_bytes, nstrips, scratch);
    k_na_strip_scan<<<dim3(xb, 2), kThreads, 0, st>>>(w, nstrips, scratch);
    k_na_planes<<<nstrips, kThreads, 0, st>>>(d_img, w, h, stride_bytes, nstrips, scratch,
                                              reinterpret_cast<uint32_t*>(d_S), reinterpret_cast<uint32_t*>(d_SQ));
    e = cudaGetLastError();
    if (e != cudaSuccess) {
        fprintf(stderr, "ugdet: ugd_integral_noalloc: launch failed: %s\n", cudaGetErrorString(e));
        return UGD_ERR_CUDA;
    }
    return UGD_OK;
}
