// Image-side kernels of the detection path (sm_100a): colour conversion + skin mask + histogram,
// equalizeHist, pyramid resize, integral / squared-integral images.  All integer, bit-exact with
// the oracle.  These passes are HBM-bound: every kernel
// moves 16-byte vectors per thread, consecutive threads touch consecutive addresses.
#include <mutex>

#include "kernels.cuh"

namespace ugd {

__device__ __forceinline__ uint32_t sat8(int v) { return (uint32_t)min(max(v, 0), 255); }

// ------------------------------------------------------------------------------------------------
// K1  BGR -> gray (a4) + YCrCb skin mask (a2,a3) + 256-bin histogram of gray (first half of a5).
// 4 pixels per thread per iteration: 12 B in, 4 B + 4 B out.
// AREA2 = true fuses the 2x2 INTER_AREA decimation of SURVEY 8(f) N1 in front (reference face_tracker.cc:112-114 with
// scan.frame_size = half the frame, as conf/apink.mrchu.720p.cfg has it): the thread reads the 2 x 8 source pixels under
// its four scan-frame pixels (three 8-byte loads per row, a warp reads 768 contiguous bytes) and the decimated BGR
// frame is never written.
// Histogram: one private 256-bin copy per warp, shared-memory atomics.  (Tried: a private histogram of 8-bit counters per
// THREAD -- 64 KB, bank-conflict-free plain load-add-store, summed with DP4A at the end -- to get rid of the atomics:
// 0.078 ms against 0.068 ms for this form on 64 fused 720p frames; zeroing and summing 64 KB per 8192 pixels costs more
// than the atomics do, and with 32768 pixels per CTA the grid gets too coarse.)
// ------------------------------------------------------------------------------------------------
constexpr int kPreIters = 8;   // quads per thread -> 8192 pixels per CTA, 256 global atomics per CTA

template <bool AREA2>
__global__ void __launch_bounds__(256) k_preprocess(DevBufs b) {
    __shared__ uint32_t sh[8][256];                         // per-warp private histograms
    const int frame = blockIdx.y, tid = threadIdx.x, warp = tid >> 5;
    for (int i = tid; i < 8 * 256; i += 256) (&sh[0][0])[i] = 0;
    __syncthreads();
    const int n = b.w * b.h;
    const uint8_t* src = AREA2 ? b.area2_src + (size_t)frame * b.area2_frame_stride : b.bgr + (size_t)frame * b.bgr_stride;
    uint8_t* gray = b.gray_raw + (size_t)frame * b.px_stride;
    uint8_t* mask = b.mask + (size_t)frame * b.px_stride;
    const int lo0 = b.skin_lo[0], lo1 = b.skin_lo[1], lo2 = b.skin_lo[2];
    const int hi0 = b.skin_hi[0], hi1 = b.skin_hi[1], hi2 = b.skin_hi[2];
    for (int it = 0; it < kPreIters; ++it) {
        int q = (blockIdx.x * kPreIters + it) * 256 + tid;  // quad index
        int p0 = q * 4;
        if (p0 >= n) break;
        uint8_t px[12];
        if (AREA2) {
            // b.w is a multiple of 4 here (launch_preprocess checks): a quad never straddles scan-frame rows
            const int y = p0 / b.w, x = p0 - y * b.w;
            const uint2* r0 = reinterpret_cast<const uint2*>(src + (size_t)(2 * y) * b.area2_row_stride + (size_t)x * 6);
            const uint2* r1 = reinterpret_cast<const uint2*>(src + (size_t)(2 * y + 1) * b.area2_row_stride + (size_t)x * 6);
            uint8_t t0[24], t1[24];
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                reinterpret_cast<uint2*>(t0)[k] = __ldg(r0 + k);
                reinterpret_cast<uint2*>(t1)[k] = __ldg(r1 + k);
            }
#pragma unroll
            for (int k = 0; k < 4; ++k)
#pragma unroll
                for (int c = 0; c < 3; ++c)
                    px[3 * k + c] = (uint8_t)(((int)t0[6 * k + c] + t0[6 * k + 3 + c] + t1[6 * k + c] + t1[6 * k + 3 + c] + 2) >> 2);
        } else if (p0 + 4 <= n) {
            const uint32_t* s32 = reinterpret_cast<const uint32_t*>(src) + (size_t)q * 3;
            uint32_t a = s32[0], c = s32[1], d = s32[2];
            *reinterpret_cast<uint32_t*>(px) = a;
            *reinterpret_cast<uint32_t*>(px + 4) = c;
            *reinterpret_cast<uint32_t*>(px + 8) = d;
        } else {
            for (int k = 0; k < 12; ++k) px[k] = (p0 * 3 + k < n * 3) ? src[(size_t)p0 * 3 + k] : 0;
        }
        uint32_t g4 = 0, m4 = 0;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            int B = px[3 * k], G = px[3 * k + 1], R = px[3 * k + 2];
            int Y = (1868 * B + 9617 * G + 4899 * R + 8192) >> 14;
            int g = b.gray14 ? Y : (3735 * B + 19235 * G + 9798 * R + 16384) >> 15;      // (OpenCV 2.4 used the 14-bit coefficients for BGR2GRAY too)
            int Cr = (int)sat8(((R - Y) * 11682 + 128 * 16384 + 8192) >> 14);
            int Cb = (int)sat8(((B - Y) * 9241 + 128 * 16384 + 8192) >> 14);
            Y = (int)sat8(Y);
            bool ok = Y >= lo0 && Y <= hi0 && Cr >= lo1 && Cr <= hi1 && Cb >= lo2 && Cb <= hi2;
            g4 |= (uint32_t)g << (8 * k);
            m4 |= (ok ? 255u : 0u) << (8 * k);
            if (p0 + k < n) atomicAdd(&sh[warp][g], 1u);
        }
        if (p0 + 4 <= n) {
            *reinterpret_cast<uint32_t*>(gray + p0) = g4;
            *reinterpret_cast<uint32_t*>(mask + p0) = m4;
        } else {
            for (int k = 0; p0 + k < n; ++k) { gray[p0 + k] = (uint8_t)(g4 >> (8 * k)); mask[p0 + k] = (uint8_t)(m4 >> (8 * k)); }
        }
    }
    __syncthreads();
    uint32_t t = 0;
#pragma unroll
    for (int wv = 0; wv < 8; ++wv) t += sh[wv][tid];
    if (t) atomicAdd(&b.hist[(size_t)frame * 256 + tid], t);
}

// The fused decimation needs whole quads per scan-frame row and 8-byte aligned source rows.
bool preprocess_can_fuse_area2(const uint8_t* src, size_t row_stride, size_t frame_stride, int scan_w) {
    return scan_w % 4 == 0 && reinterpret_cast<uintptr_t>(src) % 8 == 0 && row_stride % 8 == 0 && frame_stride % 8 == 0;
}

void launch_preprocess(const DevBufs& b, cudaStream_t st) {
    int n = b.w * b.h;
    int per_cta = 256 * 4 * kPreIters;
    dim3 grid((n + per_cta - 1) / per_cta, b.nframes);
    if (b.area2_src) k_preprocess<true><<<grid, 256, 0, st>>>(b);
    else k_preprocess<false><<<grid, 256, 0, st>>>(b);
}

// ------------------------------------------------------------------------------------------------
// K2  equalizeHist LUT (a5) + apply + AND with the skin mask (a6).  16 pixels per thread.
// Every CTA rebuilds the 256-entry LUT from the frame histogram (256 threads, one bin each).
// ------------------------------------------------------------------------------------------------
constexpr int kEqIters = 4;

__global__ void __launch_bounds__(256) k_equalize(DevBufs b) {
    __shared__ uint32_t cum[256];
    __shared__ uint8_t lut[256];
    __shared__ int s_i0;
    const int frame = blockIdx.y, tid = threadIdx.x;
    const int n = b.w * b.h;
    uint32_t hv = b.hist[(size_t)frame * 256 + tid];
    if (tid == 0) s_i0 = 256;
    __syncthreads();
    if (hv) atomicMin(&s_i0, tid);
    cum[tid] = hv;
    __syncthreads();
    const int i0 = s_i0;
    // inclusive scan over bins (Hillis-Steele, 8 steps)
    for (int d = 1; d < 256; d <<= 1) {
        uint32_t v = tid >= d ? cum[tid - d] : 0;
        __syncthreads();
        cum[tid] += v;
        __syncthreads();
    }
    {
        uint32_t h0 = i0 < 256 ? cum[i0] - (i0 ? cum[i0 - 1] : 0) : 0;
        if (i0 == 256 || h0 == (uint32_t)n) lut[tid] = (uint8_t)tid;          // constant image: unchanged
        else if (tid <= i0) lut[tid] = 0;
        else {
            float scale = 255.f / (float)(n - (int)h0);                       // IEEE division
            uint32_t s = cum[tid] - cum[i0];
            lut[tid] = (uint8_t)sat8(__float2int_rn(__fmul_rn((float)s, scale)));
        }
    }
    __syncthreads();
    const uint8_t* src = b.gray_raw + (size_t)frame * b.px_stride;
    const uint8_t* mk = b.mask + (size_t)frame * b.px_stride;
    uint8_t* dst = b.gray_eq + (size_t)frame * b.px_stride;
    for (int it = 0; it < kEqIters; ++it) {
        int p0 = ((blockIdx.x * kEqIters + it) * 256 + tid) * 16;
        if (p0 >= n) break;
        // planes are padded to a multiple of 256 bytes, so the vector access is always in bounds
        uint4 g = *reinterpret_cast<const uint4*>(src + p0);
        uint4 m = *reinterpret_cast<const uint4*>(mk + p0);
        uint32_t gi[4] = {g.x, g.y, g.z, g.w}, mi[4] = {m.x, m.y, m.z, m.w}, o[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            uint32_t v = gi[k];
            uint32_t r = (uint32_t)lut[v & 255] | (uint32_t)lut[(v >> 8) & 255] << 8 |
                         (uint32_t)lut[(v >> 16) & 255] << 16 | (uint32_t)lut[v >> 24] << 24;
            o[k] = b.use_skin ? (r & mi[k]) : r;
        }
        *reinterpret_cast<uint4*>(dst + p0) = make_uint4(o[0], o[1], o[2], o[3]);
    }
}

void launch_equalize(const DevBufs& b, cudaStream_t st) {
    int n = b.w * b.h;
    int per_cta = 256 * 16 * kEqIters;
    dim3 grid((n + per_cta - 1) / per_cta, b.nframes);
    k_equalize<<<grid, 256, 0, st>>>(b);
}

// ------------------------------------------------------------------------------------------------
// K3  pyramid resize (a7; INTER_LINEAR_EXACT, A.2) fused with the first phase of the integral (a8): per-strip column
// sums.  One thread owns FOUR output columns of one level and walks the 32 rows of a strip: the x side of the
// bilinear blend (source column, 8.8 coefficients, source pointer) is set up once and reused for every row, the pixels
// of a row leave as one 32-bit store (a warp writes 128 consecutive bytes), and the running column sums of pixel and
// pixel^2 are written once per strip -- the pyramid is never re-read for them.  ~19 instructions per output pixel
// (the separate kernels took 56 + 4); all levels of all frames in one launch.
// The tables never point at the last source column / row: there (i0 = n-2, c1 = 256) stands for (n-1, 0), so the
// "+1" neighbour is always inside the frame and needs no clamp.
// ------------------------------------------------------------------------------------------------
constexpr int kResizeThreads = 64;     // x 4 columns = 256 columns per block

__global__ void __launch_bounds__(kResizeThreads) k_resize_colsum(DevBufs b, const WorkItem* __restrict__ wk) {
    const WorkItem it = wk[blockIdx.x];
    const LevelInfo& L = b.levels[it.level];
    const int strip = it.first & 0xffff, chunk = it.first >> 16;
    const int frame = blockIdx.y;
    const int sw = L.sw, pitch = L.img_pitch, spitch = b.src_pitch;
    // A warp owns 128 consecutive output columns; lane l takes columns l, l + 32, l + 64, l + 96 of them, so that the 32
    // lanes of every load instruction read NEIGHBOURING source pixels (32 x scale bytes: one or two L1 wavefronts instead
    // of the four to six that columns 4 apart cost -- the kernel is bound by L1 wavefronts, ncu 87 % of peak before)
    const int c0 = chunk * (4 * kResizeThreads) + (threadIdx.x >> 5) * 128 + (threadIdx.x & 31);
    if (c0 - (int)(threadIdx.x & 31) >= pitch) return;      // the whole warp lies in the row padding
    const uint32_t* xtab = b.xtab + L.xtab_off;
    const uint8_t* src = b.src_gray + (size_t)frame * b.src_stride;
    const uint8_t* col[4];
    int cx1[4];
    bool live[4], inpitch[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const int c = c0 + 32 * k;
        inpitch[k] = c < pitch;
        const uint32_t xe = inpitch[k] ? __ldg(xtab + c) : 0u;
        col[k] = src + (xe & 0xffffu);
        cx1[k] = (int)(xe >> 16);
        live[k] = c < sw;                                   // columns of the row padding stay 0
    }
    const uint32_t* ytab = b.ytab + L.ytab_off;
    const int v0 = strip * kStripRows, v1 = min(v0 + kStripRows, L.sh);
    uint8_t* dst = b.pyr + (size_t)frame * b.pyr_stride + L.img_off + (size_t)v0 * pitch + c0;
    uint32_t s[4] = {0, 0, 0, 0}, q[4] = {0, 0, 0, 0};
#pragma unroll 2
    for (int v = v0; v < v1; ++v, dst += pitch) {
        const uint32_t ye = __ldg(ytab + v);
        const int cy1 = (int)(ye >> 16), cy0 = 256 - cy1;
        const size_t roff = (size_t)(ye & 0xffffu) * spitch;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const uint8_t* r0 = col[k] + roff;
            const uint8_t* r1 = r0 + spitch;
            const int cx0 = 256 - cx1[k];
            const int t0 = (int)__ldg(r0) * cx0 + (int)__ldg(r0 + 1) * cx1[k];
            const int t1 = (int)__ldg(r1) * cx0 + (int)__ldg(r1 + 1) * cx1[k];
            const uint32_t pv = live[k] ? (uint32_t)(t0 * cy0 + t1 * cy1 + 32768) >> 16 : 0u;
            if (inpitch[k]) dst[32 * k] = (uint8_t)pv;
            s[k] += pv; q[k] += pv * pv;
        }
    }
    const size_t o = (size_t)frame * b.carry_stride + L.carry_off + (size_t)strip * pitch + c0;
#pragma unroll
    for (int k = 0; k < 4; ++k)
        if (inpitch[k]) { b.carryS[o + 32 * k] = s[k]; b.carryQ[o + 32 * k] = q[k]; }
}

void launch_resize_colsum(const DevBufs& b, const WorkItem* wk, int nblocks, cudaStream_t st) {
    if (nblocks) k_resize_colsum<<<dim3(nblocks, b.nframes), kResizeThreads, 0, st>>>(b, wk);
}

// ------------------------------------------------------------------------------------------------
// K4  integral images S and SQ (a8).  The strip column sums come from k_resize_colsum; a small pass turns them into the
// starting column sums of every 32-row strip:
//   carry  : exclusive scan of the strip sums down the strips (tiny)
// and k_integral_bands (below) then writes both planes in one sweep, each output element once.
// Arithmetic is uint32 modular, which reproduces cv2's wrapping int32 SQ plane bit for bit.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_carry(DevBufs b, const WorkItem* __restrict__ wk) {
    const WorkItem it = wk[blockIdx.x];
    const LevelInfo& L = b.levels[it.level];
    const int quads = L.img_pitch >> 2;
    const int qx = it.first + threadIdx.x;
    if (qx >= quads) return;
    const int frame = blockIdx.y;
    size_t o = (size_t)frame * b.carry_stride + L.carry_off + qx * 4;
    uint4 as = make_uint4(0, 0, 0, 0), aq = as;
    for (int t0 = 0; t0 < L.nstrips; t0 += 8) {                      // 8 strips of loads in flight, then the serial prefix
        uint4 vs[8], vq[8];
#pragma unroll
        for (int u = 0; u < 8; ++u)
            if (t0 + u < L.nstrips) {
                vs[u] = *reinterpret_cast<const uint4*>(b.carryS + o + (size_t)u * L.img_pitch);
                vq[u] = *reinterpret_cast<const uint4*>(b.carryQ + o + (size_t)u * L.img_pitch);
            }
#pragma unroll
        for (int u = 0; u < 8; ++u)
            if (t0 + u < L.nstrips) {
                *reinterpret_cast<uint4*>(b.carryS + o + (size_t)u * L.img_pitch) = as;
                *reinterpret_cast<uint4*>(b.carryQ + o + (size_t)u * L.img_pitch) = aq;
                as.x += vs[u].x; as.y += vs[u].y; as.z += vs[u].z; as.w += vs[u].w;
                aq.x += vq[u].x; aq.y += vq[u].y; aq.z += vq[u].z; aq.w += vq[u].w;
            }
        o += (size_t)8 * L.img_pitch;
    }
}

// ------------------------------------------------------------------------------------------------
// K4'  the integral main pass: one warp owns a band of 512 output columns x a segment of 64 pixel rows of
// one level of one frame, seeded from the strip carries, and walks it top to bottom.  Lane l holds 16 consecutive columns:
// per row it loads 16 pixels (one 16-byte load), forms their exclusive prefix (and that of their squares),
// a warp shuffle scan supplies the offset of the lanes to its left, and the running column sums live in
// registers from the first row to the last, so nothing is ever re-read and every output element is written
// once with 16-byte stores.  Bands to the right of the first one add the row sums of the bands on their left
// (dp4a + redux.sync on pixels that sit in L2).  All arithmetic is uint32 modular, like cv2's int32 SQ plane.
// ------------------------------------------------------------------------------------------------
constexpr int kBandCols = 512;
constexpr int kSegRows = kIntSegRows;

// 256-bit store (sm_100: STG.E.256): a lane's 8 consecutive ints go out as one fully written 32-byte sector
__device__ __forceinline__ void st256(uint32_t* p, const uint32_t* v) {
    asm volatile("st.global.v8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(p), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]),
                 "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]) : "memory");
}

// One lane's 16 consecutive columns of an integral row.  Natural layout: two 32-byte stores at column X0.  De-interleaved
// layout (grid-step-2 levels, DevBufs::deint2): a row is [P/2 even columns][P/2 odd columns] -- what k_scan's two
// sub-plane tiles and its TMA boxes want -- so the 8 even and the 8 odd columns go out as one 32-byte store each.
__device__ __forceinline__ void store_row16(uint32_t* row, int X0, int P, bool deint, const uint32_t (&v)[16]) {
    if (!deint) { st256(row + X0, v); st256(row + X0 + 8, v + 8); return; }
    const uint32_t ev[8] = {v[0], v[2], v[4], v[6], v[8], v[10], v[12], v[14]};
    const uint32_t od[8] = {v[1], v[3], v[5], v[7], v[9], v[11], v[13], v[15]};
    st256(row + (X0 >> 1), ev);
    st256(row + (P >> 1) + (X0 >> 1), od);
}

__device__ __forceinline__ uint4 load_px16(const uint8_t* row, int x, int pitch) {
    return x < pitch ? *reinterpret_cast<const uint4*>(row + x) : make_uint4(0u, 0u, 0u, 0u);
}

__global__ void __launch_bounds__(32) k_integral_bands(DevBufs b, const WorkItem* __restrict__ wk) {
    // grid = (frames, bands): consecutive blocks are the SAME band of different frames, so the tall bands of
    // the large levels are dealt evenly over the SMs and start first
    const WorkItem it = wk[blockIdx.y];
    const LevelInfo& L = b.levels[it.level];
    const int frame = blockIdx.x, lane = threadIdx.x;
    const int pitch = L.img_pitch, P = L.P;
    const int band = it.first & 255, seg = it.first >> 8;                  // a band is cut into segments of kSegRows pixel rows
    const int X0 = band * kBandCols + lane * 16;                           // first output column of this lane
    const int Y0 = seg * kSegRows, Y1 = min(Y0 + kSegRows, L.sh);          // pixel rows [Y0, Y1) -> output rows Y0+1 .. Y1
    const uint8_t* img = b.pyr + (size_t)frame * b.pyr_stride + L.img_off + (size_t)Y0 * pitch;
    uint32_t* outS = reinterpret_cast<uint32_t*>(b.S + (size_t)frame * b.int_stride + L.int_off) + (size_t)Y0 * P;
    uint32_t* outQ = reinterpret_cast<uint32_t*>(b.SQ + (size_t)frame * b.int_stride + L.int_off) + (size_t)Y0 * P;
    const bool stores = X0 < P;                                            // P and X0 are multiples of 16: all 16 columns or none
    const bool deint = b.deint2 && L.step == 2;
    uint32_t accS[16], accQ[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) { accS[k] = 0; accQ[k] = 0; }
    if (seg == 0) {                                                        // row 0 of both planes is zero
        if (stores) { store_row16(outS, X0, P, deint, accS); store_row16(outQ, X0, P, deint, accQ); }
    } else {
        // output row Y0 = prefix along x of the column sums of all pixel rows above (k_colsum + k_carry, strip = 32 rows)
        const size_t co = (size_t)frame * b.carry_stride + L.carry_off + (size_t)(Y0 / kStripRows) * pitch;
        const uint32_t* cS = b.carryS + co;
        const uint32_t* cQ = b.carryQ + co;
        uint32_t leftS = 0, leftQ = 0;
        for (int lb = 0; lb < band; ++lb) {
            const int x = lb * kBandCols + lane * 16;
            uint32_t s = 0, q = 0;
            if (x < pitch) {
#pragma unroll
                for (int v = 0; v < 4; ++v) {
                    const uint4 a = *reinterpret_cast<const uint4*>(cS + x + 4 * v), c = *reinterpret_cast<const uint4*>(cQ + x + 4 * v);
                    s += a.x + a.y + a.z + a.w; q += c.x + c.y + c.z + c.w;
                }
            }
            leftS += __reduce_add_sync(0xffffffffu, s);
            leftQ += __reduce_add_sync(0xffffffffu, q);
        }
        uint32_t rs = 0, rq = 0;
        if (X0 < pitch) {
#pragma unroll
            for (int v = 0; v < 4; ++v) {
                const uint4 a = *reinterpret_cast<const uint4*>(cS + X0 + 4 * v), c = *reinterpret_cast<const uint4*>(cQ + X0 + 4 * v);
                accS[4 * v] = rs; rs += a.x; accS[4 * v + 1] = rs; rs += a.y; accS[4 * v + 2] = rs; rs += a.z; accS[4 * v + 3] = rs; rs += a.w;
                accQ[4 * v] = rq; rq += c.x; accQ[4 * v + 1] = rq; rq += c.y; accQ[4 * v + 2] = rq; rq += c.z; accQ[4 * v + 3] = rq; rq += c.w;
            }
        } else {
#pragma unroll
            for (int k = 0; k < 16; ++k) { accS[k] = 0; accQ[k] = 0; }
        }
        uint32_t is = rs, iq = rq;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const uint32_t a = __shfl_up_sync(0xffffffffu, is, d), e = __shfl_up_sync(0xffffffffu, iq, d);
            if (lane >= d) { is += a; iq += e; }
        }
        const uint32_t offS = is - rs + leftS, offQ = iq - rq + leftQ;
#pragma unroll
        for (int k = 0; k < 16; ++k) { accS[k] += offS; accQ[k] += offQ; }
    }
    const int sh = Y1 - Y0;                                                // rows of this segment (the loop below is segment-relative)
    constexpr int PF = 4;                                                  // rows in flight: loads run PF rows ahead of the stores
    uint4 nx[PF];
#pragma unroll
    for (int r = 0; r < PF; ++r) nx[r] = r < sh ? load_px16(img + (size_t)r * pitch, X0, pitch) : make_uint4(0u, 0u, 0u, 0u);
    for (int y0 = 0; y0 < sh; y0 += PF) {
        uint4 cur[PF];
        uint32_t leftS[PF], leftQ[PF];
#pragma unroll
        for (int r = 0; r < PF; ++r) {
            cur[r] = nx[r];
            const int yn = y0 + PF + r;
            if (yn < sh) nx[r] = load_px16(img + (size_t)yn * pitch, X0, pitch);
            leftS[r] = 0; leftQ[r] = 0;
        }
        // row sums of the bands to the left (pixels beyond sw are zero padding)
        for (int lb = 0; lb < band; ++lb) {
            uint4 q[PF];
#pragma unroll
            for (int r = 0; r < PF; ++r)
                q[r] = y0 + r < sh ? load_px16(img + (size_t)(y0 + r) * pitch, lb * kBandCols + lane * 16, pitch) : make_uint4(0u, 0u, 0u, 0u);
#pragma unroll
            for (int r = 0; r < PF; ++r) {
                uint32_t s = __dp4a(q[r].x, 0x01010101u, 0u), t = __dp4a(q[r].x, q[r].x, 0u);
                s = __dp4a(q[r].y, 0x01010101u, s); t = __dp4a(q[r].y, q[r].y, t);
                s = __dp4a(q[r].z, 0x01010101u, s); t = __dp4a(q[r].z, q[r].z, t);
                s = __dp4a(q[r].w, 0x01010101u, s); t = __dp4a(q[r].w, q[r].w, t);
                leftS[r] += __reduce_add_sync(0xffffffffu, s);
                leftQ[r] += __reduce_add_sync(0xffffffffu, t);
            }
        }
#pragma unroll
        for (int r = 0; r < PF; ++r) {
            const int y = y0 + r;
            if (y >= sh) break;
            const uint32_t w4[4] = {cur[r].x, cur[r].y, cur[r].z, cur[r].w};
            // exclusive prefix inside the lane, added straight into the running column sums
            uint32_t rs = 0, rq = 0;
#pragma unroll
            for (int k = 0; k < 16; ++k) {
                const uint32_t p = (w4[k >> 2] >> (8 * (k & 3))) & 255u;
                accS[k] += rs; accQ[k] += rq;
                rs += p; rq += p * p;
            }
            // exclusive scan of the lane totals across the warp
            uint32_t is = rs, iq = rq;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const uint32_t a = __shfl_up_sync(0xffffffffu, is, d), e = __shfl_up_sync(0xffffffffu, iq, d);
                if (lane >= d) { is += a; iq += e; }
            }
            const uint32_t offS = is - rs + leftS[r], offQ = iq - rq + leftQ[r];
#pragma unroll
            for (int k = 0; k < 16; ++k) { accS[k] += offS; accQ[k] += offQ; }
            if (stores) { store_row16(outS + (size_t)(y + 1) * P, X0, P, deint, accS); store_row16(outQ + (size_t)(y + 1) * P, X0, P, deint, accQ); }
        }
    }
}

// ------------------------------------------------------------------------------------------------
// K4t  tilted integral T (third output of cv::integral, int32, same (sh+1) x P layout as S):
//   T(X,Y) = sum of img(x,y) over y < Y, |x - X + 1| <= Y - y - 1   -- a 45-degree cone with apex (X-1, Y-1).
// With D1 / D2 the running sums along the two diagonals that end in a pixel,
//   cone(x,y) = cone(x,y-1) + D1(x,y) + D2(x,y) - img(x,y),   D1(x,y) = D1(x-1,y-1) + img,  D2(x,y) = D2(x+1,y-1) + img.
// One block walks one level of one frame top to bottom; a thread owns J columns (stride 256), keeps their
// cone sums in registers and hands the diagonal sums to its neighbours through double-buffered shared
// arrays (one barrier per row).  The next row's pixels are fetched before the barrier.  uint32 modular.
// Only launched for cascades with tilted features.
// ------------------------------------------------------------------------------------------------
constexpr int kTiltThreads = 256;

template <int J>
__device__ __forceinline__ void tilted_rows(const uint8_t* __restrict__ img, int pitch, int sw, int sh, uint32_t* __restrict__ T,
                                            int P, uint32_t* sm, int slots) {
    const int tid = threadIdx.x, W1 = sw + 1;
    uint32_t* d1 = sm;                    // [2][slots]: slot X + 1 holds the sum of apex column X - 1
    uint32_t* d2 = sm + 2 * slots;
    uint32_t cone[J], px[J];
#pragma unroll
    for (int j = 0; j < J; ++j) {
        const int X = tid + j * kTiltThreads;
        cone[j] = 0;
        px[j] = (X >= 1 && X < W1) ? img[X - 1] : 0u;
        if (X < W1) T[X] = 0;
    }
    for (int i = tid; i < 4 * slots; i += kTiltThreads) sm[i] = 0;
    __syncthreads();
    for (int y = 0; y < sh; ++y) {
        const int cur = (y & 1) * slots, nxt = slots - cur;
        uint32_t p[J];
#pragma unroll
        for (int j = 0; j < J; ++j) {
            const int X = tid + j * kTiltThreads;
            p[j] = px[j];
            if (y + 1 < sh) px[j] = (X >= 1 && X < W1) ? img[(size_t)(y + 1) * pitch + X - 1] : 0u;
        }
#pragma unroll
        for (int j = 0; j < J; ++j) {
            const int X = tid + j * kTiltThreads;
            if (X < W1) {
                const uint32_t n1 = d1[cur + X] + p[j];          // from the upper-left neighbour
                const uint32_t n2 = d2[cur + X + 2] + p[j];      // from the upper-right neighbour
                cone[j] += n1 + n2 - p[j];
                T[(size_t)(y + 1) * P + X] = cone[j];
                d1[nxt + X + 1] = n1;
                d2[nxt + X + 1] = n2;
            }
        }
        __syncthreads();
    }
}

__global__ void __launch_bounds__(kTiltThreads) k_tilted(DevBufs b) {
    extern __shared__ uint32_t s_tilt[];
    const LevelInfo& L = b.levels[blockIdx.x];
    const int frame = blockIdx.y;
    const uint8_t* img = b.pyr + (size_t)frame * b.pyr_stride + L.img_off;
    uint32_t* T = reinterpret_cast<uint32_t*>(b.T + (size_t)frame * b.int_stride + L.int_off);
    const int slots = L.sw + 3;
    const int need = (L.sw + 1 + kTiltThreads - 1) / kTiltThreads;
    if (need <= 1) tilted_rows<1>(img, L.img_pitch, L.sw, L.sh, T, L.P, s_tilt, slots);
    else if (need <= 2) tilted_rows<2>(img, L.img_pitch, L.sw, L.sh, T, L.P, s_tilt, slots);
    else if (need <= 4) tilted_rows<4>(img, L.img_pitch, L.sw, L.sh, T, L.P, s_tilt, slots);
    else if (need <= 8) tilted_rows<8>(img, L.img_pitch, L.sw, L.sh, T, L.P, s_tilt, slots);
    else tilted_rows<17>(img, L.img_pitch, L.sw, L.sh, T, L.P, s_tilt, slots);
}

static size_t tilted_smem(int max_w) { return (size_t)4 * (max_w + 3) * sizeof(uint32_t); }

// Called once per plan (per frame size) on the detector's device: levels wider than ~3000 pixels need more
// than the default 48 KB of dynamic shared memory.
cudaError_t init_tilted_attributes(int device, int max_w) {
    static std::mutex mu;
    static int granted[64] = {0};                       // per device, only ever raised (a smaller plan must not lower it)
    const int need = (int)tilted_smem(max_w);
    if (need <= 48 * 1024) return cudaSuccess;
    std::lock_guard<std::mutex> lk(mu);
    if (need <= granted[device & 63]) return cudaSuccess;
    const cudaError_t e = cudaFuncSetAttribute(k_tilted, cudaFuncAttributeMaxDynamicSharedMemorySize, need);
    if (e == cudaSuccess) granted[device & 63] = need;
    return e;
}

void launch_tilted(const DevBufs& b, int max_level_w, cudaStream_t st) {
    if (!b.T || !b.nlevels) return;
    k_tilted<<<dim3(b.nlevels, b.nframes), kTiltThreads, tilted_smem(max_level_w), st>>>(b);
}

void launch_integral_bands(const DevBufs& b, const WorkItem* wk, int nblocks, cudaStream_t st) {
    if (nblocks) k_integral_bands<<<dim3(b.nframes, nblocks), 32, 0, st>>>(b, wk);
}

void launch_carry(const DevBufs& b, const WorkItem* wk, int nblocks, cudaStream_t st) {
    if (nblocks) k_carry<<<dim3(nblocks, b.nframes), 256, 0, st>>>(b, wk);
}

}  // namespace ugd
