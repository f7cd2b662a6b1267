// groupRectangles + clip + skin post-filter on the device,
// one CTA per frame, so a batch never round-trips to the host between the cascade and the result.
// Candidates per frame are few (10^2..10^3); this kernel is latency-, not bandwidth-bound.
#include "kernels.cuh"

namespace ugd {

__device__ __forceinline__ bool similar_rects(const ugd_rect& a, const ugd_rect& b) {
    const double eps = 0.2;
    const double d = __dmul_rn(__dmul_rn(eps, (double)(min(a.width, b.width) + min(a.height, b.height))), 0.5);
    return (double)abs(a.x - b.x) <= d && (double)abs(a.y - b.y) <= d &&
           (double)abs(a.x + a.width - b.x - b.width) <= d && (double)abs(a.y + a.height - b.y - b.height) <= d;
}

__device__ __forceinline__ bool rect_less(const ugd_rect& a, const ugd_rect& b) {   // canonical output order
    if (a.y != b.y) return a.y < b.y;
    if (a.x != b.x) return a.x < b.x;
    if (a.height != b.height) return a.height < b.height;
    return a.width < b.width;
}

// Phase 1: label[i] = i.   Phase 2: every similar pair (i < j) is united; the pair space of a frame is cut into
// (i-block, j-chunk) tiles over many CTAs, so deep pyramids with thousands of candidates do not serialise on one SM.
constexpr int kLinkI = 128, kLinkJ = 8;

__global__ void __launch_bounds__(256) k_group_init(DevBufs b) {
    const int frame = blockIdx.y, i = blockIdx.x * 256 + threadIdx.x;
    if (i < min(b.rawcount[frame], b.rawcap)) b.grp_scratch[(size_t)frame * b.grp_stride + i] = i;
}

// find with path halving.  Only roots are ever CAS-linked, so shortening a non-root's pointer to its
// grandparent (which is in the same tree and smaller) can never undo a link.
__device__ __forceinline__ int uf_find(volatile int* parent, int x) {
    int p = parent[x];
    while (p != x) {
        const int g = parent[p];
        if (g != p) parent[x] = g;
        x = p; p = g;
    }
    return x;
}

__global__ void __launch_bounds__(kLinkI) k_group_link(DevBufs b) {
    const int frame = blockIdx.z;
    const int n = min(b.rawcount[frame], b.rawcap);
    const int i = blockIdx.x * kLinkI + threadIdx.x;
    if (blockIdx.x * kLinkI >= n || b.min_neighbors <= 0) return;
    const ugd_rect* raw = b.raw + (size_t)frame * b.rawcap;
    int* parent = b.grp_scratch + (size_t)frame * b.grp_stride;
    const int chunk = (n + kLinkJ - 1) / kLinkJ;
    const int j0 = max(blockIdx.y * chunk, blockIdx.x * kLinkI + 1), j1 = min((blockIdx.y + 1) * chunk, n);
    if (i >= n) return;
    const ugd_rect ri = raw[i];
    // integer pre-filter: SimilarRects needs |dx|, |dy| <= 0.1 * (min widths + min heights) <= 0.1 * (w_i + h_i), so a
    // pair with 10 * |dx| > w_i + h_i can never be similar (exact superset; the fp64 test decides the rest)
    const int lim = ri.width + ri.height;
    for (int j = max(j0, i + 1); j < j1; ++j) {
        const int4 q = __ldg(reinterpret_cast<const int4*>(raw + j));
        if (abs(ri.x - q.x) * 10 > lim || abs(ri.y - q.y) * 10 > lim) continue;
        ugd_rect rj; rj.x = q.x; rj.y = q.y; rj.width = q.z; rj.height = q.w;
        if (!similar_rects(ri, rj)) continue;
        int a = i, c = j;
        for (;;) {                                           // unite: the larger root is hung under the smaller one
            a = uf_find(parent, a); c = uf_find(parent, c);
            if (a == c) break;
            const int hi = max(a, c), lo = min(a, c);
            if (atomicCAS(&parent[hi], hi, lo) == hi) break;
        }
    }
}

__global__ void __launch_bounds__(256) k_group(DevBufs b) {
    __shared__ int s_nkept, s_nfinal, s_ncls;
    __shared__ unsigned long long s_red[8];
    const int frame = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int total = b.rawcount[frame];
    const int n = min(total, b.rawcap);
    unsigned long long* ctr = b.counters + (size_t)frame * kCtrPerFrame;
    const ugd_rect* raw = b.raw + (size_t)frame * b.rawcap;
    int* scratch = b.grp_scratch + (size_t)frame * b.grp_stride;
    int* label = scratch;                    // [rawcap]
    int* acc = scratch + b.rawcap;           // [rawcap][4] class sums -> class mean rect
    int* cnt = scratch + 5 * b.rawcap;       // [rawcap]
    ugd_rect* kept = reinterpret_cast<ugd_rect*>(scratch + 6 * b.rawcap);   // [rawcap]
    int* clist = scratch + 10 * b.rawcap;    // [rawcap] roots of the classes with more than g members
    const int g = b.min_neighbors;
    if (tid == 0) {
        s_nkept = 0; s_nfinal = 0; s_ncls = 0;
        ctr[CTR_RAW] = (unsigned long long)total;
        if (total > b.rawcap) ctr[CTR_RAW_OVERFLOW] = 1ull;
    }
    __syncthreads();

    if (g <= 0) {                                            // groupRectangles leaves the list unchanged
        for (int i = tid; i < n; i += 256) kept[i] = raw[i];
        if (tid == 0) s_nkept = n;
        __syncthreads();
    } else {
        // ---- connected components of the similarity graph: the links were made by k_group_link
        //      (lock-free union-find, larger root -> smaller root); flatten to the class minimum ----
        for (int i = tid; i < n; i += 256) {
            int l = ((volatile int*)label)[i];
            int r = ((volatile int*)label)[l];
            while (r != l) { l = r; r = ((volatile int*)label)[l]; }
            label[i] = l;                                    // roots never change here, so concurrent writes are benign
        }
        __syncthreads();
        // ---- per-class integer sums -> rounded means ----
        for (int i = tid; i < n; i += 256) { acc[4 * i] = acc[4 * i + 1] = acc[4 * i + 2] = acc[4 * i + 3] = 0; cnt[i] = 0; }
        __syncthreads();
        for (int i = tid; i < n; i += 256) {
            const int l = label[i];
            const ugd_rect r = raw[i];
            atomicAdd(&acc[4 * l], r.x); atomicAdd(&acc[4 * l + 1], r.y);
            atomicAdd(&acc[4 * l + 2], r.width); atomicAdd(&acc[4 * l + 3], r.height);
            atomicAdd(&cnt[l], 1);
        }
        __syncthreads();
        for (int i = tid; i < n; i += 256) {
            if (label[i] != i) continue;
            const float s = __fdiv_rn(1.f, (float)cnt[i]);
#pragma unroll
            for (int k = 0; k < 4; ++k) acc[4 * i + k] = __float2int_rn(__fmul_rn((float)acc[4 * i + k], s));
        }
        __syncthreads();
        // ---- keep classes with more than g members that are not nested in a stronger class.  The classes that
        //      qualify are first gathered into a list, so the all-pairs test runs over them and not over every
        //      candidate (thousands of candidates give a few hundred classes) ----
        for (int i = tid; i < n; i += 256)
            if (label[i] == i && cnt[i] > g) clist[atomicAdd(&s_ncls, 1)] = i;
        __syncthreads();
        const int nc = s_ncls;
        for (int a = tid; a < nc; a += 256) {
            const int i = clist[a];
            const int n1 = cnt[i];
            const int x1 = acc[4 * i], y1 = acc[4 * i + 1], w1 = acc[4 * i + 2], h1 = acc[4 * i + 3];
            bool drop = false;
            for (int c = 0; c < nc && !drop; ++c) {
                const int j = clist[c];
                if (j == i) continue;
                const int n2 = cnt[j];
                const int x2 = acc[4 * j], y2 = acc[4 * j + 1], w2 = acc[4 * j + 2], h2 = acc[4 * j + 3];
                const int dx = __double2int_rn(__dmul_rn((double)w2, 0.2)), dy = __double2int_rn(__dmul_rn((double)h2, 0.2));
                if (x1 >= x2 - dx && y1 >= y2 - dy && x1 + w1 <= x2 + w2 + dx && y1 + h1 <= y2 + h2 + dy &&
                    (n2 > max(3, n1) || n1 < 3))
                    drop = true;
            }
            if (!drop) {
                const int slot = atomicAdd(&s_nkept, 1);
                ugd_rect r; r.x = x1; r.y = y1; r.width = w1; r.height = h1;
                kept[slot] = r;
            }
        }
        __syncthreads();
    }

    // ---- clip to the image (A.5 last line), drop empties; then the skin-coverage test (detector.cc:33-42).  Survivors are
    //      flagged in `clist` (dead by now) instead of being compacted one by one: without the skin test a thread takes a
    //      rectangle (cascades that fire on thousands of windows keep hundreds of them); with it the block sums the mask
    //      under one rectangle at a time ----
    const int nk = s_nkept;
    const uint8_t* mask = b.mask + (size_t)frame * b.px_stride;
    if (!b.use_skin) {
        for (int k = tid; k < nk; k += 256) {
            ugd_rect r = kept[k];
            const int x0 = max(r.x, 0), y0 = max(r.y, 0), x1 = min(r.x + r.width, b.w), y1 = min(r.y + r.height, b.h);
            const bool ok = x1 > x0 && y1 > y0;
            if (!b.no_clip) { r.x = x0; r.y = y0; r.width = x1 - x0; r.height = y1 - y0; kept[k] = r; }   // (2.4 legacy mode does not clip)
            clist[k] = ok ? 1 : 0;
            if (ok) atomicAdd(&s_nfinal, 1);
        }
    } else {
        for (int k = 0; k < nk; ++k) {
            ugd_rect r = kept[k];
            const int x0 = max(r.x, 0), y0 = max(r.y, 0), x1 = min(r.x + r.width, b.w), y1 = min(r.y + r.height, b.h);
            bool ok = x1 > x0 && y1 > y0;                   // block-uniform
            if (ok) {
                unsigned long long sum = 0;
                for (int yy = warp; yy < y1 - y0; yy += 8) {
                    const uint8_t* row = mask + (size_t)(y0 + yy) * b.w + x0;
                    for (int xx = lane; xx < x1 - x0; xx += 32) sum += row[xx];
                }
#pragma unroll
                for (int d = 16; d; d >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, d);
                if (lane == 0) s_red[warp] = sum;
                __syncthreads();
                sum = 0;
#pragma unroll
                for (int wv = 0; wv < 8; ++wv) sum += s_red[wv];
                // cv::mean = sum * (1/N); norm of a 1-channel mean is the mean itself
                const double mean = __dmul_rn((double)sum, __ddiv_rn(1.0, __dmul_rn((double)(x1 - x0), (double)(y1 - y0))));
                ok = !(__ddiv_rn(mean, 256.0) < 0.8);
                __syncthreads();                            // s_red is reused by the next rectangle
            }
            if (tid == 0) {
                if (!b.no_clip) { r.x = x0; r.y = y0; r.width = x1 - x0; r.height = y1 - y0; kept[k] = r; }   // (the mask mean is over the part inside)
                clist[k] = ok ? 1 : 0;
                if (ok) ++s_nfinal;
            }
        }
    }
    __syncthreads();
    // ---- canonical order (rank sort over the flagged rectangles; cv2's own order depends on thread timing) ----
    const int nf = s_nfinal;
    ugd_rect* out = b.out + (size_t)frame * b.outcap;
    for (int i = tid; i < nk; i += 256) {
        if (!clist[i]) continue;
        const ugd_rect ri = kept[i];
        int rank = 0;
        for (int j = 0; j < nk; ++j) {
            if (!clist[j]) continue;
            const ugd_rect rj = kept[j];
            if (rect_less(rj, ri) || (!rect_less(ri, rj) && j < i)) ++rank;
        }
        if (rank < b.outcap) out[rank] = ri;
    }
    if (tid == 0) { b.outcount[frame] = nf; ctr[CTR_GROUPED] = (unsigned long long)nf; }
}

void launch_group(const DevBufs& b, cudaStream_t st) {
    if (b.min_neighbors > 0) {
        k_group_init<<<dim3((b.rawcap + 255) / 256, b.nframes), 256, 0, st>>>(b);
        k_group_link<<<dim3((b.rawcap + kLinkI - 1) / kLinkI, kLinkJ, b.nframes), kLinkI, 0, st>>>(b);
    }
    k_group<<<b.nframes, 256, 0, st>>>(b);
}

// ------------------------------------------------------------------------------------------------
// Multi-GPU gather, root side: the peers' fixed slots {rect[cap] per frame, then count per frame}
// arrive over NVLink in `recv` ([world] messages); this kernel packs them into one dense list so that the
// root's device-to-host copy moves the rectangles that exist, not world x frames x cap slots.
//   packed = { int total; int counts[world * slot_frames]; ugd_rect rects[total] }   (rects 16-byte aligned)
// One block; entries of rank `skip_rank` (the root's own frames: delivered by its local path) count as empty.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(1024) k_pack_lists(const uint8_t* __restrict__ recv, size_t msg_bytes, int world, int slot_frames, int cap,
                                                      int skip_rank, int* __restrict__ packed, size_t rect_off_bytes) {
    __shared__ int s_warp[32];
    __shared__ int s_base;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int n = world * slot_frames;
    ugd_rect* dst = reinterpret_cast<ugd_rect*>(reinterpret_cast<uint8_t*>(packed) + rect_off_bytes);
    if (tid == 0) s_base = 0;
    __syncthreads();
    for (int e0 = 0; e0 < n; e0 += 1024) {
        const int e = e0 + tid;
        int cnt = -1, take = 0;
        const ugd_rect* src = nullptr;
        if (e < n) {
            const int r = e / slot_frames, f = e - r * slot_frames;
            const uint8_t* msg = recv + (size_t)r * msg_bytes;
            cnt = reinterpret_cast<const int*>(msg + (size_t)slot_frames * cap * sizeof(ugd_rect))[f];
            src = reinterpret_cast<const ugd_rect*>(msg) + (size_t)f * cap;
            if (r == skip_rank) cnt = -1;
            take = max(min(cnt, cap), 0);
            packed[1 + e] = cnt;
        }
        int incl = take;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { const int v = __shfl_up_sync(0xffffffffu, incl, d); if (lane >= d) incl += v; }
        if (lane == 31) s_warp[wid] = incl;
        __syncthreads();
        if (wid == 0) {
            int v = s_warp[lane], w = v;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) { const int u = __shfl_up_sync(0xffffffffu, w, d); if (lane >= d) w += u; }
            s_warp[lane] = w - v;                       // exclusive over warps
        }
        __syncthreads();
        const int off = s_base + s_warp[wid] + incl - take;
        for (int k = 0; k < take; ++k) dst[off + k] = src[k];
        __syncthreads();
        if (tid == 1023) s_base = off + take;
        __syncthreads();
    }
    if (tid == 0) packed[0] = s_base;
}

void launch_pack_lists(const uint8_t* recv, size_t msg_bytes, int world, int slot_frames, int cap, int skip_rank, int* packed,
                       size_t rect_off_bytes, cudaStream_t st) {
    k_pack_lists<<<1, 1024, 0, st>>>(recv, msg_bytes, world, slot_frames, cap, skip_rank, packed, rect_off_bytes);
}

}  // namespace ugd
