// Window scan and cascade stage evaluation for sm_100a.
//
//   k_scan : one block = 8 grid rows of one level of one frame (one warp per row), walking the rows
//            left to right in tiles of 32 x 8 windows.  Per tile:
//              1. the (8*step+win) x (32*step+win) patch of the integral image S is staged in shared
//                 memory (coalesced loads; for grid step 2 even/odd columns are de-interleaved so
//                 that the 32 lanes of a warp always hit 32 consecutive banks);
//              2. every lane computes the variance normalisation and stage 0 of its own window;
//              3. the oracle's sequential x-skip rule is resolved in closed form with one ballot and
//                 a run-parity carried along the row in a register;
//              4. surviving windows are ballot/popc-compacted ACROSS the block into a dense list, and
//                 stage by stage the list is re-evaluated and re-compacted, so every warp that works
//                 has (almost) all lanes on live windows.  Node data is broadcast from __constant__
//                 as merged corner lists (6/8/9 LDS + IMAD per weak classifier);
//              5. when fewer than `tile_min` windows are left (or the in-tile stages are exhausted)
//                 the survivors go to the frame's global work queue.
//   k_tail : persistent warps pull queued windows; the window's (win+1)^2 patch of S is staged in
//            shared memory once, then the 32 lanes evaluate 32 weak classifiers of a stage in
//            parallel and a warp reduction forms the stage sum.
//
// Exactness: feature sums are int32 (exact: weights are small integers), the one rounding step is
// val = f32(float(r) * vnf); stage sums are fp64 sums of fp32 leaves, exact for the shipped cascades
// (ugd_cascade_info.sum_exact), hence independent of summation order.
#include "kernels.cuh"

namespace ugd {

__constant__ DNode c_dn[kMaxDenseNodes];
__constant__ int c_stage_first_node[kMaxDenseStages + 1];   // first in-tile node of stage s (s <= dense_stages)
__constant__ float c_stage_thr[kMaxDenseStages];
__constant__ ScanGeom c_geom;

ScanGeom make_scan_geom(int win_w, int win_h) {
    ScanGeom g{};
    for (int si = 0; si < 2; ++si) {
        const int step = si + 1;
        g.TC[si] = 31 * step + win_w + 1;
        g.TR[si] = (kScanRows - 1) * step + win_h + 1;
        g.TCH[si] = si ? (g.TC[si] + 1) / 2 : 0;
        g.TP[si] = si ? 2 * g.TCH[si] : g.TC[si];
    }
    g.tile_ints = std::max(g.TR[0] * g.TP[0], g.TR[1] * g.TP[1]);
    return g;
}

static inline int32_t tile_off(const ScanGeom& g, int si, int x, int y) {     // byte offset
    return 4 * (si ? y * g.TP[1] + (x & 1) * g.TCH[1] + (x >> 1) : y * g.TP[0] + x);
}

cudaError_t upload_scan_constants(const Cascade& c, int dense_stages, cudaStream_t st) {
    const int nn = c.stage_first_node[dense_stages];
    if (nn > kMaxDenseNodes || dense_stages > kMaxDenseStages) return cudaErrorInvalidValue;
    ScanGeom g = make_scan_geom(c.win_w, c.win_h);
    for (int si = 0; si < 2; ++si) {
        g.norm_off[si][0] = tile_off(g, si, 1, 1);
        g.norm_off[si][1] = tile_off(g, si, c.win_w - 1, 1);
        g.norm_off[si][2] = tile_off(g, si, 1, c.win_h - 1);
        g.norm_off[si][3] = tile_off(g, si, c.win_w - 1, c.win_h - 1);
    }
    std::vector<DNode> dn((size_t)std::max(nn, 1));
    for (int t = 0; t < c.ntrees(); ++t) {
        const int base = c.tree_first_node[t];
        for (int k = 0; k < c.tree_nnodes[t]; ++k) {
            const int n = base + k;
            if (n >= nn) continue;
            const CornerNode& cn = c.corners[n];
            const HostNode& h = c.nodes[n];
            DNode& d = dn[n];
            memset(&d, 0, sizeof d);
            for (int j = 0; j < 12; ++j) {
                d.coef[j] = cn.coef[j];
                for (int si = 0; si < 2; ++si) d.off[si][j] = tile_off(g, si, cn.cx[j], cn.cy[j]);
            }
            d.meta = (uint32_t)cn.nc | (uint32_t)h.left << 8 | (uint32_t)h.right << 16 |
                     (k == c.tree_nnodes[t] - 1 ? 1u << 24 : 0u);
            d.thr = h.thr; d.lval = h.left_val; d.rval = h.right_val;
        }
    }
    cudaError_t e;
    if ((e = cudaMemcpyToSymbolAsync(c_dn, dn.data(), sizeof(DNode) * dn.size(), 0, cudaMemcpyHostToDevice, st))) return e;
    if ((e = cudaMemcpyToSymbolAsync(c_stage_first_node, c.stage_first_node.data(), sizeof(int) * (size_t)(dense_stages + 1), 0, cudaMemcpyHostToDevice, st))) return e;
    if ((e = cudaMemcpyToSymbolAsync(c_stage_thr, c.stage_thr.data(), sizeof(float) * (size_t)dense_stages, 0, cudaMemcpyHostToDevice, st))) return e;
    if ((e = cudaMemcpyToSymbolAsync(c_geom, &g, sizeof g, 0, cudaMemcpyHostToDevice, st))) return e;
    return cudaStreamSynchronize(st);      // dn is a local: the copy must finish before it goes away
}

__device__ __forceinline__ int near_thr(double sum, float thr) {
    return fabs(sum - (double)thr) <= 1e-5 * fabs((double)thr);
}

// One in-tile stage for one window.  `t` points at the window's origin inside the shared tile; the
// loop and every node field are warp-uniform, so node data is broadcast from the constant bank.
#define TILE_AT(d, o, j) ((d).coef[j] * *reinterpret_cast<const int*>(t + (o)[j]))
// feature response r of one node: two independent IMAD chains, extra corners behind uniform branches
__device__ __forceinline__ int node_resp(const DNode& d, int si, const char* __restrict__ t) {
    const int32_t* o = d.off[si];
    int a = TILE_AT(d, o, 0) + TILE_AT(d, o, 1) + TILE_AT(d, o, 2);
    int c = TILE_AT(d, o, 3) + TILE_AT(d, o, 4) + TILE_AT(d, o, 5);
    const int nc = d.meta & 255u;
    if (nc > 6) {
        a += TILE_AT(d, o, 6); c += TILE_AT(d, o, 7);
        if (nc > 8) {
            a += TILE_AT(d, o, 8);
            if (nc > 9) { c += TILE_AT(d, o, 9); a += TILE_AT(d, o, 10); c += TILE_AT(d, o, 11); }
        }
    }
    return a + c;
}

template <bool STUMPS>
__device__ __forceinline__ double tile_stage(int n0, int n1, int si, const char* __restrict__ t, float vnf) {
    if (STUMPS) {
        // two nodes per trip, separate accumulators: twice the independent work in flight per warp
        double s0 = 0.0, s1 = 0.0;
        int n = n0;
        for (; n + 1 < n1; n += 2) {
            const DNode& d0 = c_dn[n];
            const DNode& d1 = c_dn[n + 1];
            const int r0 = node_resp(d0, si, t), r1 = node_resp(d1, si, t);
            const bool l0 = __fmul_rn(__int2float_rn(r0), vnf) < d0.thr;
            const bool l1 = __fmul_rn(__int2float_rn(r1), vnf) < d1.thr;
            s0 += (double)(l0 ? d0.lval : d0.rval);
            s1 += (double)(l1 ? d1.lval : d1.rval);
        }
        if (n < n1) {
            const DNode& d0 = c_dn[n];
            const bool l0 = __fmul_rn(__int2float_rn(node_resp(d0, si, t)), vnf) < d0.thr;
            s0 += (double)(l0 ? d0.lval : d0.rval);
        }
        return s0 + s1;
    }
    double sum = 0.0;
    float leaf = 0.f;
    int at = 0, k = 0;
    for (int n = n0; n < n1; ++n) {
        const DNode& d = c_dn[n];
        const uint32_t meta = d.meta;
        const bool lt = __fmul_rn(__int2float_rn(node_resp(d, si, t)), vnf) < d.thr;
        if (at == k) {
            const int go = lt ? ((meta >> 8) & 255u) : ((meta >> 16) & 255u);
            leaf = lt ? d.lval : d.rval;
            at = go ? go : 255;
        }
        if (meta >> 24) { sum += (double)leaf; at = 0; k = 0; } else ++k;
    }
    return sum;
}
#undef TILE_AT

__device__ __forceinline__ void emit_raw(const DevBufs& b, int frame, float scale, int rw, int rh, int x, int y, double sum) {
    int slot = atomicAdd(&b.rawcount[frame], 1);
    if (slot < b.rawcap) {
        ugd_rect r;
        r.x = __float2int_rn(__fmul_rn((float)x, scale));
        r.y = __float2int_rn(__fmul_rn((float)y, scale));
        r.width = rw; r.height = rh;
        b.raw[(size_t)frame * b.rawcap + slot] = r;
        b.rawsum[(size_t)frame * b.rawcap + slot] = sum;
    }
}

template <bool STUMPS>
__global__ void __launch_bounds__(kScanRows * 32) k_scan(DevBufs b, const WorkItem* __restrict__ wk) {
    constexpr int NW = kScanRows, NT = NW * 32;
    extern __shared__ int s_tile[];
    __shared__ float s_vnf[NT];
    __shared__ uint16_t s_id[2][NT];
    __shared__ int s_n[3];                             // rotating list counters
    __shared__ int s_qbase;
    __shared__ unsigned int s_ctr[4];                  // evaluated, lowvar, near, weak

    const WorkItem it = wk[blockIdx.x];
    const int frame = blockIdx.y, tid = threadIdx.x, w = tid >> 5, lane = tid & 31;
    const LevelInfo& Lg = b.levels[it.level];
    const int sw = Lg.sw, sh = Lg.sh, P = Lg.P, step = Lg.step, nxg = Lg.nxg, nyg = Lg.nyg, nxg_pad = Lg.nxg_pad;
    const float scale = Lg.scale;
    const int rect_w = Lg.rect_w, rect_h = Lg.rect_h;
    const uint32_t pos_off = Lg.pos_off;
    const int si = step - 1;
    const int TP = c_geom.TP[si], TCH = c_geom.TCH[si], TC = c_geom.TC[si], TR = c_geom.TR[si];
    const int r0 = it.first;
    const int row = r0 + w;
    const bool row_valid = row < nyg;
    const int32_t* S = b.S + (size_t)frame * b.int_stride + Lg.int_off;
    const int32_t* Q = b.SQ + (size_t)frame * b.int_stride + Lg.int_off;
    const int ow = b.win_w, oh = b.win_h, D = b.dense_stages, nstages = b.nstages;
    const double area = (double)((ow - 2) * (oh - 2));
    const unsigned below = lane ? (0xffffffffu >> (32 - lane)) : 0u;
    int16_t* map = (b.stage_map && frame == 0) ? b.stage_map + pos_off : nullptr;
    if (tid < 4) s_ctr[tid] = 0;
    int run_in = 0;                                    // stage-0 failure run ending just before this segment
    unsigned weak = 0;                                 // lane 0 of each warp tallies weak classifiers

    for (int xi0 = 0; xi0 < nxg; xi0 += 32) {
        __syncthreads();                               // previous tile fully consumed
        if (tid == 0) s_n[0] = 0;
        // ---- 1. stage the S tile ----
        {
            const int X0 = xi0 * step, Y0 = r0 * step;
            for (int r = w; r < TR; r += NW) {
                const int32_t* src = S + (size_t)min(Y0 + r, sh) * P;
                int* dst = s_tile + r * TP;
                for (int c = lane; c < TC; c += 32) {
                    const int v = src[min(X0 + c, sw)];
                    dst[si ? (c & 1) * TCH + (c >> 1) : c] = v;
                }
            }
        }
        __syncthreads();
        // ---- 2. variance normalisation + stage 0, every lane ----
        const int xi = xi0 + lane;
        const bool valid = row_valid && xi < nxg;
        const int x = valid ? xi * step : 0, y = valid ? row * step : 0;
        const char* t = reinterpret_cast<const char*>(s_tile + (w * step) * TP + lane);
        float vnf = 1.f;
        bool ok = false;
        {
            const int32_t* no = c_geom.norm_off[si];
            const int vs = *reinterpret_cast<const int*>(t + no[0]) - *reinterpret_cast<const int*>(t + no[1]) -
                           *reinterpret_cast<const int*>(t + no[2]) + *reinterpret_cast<const int*>(t + no[3]);
            const int32_t* q = Q + (size_t)y * P + x;
            const uint32_t vq = (uint32_t)q[P + 1] - (uint32_t)q[P + ow - 1] - (uint32_t)q[(oh - 1) * P + 1] + (uint32_t)q[(oh - 1) * P + ow - 1];
            double nf = area * (double)vq - (double)vs * (double)vs;
            if (nf > 0.0) {
                nf = sqrt(nf);
                vnf = (float)(1.0 / nf);
                ok = area * (double)vnf < 1e-1;
            }
            if (!ok) vnf = 1.f;
        }
        double sum = tile_stage<STUMPS>(c_stage_first_node[0], c_stage_first_node[1], si, t, vnf);
        const float thr0 = c_stage_thr[0];
        const bool pass0 = !(sum < (double)thr0);
        // ---- 3. x-skip rule in closed form: position i is evaluated iff the run of stage-0 failures
        //         ending at i-1 has even length (a failure is only seen at evaluated positions) ----
        const bool f0 = valid && ok && !pass0;
        const unsigned F = __ballot_sync(0xffffffffu, f0);
        const unsigned inv = ~F & below;
        const int run = inv ? (lane - 1 - (31 - __clz((int)inv))) : (lane + run_in);
        run_in = (F == 0xffffffffu) ? run_in + 32 : __clz((int)~F);
        const bool evaluated = valid && !(run & 1);
        bool alive = evaluated && ok && pass0;
        {
            const unsigned em = __ballot_sync(0xffffffffu, evaluated);
            const unsigned lm = __ballot_sync(0xffffffffu, evaluated && !ok);
            const unsigned s0 = __ballot_sync(0xffffffffu, evaluated && ok);
            const unsigned nm = __ballot_sync(0xffffffffu, evaluated && ok && near_thr(sum, thr0));
            if (lane == 0) {
                atomicAdd(&s_ctr[0], __popc(em));
                if (lm) atomicAdd(&s_ctr[1], __popc(lm));
                if (nm) atomicAdd(&s_ctr[2], __popc(nm));
                weak += __popc(s0) * (unsigned)(c_stage_first_node[1] - c_stage_first_node[0]);
            }
        }
        if (map && valid) {
            if (!evaluated) map[row * nxg_pad + xi] = kMapSkip;
            else if (!ok) map[row * nxg_pad + xi] = kMapLowVar;
            else if (!pass0) map[row * nxg_pad + xi] = 0;
        }
        s_vnf[tid] = vnf;
        // ---- 4. block-wide compaction, then stage by stage on the dense list ----
        // The list of n live windows is cut into k = ceil(n/32) chunks, one warp each; list slots are
        // handed out with one shared-memory atomic per warp, so a stage costs a single barrier.
        int cur = 0, n, s = 1, rot = 0;                // rot: which of the three list counters is being filled
        {
            const unsigned am = __ballot_sync(0xffffffffu, alive);
            int base = 0;
            if (lane == 0 && am) base = atomicAdd(&s_n[0], __popc(am));
            base = __shfl_sync(0xffffffffu, base, 0);
            if (alive) s_id[0][base + __popc(am & below)] = (uint16_t)tid;
            if (tid == 0) { s_n[1] = 0; }
            __syncthreads();
            n = s_n[0];
        }
        if (nstages == 1 && alive) {                   // one-stage cascade: accepted right here
            emit_raw(b, frame, scale, rect_w, rect_h, x, y, sum);
            if (map) map[row * nxg_pad + xi] = 1;
        }
        while (s < D && n > 0 && !(n < b.tile_min && s >= 2)) {
            const int k = (n + 31) >> 5;
            const int nxt = rot == 2 ? 0 : rot + 1, nxt2 = nxt == 2 ? 0 : nxt + 1;
            if (tid == 0) s_n[nxt2] = 0;               // counter of the round after next: nobody reads it any more
            bool pass = false;
            int my = 0;
            if (w < k) {                               // warp-uniform: this warp owns a chunk of the list
                const int i = w * 32 + lane;
                const bool act = i < n;
                my = s_id[cur][act ? i : 0];
                const char* tw = reinterpret_cast<const char*>(s_tile + ((my >> 5) * step) * TP + (my & 31));
                const int b0 = c_stage_first_node[s], b1 = c_stage_first_node[s + 1];
                sum = tile_stage<STUMPS>(b0, b1, si, tw, s_vnf[my]);
                if (lane == 0) weak += (unsigned)min(32, n - w * 32) * (unsigned)(b1 - b0);
                const float thr = c_stage_thr[s];
                pass = act && !(sum < (double)thr);
                const unsigned nm = __ballot_sync(0xffffffffu, act && near_thr(sum, thr));
                if (lane == 0 && nm) atomicAdd(&s_ctr[2], __popc(nm));
                const int mpos = (r0 + (my >> 5)) * nxg_pad + xi0 + (my & 31);
                if (act && !pass && map) map[mpos] = (int16_t)s;
                if (pass && s + 1 == nstages) {
                    emit_raw(b, frame, scale, rect_w, rect_h, (xi0 + (my & 31)) * step, (r0 + (my >> 5)) * step, sum);
                    if (map) map[mpos] = (int16_t)nstages;
                    pass = false;                       // accepted: nothing left to do for it
                }
            }
            const unsigned pm = __ballot_sync(0xffffffffu, pass);
            int base = 0;
            if (lane == 0 && pm) base = atomicAdd(&s_n[nxt], __popc(pm));
            base = __shfl_sync(0xffffffffu, base, 0);
            if (pass) s_id[cur ^ 1][base + __popc(pm & below)] = (uint16_t)my;
            __syncthreads();
            n = s_n[nxt];
            cur ^= 1; rot = nxt; ++s;
        }
        // ---- 5. survivors -> the frame's work queue (they resume at stage s) ----
        if (n > 0 && s < nstages) {
            if (tid == 0) s_qbase = atomicAdd(&b.qcount[frame], n);
            __syncthreads();
            if (tid < n) {
                const int my = s_id[cur][tid];
                const int qx = (xi0 + (my & 31)) * step, qy = (r0 + (my >> 5)) * step;
                b.queue[(size_t)frame * b.pos_stride + s_qbase + tid] =
                    make_uint4((uint32_t)qx | (uint32_t)qy << 12 | (uint32_t)it.level << 24, __float_as_uint(s_vnf[my]), (uint32_t)s, 0u);
            }
        }
    }
    if (lane == 0 && weak) atomicAdd(&s_ctr[3], weak);
    __syncthreads();
    if (tid < 4 && s_ctr[tid]) {
        const int slot[4] = {CTR_EVALUATED, CTR_LOWVAR, CTR_NEAR, CTR_WEAK};
        atomicAdd(&b.counters[(size_t)frame * kCtrPerFrame + slot[tid]], (unsigned long long)s_ctr[tid]);
    }
}

static int g_scan_smem = 0;

void launch_scan(const DevBufs& b, const WorkItem* wk, int nblocks, cudaStream_t st) {
    if (!nblocks) return;
    const size_t smem = (size_t)make_scan_geom(b.win_w, b.win_h).tile_ints * sizeof(int);
    if (b.stumps_only) k_scan<true><<<dim3(nblocks, b.nframes), kScanRows * 32, smem, st>>>(b, wk);
    else k_scan<false><<<dim3(nblocks, b.nframes), kScanRows * 32, smem, st>>>(b, wk);
}

// ------------------------------------------------------------------------------------------------
// Tail: one warp per queued window, lanes over the weak classifiers of a stage.
// Node data: merged corner lists whose patch BYTE offsets were computed at load time; two offsets
// share a 32-bit word, so one add of (patch base | patch base << 16) rebases both.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ int lds32(uint32_t addr) {
    int v;
    asm volatile("ld.shared.b32 %0, [%1];" : "=r"(v) : "r"(addr));
    return v;
}
__device__ __forceinline__ int sx8(uint32_t w, int byte) { return (int)(int8_t)(w >> (8 * byte)); }

// response of tail node n against the patch whose shared-memory byte address is packed in pb2
__device__ __forceinline__ int tail_resp(const DevBufs& b, int n, uint32_t pb2, uint32_t nc, uint4& C) {
    const uint4 A = __ldg(b.nodeA + n);
    const uint4 B = __ldg(b.nodeB + n);
    C = __ldg(b.nodeC + n);
    const uint32_t o01 = A.x + pb2, o23 = A.y + pb2, o45 = A.z + pb2;
    int a = sx8(B.z, 0) * lds32(o01 & 0xffffu) + sx8(B.z, 1) * lds32(o01 >> 16) + sx8(B.z, 2) * lds32(o23 & 0xffffu);
    int c = sx8(B.z, 3) * lds32(o23 >> 16) + sx8(B.w, 0) * lds32(o45 & 0xffffu) + sx8(B.w, 1) * lds32(o45 >> 16);
    if (nc > 6) {
        const uint32_t o67 = A.w + pb2;
        a += sx8(B.w, 2) * lds32(o67 & 0xffffu);
        c += sx8(B.w, 3) * lds32(o67 >> 16);
        if (nc > 8) {
            const uint32_t o89 = B.x + pb2;
            a += sx8(C.x, 0) * lds32(o89 & 0xffffu);
            if (nc > 9) {
                const uint32_t oab = B.y + pb2;
                c += sx8(C.x, 1) * lds32(o89 >> 16);
                a += sx8(C.x, 2) * lds32(oab & 0xffffu);
                c += sx8(C.x, 3) * lds32(oab >> 16);
            }
        }
    }
    return a + c;
}

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int d = 16; d; d >>= 1) v += __shfl_xor_sync(0xffffffffu, v, d);
    return v;
}

__global__ void __launch_bounds__(256) k_tail(DevBufs b) {
    extern __shared__ int32_t s_patch[];                       // tail_warps x (win_h+1)*(win_w+1) ints
    const int frame = blockIdx.y, tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int nwarps = blockDim.x >> 5;
    const int PW = b.win_w + 1, PH = b.win_h + 1, PN = PW * PH;
    int32_t* patch = s_patch + warp * PN;
    const uint32_t pbase = (uint32_t)__cvta_generic_to_shared(patch);
    const uint32_t pb2 = pbase | pbase << 16;                  // all patches sit below 64 KB (tail_warps)
    const int qn = min(b.qcount[frame], (int)b.pos_stride);
    const uint4* queue = b.queue + (size_t)frame * b.pos_stride;
    const int nstages = b.nstages;
    const bool stumps = b.stumps_only != 0;
    unsigned near_cnt = 0;
    unsigned long long weak_cnt = 0;                          // per-lane tally of evaluated nodes
    for (int i = blockIdx.x * nwarps + warp; i < qn; i += gridDim.x * nwarps) {
        const uint4 e = queue[i];
        const int x = e.x & 4095u, y = (e.x >> 12) & 4095u, level = e.x >> 24;
        const float vnf = __uint_as_float(e.y);
        const LevelInfo& L = b.levels[level];
        const int P = L.P;
        const int32_t* S = b.S + (size_t)frame * b.int_stride + L.int_off + (size_t)y * P + x;
        __syncwarp();
        {
            // flattened patch index -> (row, col) incrementally; four independent loads in flight per lane
            int r = 0, c = lane;
            while (c >= PW) { c -= PW; ++r; }
            for (int idx = lane; idx < PN; idx += 128) {
                int off[4], v[4];
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    off[u] = r * P + c;
                    c += 32;
                    while (c >= PW) { c -= PW; ++r; }
                }
#pragma unroll
                for (int u = 0; u < 4; ++u) v[u] = (idx + 32 * u < PN) ? __ldg(S + off[u]) : 0;
#pragma unroll
                for (int u = 0; u < 4; ++u) if (idx + 32 * u < PN) patch[idx + 32 * u] = v[u];
            }
        }
        __syncwarp();
        int s = (int)e.z;
        double sum = 0.0;
        for (; s < nstages; ++s) {
            const int t0 = b.stage_first_tree[s], nt = b.stage_ntrees[s];
            double part = 0.0;
            if (stumps) {
                const int n0 = b.stage_first_node[s];
                for (int t = lane; t < nt; t += 32) {
                    const int n = n0 + t;
                    uint4 C;
                    const int r = tail_resp(b, n, pb2, __ldg(b.nodeM + n) & 255u, C);
                    const bool lt = __fmul_rn(__int2float_rn(r), vnf) < __uint_as_float(C.y);
                    part += (double)__uint_as_float(lt ? C.z : C.w);
                    ++weak_cnt;
                }
            } else {
                for (int t = lane; t < nt; t += 32) {
                    const int base = b.tree_first_node[t0 + t];
                    int at = 0;
                    float leaf;
                    for (;;) {
                        const uint32_t M = __ldg(b.nodeM + base + at);
                        uint4 C;
                        const int r = tail_resp(b, base + at, pb2, M & 255u, C);
                        ++weak_cnt;
                        const bool lt = __fmul_rn(__int2float_rn(r), vnf) < __uint_as_float(C.y);
                        leaf = __uint_as_float(lt ? C.z : C.w);
                        const int go = lt ? ((M >> 8) & 255u) : ((M >> 16) & 255u);
                        if (!go) break;
                        at = go;
                    }
                    part += (double)leaf;
                }
            }
            sum = warp_sum(part);
            const float thr = b.stage_thr[s];
            if (lane == 0 && near_thr(sum, thr)) ++near_cnt;
            if (sum < (double)thr) break;
        }
        if (lane == 0) {
            if (s == nstages) emit_raw(b, frame, L.scale, L.rect_w, L.rect_h, x, y, sum);
            if (b.stage_map && frame == 0) {
                const int step = L.step;
                b.stage_map[L.pos_off + (y / step) * L.nxg_pad + x / step] = (int16_t)s;
            }
        }
    }
    if (lane == 0 && near_cnt) atomicAdd(&b.counters[(size_t)frame * kCtrPerFrame + CTR_NEAR], (unsigned long long)near_cnt);
#pragma unroll
    for (int d = 16; d; d >>= 1) weak_cnt += __shfl_xor_sync(0xffffffffu, weak_cnt, d);
    if (lane == 0 && weak_cnt) atomicAdd(&b.counters[(size_t)frame * kCtrPerFrame + CTR_WEAK], weak_cnt);
}

static int g_tail_smem = 0;

cudaError_t init_kernel_attributes(int win_w, int win_h) {
    int smem = 8 * (win_w + 1) * (win_h + 1) * (int)sizeof(int32_t);
    if (smem > g_tail_smem) {
        cudaError_t e = cudaFuncSetAttribute(k_tail, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        if (e != cudaSuccess) return e;
        g_tail_smem = smem;
    }
    const int scan_smem = make_scan_geom(win_w, win_h).tile_ints * (int)sizeof(int);
    if (scan_smem > g_scan_smem) {
        cudaError_t e = cudaFuncSetAttribute(k_scan<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, scan_smem);
        if (e == cudaSuccess) e = cudaFuncSetAttribute(k_scan<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, scan_smem);
        if (e != cudaSuccess) return e;
        g_scan_smem = scan_smem;
    }
    return cudaSuccess;
}

void launch_tail(const DevBufs& b, int blocks_per_frame, cudaStream_t st) {
    int smem = b.tail_warps * (b.win_w + 1) * (b.win_h + 1) * (int)sizeof(int32_t);
    k_tail<<<dim3(blocks_per_frame, b.nframes), b.tail_warps * 32, smem, st>>>(b);
}

}  // namespace ugd
