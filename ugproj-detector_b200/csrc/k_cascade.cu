// Window scan and cascade stage evaluation for sm_100a.
//
//   k_stage0 : one lane per grid position: variance normalisation + stage 0.  Lanes of a warp are
//              x-adjacent windows, so every corner gather is one coalesced 128/256-byte request.
//   k_dense  : resolves the oracle's sequential x-skip rule in closed form with ballots, runs the
//              next few stages densely (cascade nodes broadcast from __constant__), then ballot/popc
//              compacts the survivors into the per-frame work queue.
//   k_tail   : persistent warps pull queued windows; the window's (win+1)^2 patch of S is staged in
//              shared memory once, then the 32 lanes evaluate 32 weak classifiers of a stage in
//              parallel (no divergence, conflict-light shared-memory gathers) and a warp reduction
//              forms the stage sum.  Rejected windows free the whole warp at the next stage boundary.
//
// Exactness: feature sums are int32 (exact because weights are small integers), the one rounding
// step is val = f32(float(r) * vnf); stage sums are fp64 sums of fp32 leaves, which are exact for
// the shipped cascades (ugd_cascade_info.sum_exact), hence independent of summation order.
#include "kernels.cuh"

namespace ugd {

__constant__ PackedNode c_nodes[kMaxDenseNodes];
__constant__ int c_stage_first_node[64];   // first node of dense stage s, s <= dense_stages
__constant__ float c_stage_thr[64];

cudaError_t upload_dense_nodes(const PackedNode* nodes, int n, cudaStream_t st) {
    if (n > kMaxDenseNodes) return cudaErrorInvalidValue;
    return cudaMemcpyToSymbolAsync(c_nodes, nodes, sizeof(PackedNode) * (size_t)n, 0, cudaMemcpyHostToDevice, st);
}
cudaError_t upload_dense_stage_table(const int* first_node, const float* thr, int n, cudaStream_t st) {
    cudaError_t e = cudaMemcpyToSymbolAsync(c_stage_first_node, first_node, sizeof(int) * (size_t)n, 0, cudaMemcpyHostToDevice, st);
    if (e != cudaSuccess) return e;
    return cudaMemcpyToSymbolAsync(c_stage_thr, thr, sizeof(float) * (size_t)(n - 1), 0, cudaMemcpyHostToDevice, st);
}

__device__ __forceinline__ int rect_sum(const int32_t* __restrict__ p, uint32_t r, int pitch) {
    const int x = r & 255u, y = (r >> 8) & 255u, w = (r >> 16) & 255u, h = r >> 24;
    const int32_t* a = p + y * pitch + x;
    const int32_t* c = a + h * pitch;
    return a[0] - a[w] - c[0] + c[w];
}

__device__ __forceinline__ int sext8(uint32_t v) { return (int)(int8_t)(v & 255u); }

// Feature response of one node at the window whose integral origin is `p` (pitch in ints).
__device__ __forceinline__ float node_val(const int32_t* __restrict__ p, int pitch, uint32_t r0, uint32_t r1, uint32_t r2,
                                          uint32_t meta, uint32_t meta2, float vnf) {
    int r = sext8(meta >> 16) * rect_sum(p, r0, pitch) + sext8(meta >> 24) * rect_sum(p, r1, pitch);
    if (r2) r += sext8(meta2) * rect_sum(p, r2, pitch);
    return __fmul_rn(__int2float_rn(r), vnf);
}

// One densely evaluated stage: nodes [n0, n1) from __constant__ (uniform addresses -> broadcast).
// Trees are walked with a uniform loop over their nodes; a lane only "takes" the node it is at.
__device__ __forceinline__ double dense_stage(int n0, int n1, const int32_t* __restrict__ p, int pitch, float vnf) {
    double sum = 0.0, leaf = 0.0;
    int at = 0, k = 0;                       // lane's current node within the tree; uniform position
    for (int n = n0; n < n1; ++n) {
        const PackedNode& nd = c_nodes[n];
        const uint32_t meta = nd.meta, meta2 = nd.meta2;
        const float val = node_val(p, pitch, nd.r0, nd.r1, nd.r2, meta, meta2, vnf);
        const bool lt = val < __uint_as_float(nd.thr_bits);
        if (at == k) {
            const int go = lt ? (meta & 255u) : ((meta >> 8) & 255u);
            leaf = lt ? nd.lval : nd.rval;
            at = go ? go : 255;
        }
        if (meta2 & 0x100u) { sum += leaf; at = 0; k = 0; } else ++k;
    }
    return sum;
}

struct PosInfo { int level, row, xi; bool valid; };

__device__ __forceinline__ int near_thr(double sum, float thr) {
    return fabs(sum - (double)thr) <= 1e-5 * fabs((double)thr);
}

// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_stage0(DevBufs b, const WorkItem* __restrict__ wk) {
    const WorkItem it = wk[blockIdx.x];
    const LevelInfo& L = b.levels[it.level];
    const int frame = blockIdx.y;
    const int pos = it.first + threadIdx.x;
    const int nxg_pad = L.nxg_pad;
    const int row = pos / nxg_pad, xi = pos - row * nxg_pad;
    float* outp = b.posv + (size_t)frame * b.pos_stride + L.pos_off + pos;
    if (row >= L.nyg) return;                       // tail of the last block (rounded up to 256)
    if (xi >= L.nxg) { *outp = 0.f; return; }
    const int step = L.step, P = L.P;
    const int x = xi * step, y = row * step;
    const size_t o = (size_t)frame * b.int_stride + L.int_off + (size_t)y * P + x;
    const int32_t* S = b.S + o;
    const int32_t* Q = b.SQ + o;
    const int ow = b.win_w, oh = b.win_h;
    // A.3: variance normalisation on the inner (ow-2)x(oh-2) rectangle
    const int a0 = P + 1, a1 = P + ow - 1, a2 = (oh - 1) * P + 1, a3 = (oh - 1) * P + ow - 1;
    const int vs = S[a0] - S[a1] - S[a2] + S[a3];
    const uint32_t vq = (uint32_t)Q[a0] - (uint32_t)Q[a1] - (uint32_t)Q[a2] + (uint32_t)Q[a3];
    const double area = (double)((ow - 2) * (oh - 2));
    double nf = area * (double)vq - (double)vs * (double)vs;
    float res = 0.f;                                // 0 = low variance
    if (nf > 0.0) {
        nf = sqrt(nf);
        const float vnf = (float)(1.0 / nf);
        if (area * (double)vnf < 1e-1) {
            const double sum = dense_stage(c_stage_first_node[0], c_stage_first_node[1], S, P, vnf);
            const float thr = c_stage_thr[0];
            res = sum < (double)thr ? -1.f : vnf;
            if (near_thr(sum, thr)) atomicAdd(&b.counters[(size_t)frame * kCtrPerFrame + CTR_NEAR], 1ull);
        }
    }
    *outp = res;
}

void launch_stage0(const DevBufs& b, const WorkItem* wk, int nblocks, cudaStream_t st) {
    if (nblocks) k_stage0<<<dim3(nblocks, b.nframes), 256, 0, st>>>(b, wk);
}

// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ void emit_raw(const DevBufs& b, int frame, const LevelInfo& L, int x, int y, double sum) {
    int slot = atomicAdd(&b.rawcount[frame], 1);
    if (slot < b.rawcap) {
        const float s = L.scale;
        ugd_rect r;
        r.x = __float2int_rn(__fmul_rn((float)x, s));
        r.y = __float2int_rn(__fmul_rn((float)y, s));
        r.width = L.rect_w; r.height = L.rect_h;
        b.raw[(size_t)frame * b.rawcap + slot] = r;
        b.rawsum[(size_t)frame * b.rawcap + slot] = sum;
    }
}

__global__ void __launch_bounds__(256) k_dense(DevBufs b, const WorkItem* __restrict__ wk) {
    __shared__ unsigned int s_eval, s_lowvar, s_near, s_weak;
    const WorkItem it = wk[blockIdx.x];
    const LevelInfo& L = b.levels[it.level];
    const int frame = blockIdx.y, tid = threadIdx.x, lane = tid & 31;
    if (tid == 0) { s_eval = 0; s_lowvar = 0; s_near = 0; s_weak = 0; }
    __syncthreads();
    const int pos = it.first + tid;
    const int nxg_pad = L.nxg_pad;
    const int row = pos / nxg_pad, xi = pos - row * nxg_pad;     // row is warp-uniform (nxg_pad % 32 == 0)
    const float* pv = b.posv + (size_t)frame * b.pos_stride + L.pos_off;
    const bool valid = row < L.nyg && xi < L.nxg;
    const float v = valid ? pv[pos] : 0.f;
    if (row < L.nyg) {
        // ---- x-skip rule in closed form: position i is evaluated iff the run of stage-0 failures
        //      that ends at i-1 has even length (A.3; a failure is only seen at evaluated positions,
        //      and the position after a non-failure is always evaluated) ----
        const unsigned F = __ballot_sync(0xffffffffu, v < 0.f);
        int run_in = 0;                                          // length of the failure run ending just before this segment
        for (int back = xi - lane - 32; back >= 0; back -= 32) {
            const float pvb = pv[row * nxg_pad + back + lane];
            const unsigned m = __ballot_sync(0xffffffffu, pvb < 0.f);
            const int t = __clz((int)~m);                       // consecutive failures at the top of the word
            run_in += t;
            if (t < 32) break;
        }
        const unsigned below = lane ? (0xffffffffu >> (32 - lane)) : 0u;
        const unsigned inv = ~F & below;
        const int run = inv ? (lane - 1 - (31 - __clz((int)inv))) : (lane + run_in);
        const bool evaluated = valid && !(run & 1);
        bool alive = evaluated && v > 0.f;
        int16_t code = !evaluated ? kMapSkip : (v == 0.f ? kMapLowVar : (v < 0.f ? (int16_t)0 : (int16_t)-100));
        const unsigned em = __ballot_sync(0xffffffffu, evaluated);
        const unsigned lm = __ballot_sync(0xffffffffu, evaluated && v == 0.f);
        unsigned weak = 0;                                       // weak classifiers evaluated by this warp (lane 0's tally)
        {   // stage 0 ran in k_stage0 for every evaluated, non-low-variance window
            const unsigned s0 = __ballot_sync(0xffffffffu, evaluated && v != 0.f);
            weak = __popc(s0) * (unsigned)(c_stage_first_node[1] - c_stage_first_node[0]);
        }
        if (lane == 0) { atomicAdd(&s_eval, __popc(em)); if (lm) atomicAdd(&s_lowvar, __popc(lm)); }
        // ---- dense stages 1 .. D-1 ----
        const int step = L.step, P = L.P;
        const int x = xi * step, y = row * step;
        const int32_t* S = b.S + (size_t)frame * b.int_stride + L.int_off + (size_t)y * P + x;
        const int D = b.dense_stages;
        double sum = 0.0;
        for (int s = 1; s < D; ++s) {
            const unsigned al = __ballot_sync(0xffffffffu, alive);
            if (!al) break;
            weak += __popc(al) * (unsigned)(c_stage_first_node[s + 1] - c_stage_first_node[s]);
            if (alive) {
                sum = dense_stage(c_stage_first_node[s], c_stage_first_node[s + 1], S, P, v);
                const float thr = c_stage_thr[s];
                if (near_thr(sum, thr)) atomicAdd(&s_near, 1u);
                if (sum < (double)thr) { alive = false; code = (int16_t)s; }
            }
        }
        if (D >= b.nstages) {
            if (alive) {
                if (D == 1) sum = dense_stage(c_stage_first_node[0], c_stage_first_node[1], S, P, v);   // one-stage cascade
                emit_raw(b, frame, L, x, y, sum); code = (int16_t)b.nstages;
            }
        } else {
            // ---- ballot/popc compaction into the frame's work queue ----
            const unsigned am = __ballot_sync(0xffffffffu, alive);
            if (am) {
                int base = 0;
                if (lane == 0) base = atomicAdd(&b.qcount[frame], __popc(am));
                base = __shfl_sync(0xffffffffu, base, 0);
                if (alive) {
                    const int slot = base + __popc(am & below);
                    b.queue[(size_t)frame * b.pos_stride + slot] =
                        make_uint2((uint32_t)x | (uint32_t)y << 12 | (uint32_t)it.level << 24, __float_as_uint(v));
                }
            }
        }
        if (lane == 0 && weak) atomicAdd(&s_weak, weak);
        if (b.stage_map && frame == 0 && valid && code != -100) b.stage_map[L.pos_off + pos] = code;
    }
    __syncthreads();
    if (tid == 0) {
        unsigned long long* c = b.counters + (size_t)frame * kCtrPerFrame;
        if (s_eval) atomicAdd(&c[CTR_EVALUATED], (unsigned long long)s_eval);
        if (s_lowvar) atomicAdd(&c[CTR_LOWVAR], (unsigned long long)s_lowvar);
        if (s_near) atomicAdd(&c[CTR_NEAR], (unsigned long long)s_near);
        if (s_weak) atomicAdd(&c[CTR_WEAK], (unsigned long long)s_weak);
    }
}

void launch_dense(const DevBufs& b, const WorkItem* wk, int nblocks, cudaStream_t st) {
    if (nblocks) k_dense<<<dim3(nblocks, b.nframes), 256, 0, st>>>(b, wk);
}

// ------------------------------------------------------------------------------------------------
// Tail: one warp per queued window, lanes over the weak classifiers of a stage.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int d = 16; d; d >>= 1) v += __shfl_xor_sync(0xffffffffu, v, d);
    return v;
}

__global__ void __launch_bounds__(256) k_tail(DevBufs b) {
    extern __shared__ int32_t s_patch[];                       // 8 warps x (win_h+1)*(win_w+1) ints
    const int frame = blockIdx.y, tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int PW = b.win_w + 1, PH = b.win_h + 1, PN = PW * PH;
    int32_t* patch = s_patch + warp * PN;
    const int qn = min(b.qcount[frame], (int)b.pos_stride);
    const uint2* queue = b.queue + (size_t)frame * b.pos_stride;
    const int nstages = b.nstages, D = b.dense_stages;
    unsigned near_cnt = 0;
    unsigned long long weak_cnt = 0;                          // per-lane tally of evaluated nodes
    for (int i = blockIdx.x * 8 + warp; i < qn; i += gridDim.x * 8) {
        const uint2 e = queue[i];
        const int x = e.x & 4095u, y = (e.x >> 12) & 4095u, level = e.x >> 24;
        const float vnf = __uint_as_float(e.y);
        const LevelInfo& L = b.levels[level];
        const int P = L.P;
        const int32_t* S = b.S + (size_t)frame * b.int_stride + L.int_off + (size_t)y * P + x;
        __syncwarp();
        for (int idx = lane; idx < PN; idx += 32) {
            const int r = idx / PW, c = idx - r * PW;
            patch[idx] = S[r * P + c];
        }
        __syncwarp();
        int s = D;
        double sum = 0.0;
        for (; s < nstages; ++s) {
            const int t0 = b.stage_first_tree[s], nt = b.stage_ntrees[s];
            double part = 0.0;
            for (int t = lane; t < nt; t += 32) {
                const int base = b.stumps_only ? (t0 + t) : b.tree_first_node[t0 + t];
                int at = 0;
                double leaf;
                for (;;) {
                    const uint4 A = __ldg(b.nodeA + base + at);
                    const uint2 C = __ldg(b.nodeC + base + at);
                    const float val = node_val(patch, PW, A.x, A.y, A.z, C.x, C.y, vnf);
                    ++weak_cnt;
                    const bool lt = val < __uint_as_float(A.w);
                    const uint4 Bv = __ldg(b.nodeB + base + at);
                    leaf = lt ? __hiloint2double((int)Bv.y, (int)Bv.x) : __hiloint2double((int)Bv.w, (int)Bv.z);
                    const int go = lt ? (C.x & 255u) : ((C.x >> 8) & 255u);
                    if (!go) break;
                    at = go;
                }
                part += leaf;
            }
            sum = warp_sum(part);
            const float thr = b.stage_thr[s];
            if (lane == 0 && near_thr(sum, thr)) ++near_cnt;
            if (sum < (double)thr) break;
        }
        if (lane == 0) {
            if (s == nstages) emit_raw(b, frame, L, x, y, sum);
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
    return cudaSuccess;
}

void launch_tail(const DevBufs& b, int blocks_per_frame, cudaStream_t st) {
    int smem = 8 * (b.win_w + 1) * (b.win_h + 1) * (int)sizeof(int32_t);
    k_tail<<<dim3(blocks_per_frame, b.nframes), 256, smem, st>>>(b);
}

}  // namespace ugd
