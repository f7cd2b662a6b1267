// Window scan and cascade stage evaluation for sm_100a.
//
//   k_scan : one block (8 warps) = 16 grid rows of one level of one frame, walking the rows left to
//            right in tiles of 64 x 16 windows.  Per tile:
//              1. the patch of the integral image S under the tile is staged in shared memory
//                 (coalesced loads; for grid step 2 even/odd columns are de-interleaved so that
//                 x-adjacent windows always sit in consecutive banks);
//              2. row by row, every lane computes the variance normalisation and stage 0 of its own
//                 window; the oracle's sequential x-skip rule is resolved in closed form with one
//                 ballot and a run length carried along the row in a register;
//              3. survivors are kept as 32 bit masks, one per BANK CLASS q = (col + SH*row) mod 32
//                 (the bank of the window's origin in the tile; each class has exactly 32 candidate
//                 windows).  A warp-round takes the j-th live window of every class: lane q always
//                 works on a window of class q, so every corner load of a round is free of bank
//                 conflicts whatever the survivor pattern is.  Rounds are dealt to the 8 warps;
//                 passing windows set their bit in the next stage's masks (shared-memory atomicOr,
//                 one address per lane) -- this is the ballot/popc compaction of the design, done
//                 per bank class.  Node data is broadcast from __constant__ as merged corner lists
//                 (6/8/9 LDS + IMAD per weak classifier);
//              4. this goes on for the `dense_stages` first stages, or until fewer than `tile_min` windows
//                 are left in the tile;
//              5. the few windows that are left are finished by a block-local work queue: each warp pops
//                 one window at a time and its 32 lanes evaluate 32 weak classifiers of a stage in
//                 parallel against the same shared tile (node data from global memory through L1), a
//                 warp reduction forms the stage sum.
//
// Exactness: feature sums are int32 (exact: weights are small integers), the one rounding step is
// val = f32(float(r) * vnf); stage sums are fp64 sums of fp32 leaves, exact for the shipped cascades
// (ugd_cascade_info.sum_exact), hence independent of summation order.
#include <algorithm>
#include <mutex>
#include <vector>

#include "kernels.cuh"

namespace ugd {

__constant__ DNode c_dn[kMaxDenseNodes];
__constant__ int c_stage_first_node[kMaxDenseStages + 1];   // first in-tile node of stage s (s <= dense_stages)
__constant__ float c_stage_thr[kMaxDenseStages];
__constant__ double c_stage_base[kMaxDenseStages];
__constant__ int c_stage_nc_first[kMaxDenseStages][9];      // exact stump evaluator: first node of each evaluator class (4 patterns, then 6 / 8 / 9 / 12 corners), end
__constant__ ScanGeom c_geom;

// Tile pitches are chosen so that the bank shift between consecutive window rows, (step * pitch) mod 32,
// is one of {8, 12, 20, 24}: a window's bank class q = (col + SH * row) mod 32 then spreads 2-D clusters
// of surviving windows over many classes (tools/sim_compaction3.py).
static bool good_shift(int sh) { return sh == 8 || sh == 12 || sh == 20 || sh == 24; }

// Tile width in windows.  Grid-step-1 levels get 128-column tiles (64 windows per bank class instead of 32: the
// per-class survivor counts even out, tools/sim_stage_util.py) when the wider S tile still leaves room for six
// blocks per SM; grid-step-2 tiles (twice the pixels per window) and two-plane tiles (tilted cascades) stay at 64.
static int tile_cols(int si, int win_w, int win_h, bool tilted) {
    if (si != 0 || tilted || getenv("UGD_SCAN_NARROW")) return kTileCols;
    const int tc = (2 * kTileCols - 1) + win_w + 1, tr = (kTileRows - 1) + win_h + 1;
    return (tc + 7) * tr * 4 <= 25 * 1024 ? 2 * kTileCols : kTileCols;
}

ScanGeom make_scan_geom(int win_w, int win_h, bool tilted) {
    ScanGeom g{};
    for (int si = 0; si < 2; ++si) {
        const int step = si + 1;
        g.cols[si] = tile_cols(si, win_w, win_h, tilted);
        g.TC[si] = (g.cols[si] - 1) * step + win_w + 1;
        g.TR[si] = (kTileRows - 1) * step + win_h + 1;
        if (si == 0) {
            int tp = g.TC[0];
            while (!good_shift(tp & 31)) ++tp;
            g.TP[0] = tp; g.TCH[0] = 0; g.SH[0] = tp & 31;
        } else {
            int tch = (g.TC[1] + 1) / 2;
            while (!good_shift((4 * tch) & 31)) ++tch;
            g.TCH[1] = tch; g.TP[1] = 2 * tch; g.SH[1] = (4 * tch) & 31;
        }
    }
    g.tile_ints = std::max(g.TR[0] * g.TP[0], g.TR[1] * g.TP[1]);
    return g;
}

static inline int32_t tile_off(const ScanGeom& g, int si, int x, int y) {     // byte offset
    return 4 * (si ? y * g.TP[1] + (x & 1) * g.TCH[1] + (x >> 1) : y * g.TP[0] + x);
}
int32_t scan_tile_off(const ScanGeom& g, int si, int x, int y) { return tile_off(g, si, x, y); }
// The tilted-integral tile (cascades with tilted features only) sits right behind the S tile, same geometry.
int32_t scan_plane_bytes(const ScanGeom& g, int si) { return 4 * g.TR[si] * g.TP[si]; }

// 16-bit byte addresses cover the static shared arrays plus one tile of a small window; anything larger
// (two tiles for tilted cascades, wide windows) switches the per-lane tables to word offsets.
bool scan_word_offsets(const Cascade& c) {
    const ScanGeom g = make_scan_geom(c.win_w, c.win_h, c.has_tilted);
    return c.has_tilted || g.tile_ints * 4 > 48 * 1024;
}

// Coefficient patterns of the exact stump evaluator.  With a node's merged corners sorted by coefficient, every node of
// the four frontalface cascades is one of four vectors (two-rect features: P0; three-rect: P1; centre-surround and its
// complement: P2, P3).  A node that matches gets an evaluator with the coefficients as compile-time constants: no
// coefficient fetches, and the +-1 terms fold into three-input adds.  Packed as nibbles, coefficient + 8, corner 0 lowest.
#ifndef UGD_NPAT
#define UGD_NPAT 4
#endif
constexpr int kPatterns = UGD_NPAT;
__host__ __device__ constexpr int pat_nc(int p) { return p == 0 ? 6 : p == 1 ? 8 : 9; }
__host__ __device__ constexpr int pat_coef(int p, int j) {
    const unsigned long long v = p == 0 ? 0xA99776ULL : p == 1 ? 0xBB997755ULL : p == 2 ? 0xAAAA77774ULL : 0xC99996666ULL;
    return (int)((v >> (4 * j)) & 15ULL) - 8;
}

// Which in-tile evaluator a cascade gets: 1 = stumps, 2 = two-node trees (root + one child), both with
// exact "base + deltas" stage sums; 0 = the generic evaluator (any tree shape, plain leaf sums).
int scan_mode(const Cascade& c) {
    if (!c.delta_exact) return 0;
    if (c.max_tree_nodes == 1) return 1;
    for (int t = 0; t < c.ntrees(); ++t) {
        if (c.tree_nnodes[t] != 2) return 0;
        const HostNode& r = c.nodes[c.tree_first_node[t]];
        const HostNode& k = c.nodes[c.tree_first_node[t] + 1];
        const bool one_child = (r.left == 1 && r.right == 0) || (r.left == 0 && r.right == 1);
        if (!one_child || k.left != 0 || k.right != 0) return 0;
    }
    return 2;
}

// Sorts a node's merged corners by coefficient and returns its evaluator class: 0..3 = coefficient pattern,
// 4..7 = no pattern, 6 / 8 / 9 / 12 corners (coefficients from the table), 8 = unsupported corner count.
static int sort_and_classify(DNode& d) {
    const int nc = (int)(d.meta & 255u);
    for (int i = 1; i < nc; ++i)                                   // insertion sort of the corners by coefficient
        for (int j = i; j > 0 && d.coef[j - 1] > d.coef[j]; --j) {
            std::swap(d.coef[j - 1], d.coef[j]); std::swap(d.off[0][j - 1], d.off[0][j]); std::swap(d.off[1][j - 1], d.off[1][j]);
        }
    for (int p = 0; p < kPatterns; ++p) {
        bool same = pat_nc(p) == nc;
        for (int j = 0; same && j < nc; ++j) same = d.coef[j] == pat_coef(p, j);
        if (same) return p;
    }
    return nc == 6 ? 4 : nc == 8 ? 5 : nc == 9 ? 6 : nc == 12 ? 7 : 8;
}

// Host image of everything k_scan reads from __constant__ for one (cascade, dense_stages) pair.  Built once per
// detector (ugd_create); uploaded whenever the detector takes the device's constant bank over (api.cu: bind_constants).
cudaError_t build_scan_constants(const Cascade& c, int dense_stages, ScanConstImage& img) {
    const int nn = c.stage_first_node[dense_stages];
    if (nn > kMaxDenseNodes || dense_stages > kMaxDenseStages) return cudaErrorInvalidValue;
    ScanGeom g = make_scan_geom(c.win_w, c.win_h, c.has_tilted);
    for (int si = 0; si < 2; ++si) {
        g.norm_off[si][0] = tile_off(g, si, 1, 1);
        g.norm_off[si][1] = tile_off(g, si, c.win_w - 1, 1);
        g.norm_off[si][2] = tile_off(g, si, 1, c.win_h - 1);
        g.norm_off[si][3] = tile_off(g, si, c.win_w - 1, c.win_h - 1);
    }
    const int mode = scan_mode(c);
    const bool fast = mode == 1;
    std::vector<DNode>& dn = img.dn;
    dn.assign((size_t)std::max(nn, 1), DNode{});
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
                for (int si = 0; si < 2; ++si) d.off[si][j] = tile_off(g, si, cn.cx[j], cn.cy[j]) + (cn.plane ? scan_plane_bytes(g, si) : 0);
            }
            d.meta = (uint32_t)cn.nc | (uint32_t)h.left << 8 | (uint32_t)h.right << 16 |
                     (k == c.tree_nnodes[t] - 1 ? 1u << 24 : 0u);
            d.thr = h.thr;
            if (mode == 1) d.delta = ((double)h.left_val - (double)h.right_val) * 128.0;   // see add_if_lt
            else if (mode == 2) {
                // tree value = cr + [root takes its leaf](root leaf - cr) + [root takes the child and child goes left](cl - cr)
                const HostNode& kid = c.nodes[base + 1];
                if (k == 0) {
                    const bool leaf_is_left = h.left == 0;
                    d.delta = ((double)(leaf_is_left ? h.left_val : h.right_val) - (double)kid.right_val) * 128.0;   // see add_if_lt
                    d.meta |= leaf_is_left ? 0u : 1u << 25;          // bit 25: the root's leaf is on its RIGHT
                } else d.delta = ((double)h.left_val - (double)h.right_val) * 128.0;
            } else { d.lval = h.left_val; d.rval = h.right_val; }
        }
    }
    std::vector<double>& base = img.base;
    std::vector<int>& ncf = img.ncf;
    base.assign((size_t)std::max(dense_stages, 1), 0.0);
    ncf.assign((size_t)std::max(dense_stages, 1) * 9, 0);
    if (mode == 2) {
        for (int s = 0; s < dense_stages; ++s)
            for (int t = c.stage_first_tree[s]; t < c.stage_first_tree[s] + c.stage_ntrees[s]; ++t)
                base[s] += (double)c.nodes[c.tree_first_node[t] + 1].right_val;
        // corners sorted by coefficient; a node that is pattern 0 or 1 says so in meta bits 26-27 (1 / 2) and the
        // evaluator takes a warp-uniform branch to compile-time coefficients (pairs stay in file order)
        for (int n = 0; n < nn; ++n) {
            const int k = sort_and_classify(dn[n]);
            if (k > 7) return cudaErrorInvalidValue;
            if (k < 2) dn[n].meta |= (uint32_t)(k + 1) << 26;
        }
    }
    if (fast)
        for (int s = 0; s < dense_stages; ++s) {
            const int a = c.stage_first_node[s], z = c.stage_first_node[s + 1];
            for (int n = a; n < z; ++n) base[s] += (double)c.nodes[n].right_val;
            std::vector<std::pair<int, DNode>> keyed;
            for (int n = a; n < z; ++n) { const int k = sort_and_classify(dn[n]); if (k > 7) return cudaErrorInvalidValue; keyed.push_back({k, dn[n]}); }   // a corner count outside {6, 8, 9, 12}
            std::stable_sort(keyed.begin(), keyed.end(), [](const std::pair<int, DNode>& x, const std::pair<int, DNode>& y) { return x.first < y.first; });
            int n = a;
            for (int k = 0; k < 8; ++k) {
                ncf[s * 9 + k] = n;
                while (n < z && keyed[n - a].first == k) { dn[n] = keyed[n - a].second; ++n; }
            }
            ncf[s * 9 + 8] = n;
        }
    img.first_node.assign(c.stage_first_node.begin(), c.stage_first_node.begin() + dense_stages + 1);
    img.thr.assign(c.stage_thr.begin(), c.stage_thr.begin() + dense_stages);
    img.geom = g;
    return cudaSuccess;
}

// Asynchronous on `st`: the image belongs to the detector and outlives the copies.
cudaError_t upload_scan_constants(const ScanConstImage& img, cudaStream_t st) {
    cudaError_t e;
    if ((e = cudaMemcpyToSymbolAsync(c_dn, img.dn.data(), sizeof(DNode) * img.dn.size(), 0, cudaMemcpyHostToDevice, st))) return e;
    if ((e = cudaMemcpyToSymbolAsync(c_stage_first_node, img.first_node.data(), sizeof(int) * img.first_node.size(), 0, cudaMemcpyHostToDevice, st))) return e;
    if ((e = cudaMemcpyToSymbolAsync(c_stage_thr, img.thr.data(), sizeof(float) * img.thr.size(), 0, cudaMemcpyHostToDevice, st))) return e;
    if ((e = cudaMemcpyToSymbolAsync(c_stage_base, img.base.data(), sizeof(double) * img.base.size(), 0, cudaMemcpyHostToDevice, st))) return e;
    if ((e = cudaMemcpyToSymbolAsync(c_stage_nc_first, img.ncf.data(), sizeof(int) * img.ncf.size(), 0, cudaMemcpyHostToDevice, st))) return e;
    return cudaMemcpyToSymbolAsync(c_geom, &img.geom, sizeof img.geom, 0, cudaMemcpyHostToDevice, st);
}

__device__ __forceinline__ int near_thr(double sum, float thr) {
    return fabs(sum - (double)thr) <= 1e-5 * fabs((double)thr);
}

// One in-tile stage for one window.  `t` points at the window's origin inside the shared tile; the
// loop and every node field are warp-uniform, so node data is broadcast from the constant bank.
#define TILE_AT(cf, of) ((cf) * *reinterpret_cast<const int*>(t + (of)))
// feature response r of one node: two independent IMAD chains, extra corners behind uniform branches
__device__ __forceinline__ int node_resp(const DNode& d, int si, const char* __restrict__ t) {
    const int4* o4 = reinterpret_cast<const int4*>(d.off[si]);
    const int4* c4 = reinterpret_cast<const int4*>(d.coef);
    const int4 o0 = o4[0], k0 = c4[0], o1 = o4[1], k1 = c4[1];
    int a = TILE_AT(k0.x, o0.x) + TILE_AT(k0.y, o0.y) + TILE_AT(k0.z, o0.z);
    int c = TILE_AT(k0.w, o0.w) + TILE_AT(k1.x, o1.x) + TILE_AT(k1.y, o1.y);
    const int nc = d.meta & 255u;
    if (nc > 6) {
        a += TILE_AT(k1.z, o1.z); c += TILE_AT(k1.w, o1.w);
        if (nc > 8) {
            const int4 o2 = o4[2], k2 = c4[2];
            a += TILE_AT(k2.x, o2.x);
            if (nc > 9) { c += TILE_AT(k2.y, o2.y); a += TILE_AT(k2.z, o2.z); c += TILE_AT(k2.w, o2.w); }
        }
    }
    return a + c;
}
#undef TILE_AT

// Generic evaluator (any tree shape, plain fp64 sum of leaves in file order), one window per lane.
__device__ __forceinline__ double tile_stage_generic(int s, int si, const char* __restrict__ t, float vnf) {
    const int n0 = c_stage_first_node[s], n1 = c_stage_first_node[s + 1];
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

// Exact stump evaluator for K windows per lane (the same node is applied to the K windows, so its
// constant-bank fetches are paid once per K feature evaluations).  Nodes of a stage are sorted by corner
// count at upload time (sums are exact, hence order-independent), so NC is a compile-time constant.
// sum += delta where val < thr, in two instructions: FSET.BF writes 1.0f / 0.0f, and those bits read as the HIGH word
// of a double (low word 0) are 2^-7 / 0.0, so one DFMA with delta * 128 (stored that way at upload, exact) adds
// delta or nothing -- rounded once, exactly like the DADD it replaces (the compiler's own choice for the
// conditional add is DADD + two FSEL behind an FSETP).
__device__ __forceinline__ void add_if_lt(double& sum, float val, float thr, double delta128) {
    asm("{\n\t.reg .f32 m;\n\t.reg .b32 z;\n\t.reg .f64 md;\n\tset.lt.f32.f32 m, %1, %2;\n\tmov.b32 z, 0;\n\t"
        "mov.b64 md, {z, m};\n\tfma.rn.f64 %0, md, %3, %0;\n\t}" : "+d"(sum) : "f"(val), "f"(thr), "d"(delta128));
}

template <int NC, int K>
__device__ __forceinline__ void stage_nodes(int n0, int n1, int si, const char* const (&t)[K], const float (&vnf)[K], double (&sum)[K]) {
#pragma unroll 1
    for (int n = n0; n < n1; ++n) {
        const DNode& d = c_dn[n];
        // 16-byte fetches (LDCU.128): four offsets / four coefficients / {thr, meta, delta} per instruction
        const int4* o4 = reinterpret_cast<const int4*>(d.off[si]);
        const int4* c4 = reinterpret_cast<const int4*>(d.coef);
        int o[12], cf[12];
#pragma unroll
        for (int q = 0; q < (NC + 3) / 4; ++q) {
            const int4 ov = o4[q], cv = c4[q];
            o[4 * q] = ov.x; o[4 * q + 1] = ov.y; o[4 * q + 2] = ov.z; o[4 * q + 3] = ov.w;
            cf[4 * q] = cv.x; cf[4 * q + 1] = cv.y; cf[4 * q + 2] = cv.z; cf[4 * q + 3] = cv.w;
        }
        int acc[K];
#pragma unroll
        for (int k = 0; k < K; ++k) acc[k] = 0;
#pragma unroll
        for (int j = 0; j < NC; ++j) {                  // corner by corner: offset and coefficient fetched once for the K windows
#pragma unroll
            for (int k = 0; k < K; ++k) acc[k] += cf[j] * *reinterpret_cast<const int*>(t[k] + o[j]);
        }
        const int4 tail = *reinterpret_cast<const int4*>(&d.thr);
        const float thr = __int_as_float(tail.x);
        const double delta = __hiloint2double(tail.w, tail.z);
#pragma unroll
        for (int k = 0; k < K; ++k) add_if_lt(sum[k], __fmul_rn(__int2float_rn(acc[k]), vnf[k]), thr, delta);
    }
}

// The same for nodes whose coefficient vector is pattern PAT: coefficients are compile-time constants.
template <int PAT, int K>
__device__ __forceinline__ void stage_nodes_pat(int n0, int n1, int si, const char* const (&t)[K], const float (&vnf)[K], double (&sum)[K]) {
    constexpr int NC = pat_nc(PAT);
#pragma unroll 1
    for (int n = n0; n < n1; ++n) {
        const DNode& d = c_dn[n];
        const int32_t* o = d.off[si];
        int acc[K];
#pragma unroll
        for (int k = 0; k < K; ++k) acc[k] = 0;
#pragma unroll
        for (int j = 0; j < NC; ++j) {
            const int oj = o[j];
#pragma unroll
            for (int k = 0; k < K; ++k) acc[k] += pat_coef(PAT, j) * *reinterpret_cast<const int*>(t[k] + oj);
        }
        const float thr = d.thr;
        const double delta = d.delta;
#pragma unroll
        for (int k = 0; k < K; ++k) add_if_lt(sum[k], __fmul_rn(__int2float_rn(acc[k]), vnf[k]), thr, delta);
    }
}

template <int K>
__device__ __forceinline__ void tile_stage_k(int s, int si, const char* const (&t)[K], const float (&vnf)[K], double (&sum)[K]) {
    const int* f = c_stage_nc_first[s];
#pragma unroll
    for (int k = 0; k < K; ++k) sum[k] = k == 0 ? c_stage_base[s] : sum[0];
    stage_nodes_pat<0, K>(f[0], f[1], si, t, vnf, sum);
    stage_nodes_pat<1, K>(f[1], f[2], si, t, vnf, sum);
    if (kPatterns > 2) {
        stage_nodes_pat<2, K>(f[2], f[3], si, t, vnf, sum);
        stage_nodes_pat<3, K>(f[3], f[4], si, t, vnf, sum);
    }
    if (f[4] != f[8]) {                                 // nodes outside the four patterns: coefficients from the table
        stage_nodes<6, K>(f[4], f[5], si, t, vnf, sum);
        stage_nodes<8, K>(f[5], f[6], si, t, vnf, sum);
        stage_nodes<9, K>(f[6], f[7], si, t, vnf, sum);
        stage_nodes<12, K>(f[7], f[8], si, t, vnf, sum);
    }
}

// Exact evaluator for two-node trees (scan_mode 2), K windows per lane.  Corner counts differ from node to
// node, so the extra corners sit behind warp-uniform branches (paid once per K windows).
template <int K>
__device__ __forceinline__ void node_resp_k(const DNode& d, int si, const char* const (&t)[K], int (&acc)[K]) {
    const int32_t* o = d.off[si];
    const int nc = d.meta & 255u;
    const unsigned pat = (d.meta >> 26) & 3u;
#pragma unroll
    for (int k = 0; k < K; ++k) acc[k] = 0;
#define PCORNER(P, j)                                                                               \
    {                                                                                               \
        const int oj = o[j];                                                                        \
        _Pragma("unroll") for (int k = 0; k < K; ++k) acc[k] += pat_coef(P, j) * *reinterpret_cast<const int*>(t[k] + oj); \
    }
    if (pat == 1) { PCORNER(0, 0) PCORNER(0, 1) PCORNER(0, 2) PCORNER(0, 3) PCORNER(0, 4) PCORNER(0, 5) return; }
    if (pat == 2) { PCORNER(1, 0) PCORNER(1, 1) PCORNER(1, 2) PCORNER(1, 3) PCORNER(1, 4) PCORNER(1, 5) PCORNER(1, 6) PCORNER(1, 7) return; }
#undef PCORNER
#define CORNER(j)                                                                                   \
    {                                                                                               \
        const int oj = o[j], cj = d.coef[j];                                                        \
        _Pragma("unroll") for (int k = 0; k < K; ++k) acc[k] += cj * *reinterpret_cast<const int*>(t[k] + oj); \
    }
    CORNER(0) CORNER(1) CORNER(2) CORNER(3) CORNER(4) CORNER(5)
    if (nc > 6) {
        CORNER(6) CORNER(7)
        if (nc > 8) {
            CORNER(8)
            if (nc > 9) { CORNER(9) CORNER(10) CORNER(11) }
        }
    }
#undef CORNER
}

// sum += m * delta128 with m = 2^-7 or 0 given as the high word of a double (see add_if_lt)
__device__ __forceinline__ void add_masked(double& sum, unsigned m_hi, double delta128) {
    asm("{\n\t.reg .b32 z;\n\t.reg .f64 md;\n\tmov.b32 z, 0;\n\tmov.b64 md, {z, %1};\n\tfma.rn.f64 %0, md, %2, %0;\n\t}" : "+d"(sum) : "r"(m_hi), "d"(delta128));
}

// (Pattern classes were tried here too -- pairs ordered by (root, child) pattern, nine loops: alt2 13.7 ms against 11.7
// for this single loop; its stages hold 3 to 34 trees, too few per class to pay for nine loop set-ups.)
template <int K>
__device__ __forceinline__ void tree_pairs(int n0, int n1, int si, const char* const (&t)[K], const float (&vnf)[K], double (&sum)[K]) {
#pragma unroll 1
    for (int n = n0; n < n1; n += 2) {
        const DNode& root = c_dn[n];
        const DNode& kid = c_dn[n + 1];
        int ar[K], ak[K];
        node_resp_k<K>(root, si, t, ar);
        node_resp_k<K>(kid, si, t, ak);
        const unsigned flip = (root.meta >> 25) & 1u ? 0x3f800000u : 0u;     // the root's leaf is on its right: taken when NOT lt
        const float thr_r = root.thr, thr_k = kid.thr;
        const double d_r = root.delta, d_k = kid.delta;
#pragma unroll
        for (int k = 0; k < K; ++k) {
            // 1.0f / 0.0f masks (FSET.BF): the tree adds the root's leaf difference when the root takes its leaf,
            // else the child's when the child goes left
            const unsigned lt_r = __float_as_uint(__fmul_rn(__int2float_rn(ar[k]), vnf[k]) < thr_r ? 1.f : 0.f);
            const unsigned lt_k = __float_as_uint(__fmul_rn(__int2float_rn(ak[k]), vnf[k]) < thr_k ? 1.f : 0.f);
            const unsigned takes_leaf = lt_r ^ flip;
            add_masked(sum[k], takes_leaf, d_r);
            add_masked(sum[k], lt_k & ~takes_leaf, d_k);
        }
    }
}

template <int K>
__device__ __forceinline__ void tile_trees_k(int s, int si, const char* const (&t)[K], const float (&vnf)[K], double (&sum)[K]) {
#pragma unroll
    for (int k = 0; k < K; ++k) sum[k] = c_stage_base[s];
    tree_pairs<K>(c_stage_first_node[s], c_stage_first_node[s + 1], si, t, vnf, sum);
}

template <int MODE, int K>
__device__ __forceinline__ void stage_eval(int s, int si, const char* const (&t)[K], const float (&vnf)[K], double (&sum)[K]) {
    if (MODE == 1) tile_stage_k<K>(s, si, t, vnf, sum);
    else tile_trees_k<K>(s, si, t, vnf, sum);
}

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

// ---- per-lane node evaluation against a window in shared memory (lanes = weak classifiers) ----
// Node data: merged corner lists whose BYTE offsets (from the window's origin) were computed at load time
// for each memory layout; two offsets share a 32-bit word, so one add of (base | base << 16) rebases both.
__device__ __forceinline__ int lds32(uint32_t addr) {
    int v;
    asm volatile("ld.shared.b32 %0, [%1];" : "=r"(v) : "r"(addr));
    return v;
}
__device__ __forceinline__ int sx8(uint32_t w, int byte) { return (int)(int8_t)(w >> (8 * byte)); }

// With two tiles (tilted cascades) or a large window the addresses no longer fit 16 bits: the tables then
// hold WORD offsets and pb2 the window's word address (WORDS = true).
template <bool WORDS> __device__ __forceinline__ uint32_t lo16(uint32_t o) { return WORDS ? (o << 2) & 0x3fffcu : o & 0xffffu; }
template <bool WORDS> __device__ __forceinline__ uint32_t hi16(uint32_t o) { return WORDS ? (o >> 14) & 0x3fffcu : o >> 16; }

// response of node n against the window whose shared-memory byte (or word) address is packed in pb2
template <bool WORDS>
__device__ __forceinline__ int lane_resp(const DevBufs& b, const uint4* __restrict__ nA, const uint4* __restrict__ nB, int n, uint32_t pb2, uint32_t nc, uint4& C) {
    const uint4 A = __ldg(nA + n);
    const uint4 B = __ldg(nB + n);
    C = __ldg(b.nodeC + n);
    const uint32_t o01 = A.x + pb2, o23 = A.y + pb2, o45 = A.z + pb2;
    int a = sx8(B.z, 0) * lds32(lo16<WORDS>(o01)) + sx8(B.z, 1) * lds32(hi16<WORDS>(o01)) + sx8(B.z, 2) * lds32(lo16<WORDS>(o23));
    int c = sx8(B.z, 3) * lds32(hi16<WORDS>(o23)) + sx8(B.w, 0) * lds32(lo16<WORDS>(o45)) + sx8(B.w, 1) * lds32(hi16<WORDS>(o45));
    if (nc > 6) {
        const uint32_t o67 = A.w + pb2;
        a += sx8(B.w, 2) * lds32(lo16<WORDS>(o67));
        c += sx8(B.w, 3) * lds32(hi16<WORDS>(o67));
        if (nc > 8) {
            const uint32_t o89 = B.x + pb2;
            a += sx8(C.x, 0) * lds32(lo16<WORDS>(o89));
            if (nc > 9) {
                const uint32_t oab = B.y + pb2;
                c += sx8(C.x, 1) * lds32(hi16<WORDS>(o89));
                a += sx8(C.x, 2) * lds32(lo16<WORDS>(oab));
                c += sx8(C.x, 3) * lds32(hi16<WORDS>(oab));
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

// ---- tile staging: asynchronous global -> shared copies (LDGSTS), 4 bytes per element because the source
// index is clamped per element and the grid-step-2 layout de-interleaves even / odd columns ----
__device__ __forceinline__ void cp_async4(int* dst, const int32_t* src) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"((uint32_t)__cvta_generic_to_shared(dst)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit_wait() {
    asm volatile("cp.async.commit_group;\ncp.async.wait_group 0;" ::: "memory");
}

// Copies the patch of one integral plane under the tile whose first window is (xi0, r0) (grid units).
template <int NW>
__device__ __forceinline__ void stage_tile(int* tile, const int32_t* __restrict__ plane, int P, int sw, int sh, int X0, int Y0,
                                           int TR, int TC, int TP, int TCH, int si, int w, int lane) {
    for (int r = w; r < TR; r += NW) {
        const int32_t* src = plane + (size_t)min(Y0 + r, sh) * P;
        int* dst = tile + r * TP;
        for (int c = lane; c < TC; c += 32) cp_async4(dst + (si ? (c & 1) * TCH + (c >> 1) : c), src + min(X0 + c, sw));
    }
}

// Windows per lane that share a node fetch in the exact evaluators, and resident blocks per SM the register
// budget is cut for.  Measured on B200 (alt, 720p, ms per 64 frames): K=4 / 4 blocks (64 regs) 10.7, K=4 / 5 blocks 11.2,
// K=3 / 5 blocks 11.8, K=2 / 5 blocks 10.2, K=2 / 6 blocks (40 regs) 9.9, K=2 / 8 blocks (32 regs, spills) 10.8,
// K=1 / 6 blocks 11.3: two windows per fetch already buy most of the amortisation (24.5 vs 22 instructions per
// window and node), and the 12 extra warps per SM fill the issue slots the stage barriers leave empty.
// Warps per block (NW): 16 warps on the same 16 x 128 tile (three blocks per SM, same 48 resident warps) measured 9.5 %
// slower than 8 (9.78 vs 8.93 ms): half as many rounds per warp leaves fewer window pairs, and twice as many warps
// wait at every end-of-stage barrier.
#ifndef UGD_SCAN_KMAX
#define UGD_SCAN_KMAX 2
#endif
#ifndef UGD_SCAN_MINBLOCKS
#define UGD_SCAN_MINBLOCKS 6
#endif

// The survivors of one bank class: bit rl * (C / 32) + h of the mask is the window in tile row rl, column segment h
// (32 bits for 64-column tiles, 64 bits as two words for 128-column tiles).
template <int NMW>
struct Live {
    unsigned w[NMW];
    __device__ __forceinline__ int count() const { return NMW == 1 ? __popc(w[0]) : __popc(w[0]) + __popc(w[1]); }
    __device__ __forceinline__ bool any() const { return NMW == 1 ? w[0] != 0 : (w[0] | w[1]) != 0; }
    __device__ __forceinline__ void skip(int n) {      // drop the n lowest live windows
        if (NMW == 2) {
            const int p = __popc(w[0]);
            const bool over = n >= p;
            unsigned v = over ? w[1] : w[0];
            for (int i = over ? n - p : n; i > 0 && v; --i) v &= v - 1;
            if (over) { w[0] = 0; w[1] = v; } else w[0] = v;
        } else {
            for (int i = 0; i < n; ++i) w[0] &= w[0] - 1;
        }
    }
    __device__ __forceinline__ bool pop(int& bit) {    // take the lowest live window; bit = 0 when none is left
        const bool lo = NMW == 1 || w[0] != 0;
        unsigned v = lo ? w[0] : w[NMW - 1];
        const bool act = v != 0;
        bit = act ? __ffs((int)v) - 1 + (lo ? 0 : 32) : 0;
        v &= v - 1;
        if (lo) w[0] = v; else w[NMW - 1] = v;
        return act;
    }
};

template <int MODE, bool WORDS, int C, int NW = kScanWarps>
__global__ void __launch_bounds__(NW * 32, MODE ? UGD_SCAN_MINBLOCKS * kScanWarps / NW : 1536 / (NW * 32)) k_scan(DevBufs b, const WorkItem* __restrict__ wk, int dbl) {
    constexpr int R = kTileRows, KMAX = UGD_SCAN_KMAX;
    constexpr int SEG = C / 32, LSEG = SEG == 2 ? 1 : 2, NMW = R * C / 1024;   // column segments per row, survivor mask words per class
    static_assert((C == 64 || C == 128) && R == 16 && R % NW == 0, "32 or 64 windows per bank class");
    extern __shared__ int s_tiles[];
    __shared__ float s_vnf[R * C];
    __shared__ unsigned s_alive[3][NMW][32];           // rotating survivor masks, NMW words per bank class
    __shared__ uint16_t s_list[R * C];                 // survivors of the in-tile stages (window ids)
    __shared__ int s_pop;
    __shared__ unsigned int s_ctr[5];                  // evaluated, lowvar, near, weak, queued

    const WorkItem it = wk[blockIdx.x];
    const int frame = blockIdx.y, tid = threadIdx.x, lane = tid & 31;
    const int w = __shfl_sync(0xffffffffu, tid >> 5, 0);     // warp index, provably warp-uniform
    const LevelInfo& Lg = b.levels[it.level];
    const int sw = Lg.sw, sh = Lg.sh, P = Lg.P, step = Lg.step, nxg = Lg.nxg, nyg = Lg.nyg, nxg_pad = Lg.nxg_pad;
    const float scale = Lg.scale;
    const int rect_w = Lg.rect_w, rect_h = Lg.rect_h;
    const uint32_t pos_off = Lg.pos_off;
    const int si = step - 1;
    const int TP = c_geom.TP[si], TCH = c_geom.TCH[si], TC = c_geom.TC[si], TR = c_geom.TR[si], SH = c_geom.SH[si];
    const int r0 = it.first;
    const int32_t* S = b.S + (size_t)frame * b.int_stride + Lg.int_off;
    const int32_t* Q = b.SQ + (size_t)frame * b.int_stride + Lg.int_off;
    const int ow = b.win_w, oh = b.win_h, D = b.dense_stages, nstages = b.nstages;
    const double area = (double)((ow - 2) * (oh - 2));
    const unsigned below = lane ? (0xffffffffu >> (32 - lane)) : 0u;
    int16_t* map = (b.stage_map && frame == 0) ? b.stage_map + pos_off : nullptr;
    if (tid < 5) s_ctr[tid] = 0;
    unsigned run_par = 0;                              // per owned row: parity of the stage-0 failure run ending just before the next segment
    unsigned weak = 0;                                 // lane 0 of each warp tallies weak classifiers
    const int nw0 = c_stage_first_node[1] - c_stage_first_node[0];
    const float thr0 = c_stage_thr[0];

    // The tile buffer is double (dbl = 1) when shared memory allows: the next tile's patch is then copied
    // asynchronously while this tile is being evaluated, and its latency disappears behind the stages.
    const int planes = b.T ? 2 : 1, buf_ints = planes * TR * TP;
    const int32_t* Tp = b.T ? b.T + (size_t)frame * b.int_stride + Lg.int_off : nullptr;
    int* s_tile = s_tiles;
    if (dbl) {
        stage_tile<NW>(s_tile, S, P, sw, sh, 0, r0 * step, TR, TC, TP, TCH, si, w, lane);
        if (Tp) stage_tile<NW>(s_tile + TR * TP, Tp, P, sw, sh, 0, r0 * step, TR, TC, TP, TCH, si, w, lane);
    }
    for (int xi0 = 0; xi0 < nxg; xi0 += C) {
        __syncthreads();                               // previous tile fully consumed
        for (int i = tid; i < 96 * NMW; i += NW * 32) (&s_alive[0][0][0])[i] = 0;
        // ---- 1. the S tile (and, for cascades with tilted features, the T tile behind it) ----
        if (!dbl) {
            stage_tile<NW>(s_tile, S, P, sw, sh, xi0 * step, r0 * step, TR, TC, TP, TCH, si, w, lane);
            if (Tp) stage_tile<NW>(s_tile + TR * TP, Tp, P, sw, sh, xi0 * step, r0 * step, TR, TC, TP, TCH, si, w, lane);
        }
        cp_async_commit_wait();
        __syncthreads();
        if (dbl && xi0 + C < nxg) {                    // prefetch the next tile into the other buffer
            int* nxt = s_tile == s_tiles ? s_tiles + buf_ints : s_tiles;
            stage_tile<NW>(nxt, S, P, sw, sh, (xi0 + C) * step, r0 * step, TR, TC, TP, TCH, si, w, lane);
            if (Tp) stage_tile<NW>(nxt + TR * TP, Tp, P, sw, sh, (xi0 + C) * step, r0 * step, TR, TC, TP, TCH, si, w, lane);
        }
        // ---- 2. variance normalisation + stage 0 + x-skip, row segment by row segment ----
#pragma unroll 1
        for (int seg = 0; seg < (R / NW) * SEG; ++seg) {
            const int k = seg >> LSEG, h = seg & (SEG - 1);
            const int rl = w + k * NW, row = r0 + rl;
            const bool row_valid = row < nyg;
            {
                const int c = 32 * h + lane, xi = xi0 + c;
                if (xi0 + 32 * h >= nxg) continue;     // warp-uniform: segment past the end of the row
                const int run_in = (run_par >> k) & 1;  // only the parity of the run length matters
                const bool valid = row_valid && xi < nxg;
                const int x = valid ? xi * step : 0, y = valid ? row * step : 0;
                const char* t = reinterpret_cast<const char*>(s_tile + (rl * step) * TP + c);
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
                double sum;
                if (MODE) {
                    const char* t1[1] = {t}; const float v1[1] = {vnf}; double s1[1];
                    stage_eval<MODE, 1>(0, si, t1, v1, s1);
                    sum = s1[0];
                } else sum = tile_stage_generic(0, si, t, vnf);
                const bool pass0 = !(sum < (double)thr0);
                // x-skip rule in closed form: position i is evaluated iff the run of stage-0 failures
                // ending at i-1 has even length (a failure is only seen at evaluated positions)
                const bool f0 = valid && ok && !pass0;
                const unsigned F = __ballot_sync(0xffffffffu, f0);
                const unsigned inv = ~F & below;
                const int run = inv ? (lane - 1 - (31 - __clz((int)inv))) : (lane + run_in);
                const int run_out = (F == 0xffffffffu) ? run_in + 32 : __clz((int)~F);
                run_par = (run_par & ~(1u << k)) | ((unsigned)(run_out & 1) << k);
                const bool evaluated = valid && !(run & 1);
                const bool alive = evaluated && ok && pass0;
                {
                    const unsigned em = __ballot_sync(0xffffffffu, evaluated);
                    const unsigned lm = __ballot_sync(0xffffffffu, evaluated && !ok);
                    const unsigned s0 = __ballot_sync(0xffffffffu, evaluated && ok);
                    const unsigned nm = __ballot_sync(0xffffffffu, evaluated && ok && near_thr(sum, thr0));
                    if (lane == 0) {
                        atomicAdd(&s_ctr[0], __popc(em));
                        if (lm) atomicAdd(&s_ctr[1], __popc(lm));
                        if (nm) atomicAdd(&s_ctr[2], __popc(nm));
                        weak += __popc(s0) * (unsigned)nw0;
                    }
                }
                if (map && valid) {
                    if (!evaluated) map[row * nxg_pad + xi] = kMapSkip;
                    else if (!ok) map[row * nxg_pad + xi] = kMapLowVar;
                    else if (!pass0) map[row * nxg_pad + xi] = 0;
                }
                s_vnf[rl * C + c] = vnf;
                if (alive) {
                    if (nstages == 1) {                // one-stage cascade: accepted right here
                        emit_raw(b, frame, scale, rect_w, rect_h, x, y, sum);
                        if (map) map[row * nxg_pad + xi] = 1;
                    } else {
                        atomicOr(&s_alive[0][(rl * SEG + h) >> 5][(c + SH * rl) & 31], 1u << ((rl * SEG + h) & 31));
                    }
                }
            }
        }
        __syncthreads();
        // ---- 3. stage by stage on the per-class survivor masks ----
        int cur = 0, s = 1, total;
        Live<NMW> m;
        for (;;) {
#pragma unroll
            for (int i = 0; i < NMW; ++i) m.w[i] = s_alive[cur][i][lane];
            const int cnt = m.count();
            total = __reduce_add_sync(0xffffffffu, cnt);
            if (s >= D || total == 0 || (total < b.tile_min && s >= 2)) break;   // block-uniform
            const int rounds = __reduce_max_sync(0xffffffffu, cnt);
            const int nxt = cur == 2 ? 0 : cur + 1, nxt2 = nxt == 2 ? 0 : nxt + 1;
            if (w == 0)                                 // masks of the stage before: nobody reads them any more
#pragma unroll
                for (int i = 0; i < NMW; ++i) s_alive[nxt2][i][lane] = 0;
            const int nws = c_stage_first_node[s + 1] - c_stage_first_node[s];
            const float thr = c_stage_thr[s];
            if (MODE) {
                // exact evaluators: the rounds are dealt to the warps in runs of `per` (at least scan_kmin, so that
                // late stages with fewer rounds than warps still pair windows up), KMAX consecutive rounds at
                // once: the windows a lane holds share every node fetch
                Live<NMW> mm = m;
                const int per = max((rounds + NW - 1) / NW, b.scan_kmin);
                const int g_beg = min(w * per, rounds), g_end = min(g_beg + per, rounds);
                mm.skip(g_beg);
                for (int g = g_beg; g < g_end; g += KMAX) {
                    const int K = min(KMAX, g_end - g);
                    int bit[4]; bool act[4]; const char* tw[4]; float vn[4]; double sum[4];
#pragma unroll
                    for (int k = 0; k < KMAX; ++k) {
                        act[k] = mm.pop(bit[k]);
                        const int rl = bit[k] >> LSEG, c = ((lane - SH * rl) & 31) + ((bit[k] & (SEG - 1)) << 5);
                        tw[k] = reinterpret_cast<const char*>(s_tile + (rl * step) * TP + c);
                        vn[k] = s_vnf[rl * C + c];
                    }
                    if (K == 1 || KMAX == 1) {
                        const char* t1[1] = {tw[0]}; float v1[1] = {vn[0]}; double s1[1];
                        stage_eval<MODE, 1>(s, si, t1, v1, s1);
                        sum[0] = s1[0];
                    } else if (K == 2 || KMAX == 2) {
                        const char* t2[2] = {tw[0], tw[1]}; float v2[2] = {vn[0], vn[1]}; double s2[2];
                        stage_eval<MODE, 2>(s, si, t2, v2, s2);
                        sum[0] = s2[0]; sum[1] = s2[1];
                    } else if (K == 3 || KMAX == 3) {
                        const char* t3[3] = {tw[0], tw[1], tw[2]}; float v3[3] = {vn[0], vn[1], vn[2]}; double s3[3];
                        stage_eval<MODE, 3>(s, si, t3, v3, s3);
                        sum[0] = s3[0]; sum[1] = s3[1]; sum[2] = s3[2];
                    } else {
                        stage_eval<MODE, 4>(s, si, tw, vn, sum);
                    }
                    unsigned passbits[NMW];
#pragma unroll
                    for (int i = 0; i < NMW; ++i) passbits[i] = 0;
#pragma unroll
                    for (int k = 0; k < KMAX; ++k) {
                        if (k < K) {                    // warp-uniform
                            bool pass = act[k] && !(sum[k] < (double)thr);
                            const unsigned am = __ballot_sync(0xffffffffu, act[k]);
                            const unsigned nm = __ballot_sync(0xffffffffu, act[k] && near_thr(sum[k], thr));
                            if (lane == 0) {
                                weak += (unsigned)__popc(am) * (unsigned)nws;
                                if (nm) atomicAdd(&s_ctr[2], __popc(nm));
                            }
                            const int rl = bit[k] >> LSEG, c = ((lane - SH * rl) & 31) + ((bit[k] & (SEG - 1)) << 5);
                            const int mpos = (r0 + rl) * nxg_pad + xi0 + c;
                            if (act[k] && !pass && map) map[mpos] = (int16_t)s;
                            if (pass && s + 1 == nstages) {
                                emit_raw(b, frame, scale, rect_w, rect_h, (xi0 + c) * step, (r0 + rl) * step, sum[k]);
                                if (map) map[mpos] = (int16_t)nstages;
                                pass = false;           // accepted: nothing left to do for it
                            }
                            if (pass) {
                                if (NMW == 1 || bit[k] < 32) passbits[0] |= 1u << bit[k];
                                else passbits[NMW - 1] |= 1u << (bit[k] - 32);
                            }
                        }
                    }
#pragma unroll
                    for (int i = 0; i < NMW; ++i)
                        if (passbits[i]) atomicOr(&s_alive[nxt][i][lane], passbits[i]);
                }
            } else {
            Live<NMW> mm = m;
            mm.skip(w);                                 // this warp takes rounds w, w+8, ...
            for (int j = w; j < rounds; j += NW) {
                int bit;
                const bool act = mm.pop(bit);
                const int rl = bit >> LSEG, c = ((lane - SH * rl) & 31) + ((bit & (SEG - 1)) << 5);
                const char* tw = reinterpret_cast<const char*>(s_tile + (rl * step) * TP + c);
                const double sum = tile_stage_generic(s, si, tw, s_vnf[rl * C + c]);
                bool pass = act && !(sum < (double)thr);
                const unsigned am = __ballot_sync(0xffffffffu, act);
                const unsigned nm = __ballot_sync(0xffffffffu, act && near_thr(sum, thr));
                if (lane == 0) {
                    weak += (unsigned)__popc(am) * (unsigned)nws;
                    if (nm) atomicAdd(&s_ctr[2], __popc(nm));
                }
                const int mpos = (r0 + rl) * nxg_pad + xi0 + c;
                if (act && !pass && map) map[mpos] = (int16_t)s;
                if (pass && s + 1 == nstages) {
                    emit_raw(b, frame, scale, rect_w, rect_h, (xi0 + c) * step, (r0 + rl) * step, sum);
                    if (map) map[mpos] = (int16_t)nstages;
                    pass = false;                       // accepted: nothing left to do for it
                }
                if (pass) atomicOr(&s_alive[nxt][bit >> 5][lane], 1u << (bit & 31));
                mm.skip(NW - 1);
            }
            }
            __syncthreads();
            cur = nxt; ++s;
        }
        // ---- 4. the few survivors of the in-tile stages: block-local work queue, one window per warp at a
        //         time, the 32 lanes evaluate 32 weak classifiers of a stage against the tile in shared memory ----
        if (total > 0 && s < nstages) {
            if (w == 0) {
                const int cnt = m.count();
                int incl = cnt;
#pragma unroll
                for (int d = 1; d < 32; d <<= 1) {
                    const int v = __shfl_up_sync(0xffffffffu, incl, d);
                    if (lane >= d) incl += v;
                }
                int pos = incl - cnt;
                Live<NMW> mm = m;
                int bit;
                while (mm.pop(bit)) {
                    const int rl = bit >> LSEG, c = ((lane - SH * rl) & 31) + ((bit & (SEG - 1)) << 5);
                    s_list[pos++] = (uint16_t)(rl * C + c);
                }
                if (lane == 0) { s_pop = 0; s_ctr[4] += (unsigned)total; }
            }
            __syncthreads();
            const uint4* __restrict__ nA = b.nodeTA[si];
            const uint4* __restrict__ nB = b.nodeTB[si];
            for (;;) {
                int i = 0;
                if (lane == 0) i = atomicAdd(&s_pop, 1);
                i = __shfl_sync(0xffffffffu, i, 0);
                if (i >= total) break;
                const int id = s_list[i], rl = id / C, c = id & (C - 1);
                const float vnf = s_vnf[id];
                const uint32_t pbase = (uint32_t)__cvta_generic_to_shared(s_tile + (rl * step) * TP + c);
                const uint32_t pw = WORDS ? pbase >> 2 : pbase;
                const uint32_t pb2 = pw | pw << 16;
                int st = s;
                double sum = 0.0;
                for (; st < nstages; ++st) {
                    const int nt = b.stage_ntrees[st];
                    double part = 0.0, seq = 0.0;
                    if (MODE == 1) {
                        const int n0 = b.stage_first_node[st];
                        for (int t = lane; t < nt; t += 32) {
                            const int n = n0 + t;
                            uint4 Cq;
                            const int r = lane_resp<WORDS>(b, nA, nB, n, pb2, __ldg(b.nodeM + n) & 255u, Cq);
                            const bool lt = __fmul_rn(__int2float_rn(r), vnf) < __uint_as_float(Cq.y);
                            part += (double)__uint_as_float(lt ? Cq.z : Cq.w);
                        }
                        if (lane == 0) weak += (unsigned)nt;
                    } else {
                        const int t0 = b.stage_first_tree[st];
                        unsigned visited = 0;
                        // walks tree t for this lane's window; returns the leaf it ends in
                        auto tree_leaf = [&](int t) {
                            const int base = b.tree_first_node[t0 + t];
                            int at = 0;
                            float leaf;
                            for (;;) {
                                const uint32_t M = __ldg(b.nodeM + base + at);
                                uint4 Cq;
                                const int r = lane_resp<WORDS>(b, nA, nB, base + at, pb2, M & 255u, Cq);
                                ++visited;
                                const bool lt = __fmul_rn(__int2float_rn(r), vnf) < __uint_as_float(Cq.y);
                                leaf = __uint_as_float(lt ? Cq.z : Cq.w);
                                const int go = lt ? ((M >> 8) & 255u) : ((M >> 16) & 255u);
                                if (!go) break;
                                at = go;
                            }
                            return leaf;
                        };
                        if (MODE == 0 && b.seq_sum) {   // (the exact evaluators, MODE 1 and 2, are only chosen for cascades with exact sums)
                            // the stage sums of this cascade are not exactly representable (ugd_cascade_info.sum_exact == 0):
                            // add the leaves one by one in file order, exactly as the oracle's loop does
                            for (int c0 = 0; c0 < nt; c0 += 32) {
                                const float leaf = c0 + lane < nt ? tree_leaf(c0 + lane) : 0.f;
                                const int m = min(32, nt - c0);
                                for (int i = 0; i < m; ++i) seq += (double)__shfl_sync(0xffffffffu, leaf, i);
                            }
                        } else {
                            for (int t = lane; t < nt; t += 32) part += (double)tree_leaf(t);
                        }
                        visited = __reduce_add_sync(0xffffffffu, visited);
                        if (lane == 0) weak += visited;
                    }
                    sum = (MODE == 0 && b.seq_sum) ? seq : warp_sum(part);
                    const float thr = b.stage_thr[st];
                    if (lane == 0 && near_thr(sum, thr)) atomicAdd(&s_ctr[2], 1u);
                    if (sum < (double)thr) break;
                }
                if (lane == 0) {
                    if (st == nstages) emit_raw(b, frame, scale, rect_w, rect_h, (xi0 + c) * step, (r0 + rl) * step, sum);
                    if (map) map[(r0 + rl) * nxg_pad + xi0 + c] = (int16_t)st;
                }
            }
        }
        if (dbl) s_tile = s_tile == s_tiles ? s_tiles + buf_ints : s_tiles;
    }
    if (lane == 0 && weak) atomicAdd(&s_ctr[3], weak);
    __syncthreads();
    if (tid < 5 && s_ctr[tid]) {
        constexpr unsigned slots = CTR_EVALUATED | CTR_LOWVAR << 4 | CTR_NEAR << 8 | CTR_WEAK << 12 | CTR_QUEUED << 16;   // s_ctr[i] -> counter
        atomicAdd(&b.counters[(size_t)frame * kCtrPerFrame + ((slots >> (4 * tid)) & 15u)], (unsigned long long)s_ctr[tid]);
    }
}

// cudaFuncSetAttribute applies per device: the limit already granted is tracked per device and only ever raised.
static std::mutex g_attr_mu;
static int g_scan_smem[64] = {0};

// Shared memory of one launch: the tile (x2 planes for tilted cascades), doubled when two buffers still leave
// room for six blocks per SM next to the static arrays (the asynchronous prefetch of the next tile).
static size_t scan_smem_bytes(const ScanGeom& g, int si, bool tilted, int* dbl) {
    const size_t one = (size_t)g.TR[si] * g.TP[si] * sizeof(int) * (tilted ? 2 : 1);
    *dbl = 2 * one + 7 * 1024 <= 37 * 1024 ? 1 : 0;          // 6 blocks x 37 KB < 228 KB
    if (getenv("UGD_SCAN_SINGLE")) *dbl = 0;
    return one * (*dbl ? 2 : 1);
}

// The work table is ordered by level and the grid-step-2 levels (scale < 2) come first, so the launch is
// split in two: each half asks only for the shared memory its own tile layout needs.
void launch_scan(const DevBufs& b, const WorkItem* wk, int nblocks, int nblocks_step2, cudaStream_t st) {
    if (!nblocks) return;
    const ScanGeom g = make_scan_geom(b.win_w, b.win_h, b.T != nullptr);
    for (int half = 0; half < 2; ++half) {
        const int first = half ? nblocks_step2 : 0, n = half ? nblocks - nblocks_step2 : nblocks_step2;
        if (!n) continue;
        int dbl = 0;
        const size_t smem = scan_smem_bytes(g, half ? 0 : 1, b.T != nullptr, &dbl);
        const dim3 grid(n, b.nframes);
        const bool wide = g.cols[half ? 0 : 1] == 2 * kTileCols;          // only ever with byte offsets (one plane, small window)
        const int variant = wide ? 6 + b.fast_scan : b.fast_scan * 2 + (b.word_off ? 1 : 0);
        switch (variant) {
            case 0: k_scan<0, false, 64><<<grid, kScanWarps * 32, smem, st>>>(b, wk + first, dbl); break;
            case 1: k_scan<0, true, 64><<<grid, kScanWarps * 32, smem, st>>>(b, wk + first, dbl); break;
            case 2: k_scan<1, false, 64><<<grid, kScanWarps * 32, smem, st>>>(b, wk + first, dbl); break;
            case 3: k_scan<1, true, 64><<<grid, kScanWarps * 32, smem, st>>>(b, wk + first, dbl); break;
            case 4: k_scan<2, false, 64><<<grid, kScanWarps * 32, smem, st>>>(b, wk + first, dbl); break;
            case 5: k_scan<2, true, 64><<<grid, kScanWarps * 32, smem, st>>>(b, wk + first, dbl); break;
            case 6: k_scan<0, false, 128><<<grid, kScanWarps * 32, smem, st>>>(b, wk + first, dbl); break;
            case 7: k_scan<1, false, 128><<<grid, kScanWarps * 32, smem, st>>>(b, wk + first, dbl); break;
            default: k_scan<2, false, 128><<<grid, kScanWarps * 32, smem, st>>>(b, wk + first, dbl); break;
        }
    }
}

cudaError_t init_kernel_attributes(int device, int win_w, int win_h, bool has_tilted) {
    const ScanGeom g = make_scan_geom(win_w, win_h, has_tilted);
    int dbl;
    const int scan_smem = (int)std::max(scan_smem_bytes(g, 0, has_tilted, &dbl), scan_smem_bytes(g, 1, has_tilted, &dbl));
    std::lock_guard<std::mutex> lk(g_attr_mu);
    int& granted = g_scan_smem[device & 63];
    if (scan_smem > granted) {
        const void* variants[9] = {(const void*)k_scan<0, false, 64>, (const void*)k_scan<0, true, 64>, (const void*)k_scan<1, false, 64>,
                                   (const void*)k_scan<1, true, 64>, (const void*)k_scan<2, false, 64>, (const void*)k_scan<2, true, 64>,
                                   (const void*)k_scan<0, false, 128>, (const void*)k_scan<1, false, 128>, (const void*)k_scan<2, false, 128>};
        for (const void* f : variants) {
            cudaError_t e = cudaFuncSetAttribute(f, cudaFuncAttributeMaxDynamicSharedMemorySize, scan_smem);
            if (e != cudaSuccess) return e;
        }
        granted = scan_smem;
    }
    return cudaSuccess;
}


}  // namespace ugd
