// Window scan and cascade stage evaluation for sm_100a.
//
//   k_scan : one block (8 warps) = 32 grid rows of one level of one frame on the grid-step-1 levels (tiles of 128 x 32
//            windows, 128 candidate windows per bank class) or 16 grid rows on the grid-step-2 levels, two-plane
//            (tilted) tiles and large windows (tiles of 64 x 16), walking the rows left to right.  Per tile:
//              1. the patch of the integral image S under the tile is staged in shared memory by TMA (one
//                 cp.async.bulk.tensor box per plane; for grid step 2 the planes are stored with even / odd columns
//                 de-interleaved, so that x-adjacent windows always sit in consecutive banks and both halves are plain
//                 boxes);
//              2. a warp takes its rows two at a time: every lane computes the variance normalisation of its two windows
//                 (same column, two rows) and runs stage 0 on both at once; the oracle's sequential x-skip rule is resolved
//                 in closed form, per row, with one ballot and a run length carried along the row in a register;
//              3. survivors are kept as bit masks, one per BANK CLASS q = (col + SH*row) mod 32 (the bank of the window's
//                 origin in the tile).  A warp-round takes the j-th live window of every class: lane q always works on a
//                 window of class q, so every corner load of a round is free of bank conflicts whatever the survivor
//                 pattern is.  Rounds are dealt to the 8 warps, two rounds (two windows per lane) share every node fetch;
//                 passing windows set their bit in the next stage's masks (shared-memory atomicOr, one address per
//                 lane) -- this is the ballot/popc compaction of the design, done per bank class.  Node data is
//                 broadcast from __constant__ as merged corner lists (6/8/9 LDS per weak classifier and window);
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
#include <cmath>
#include <cstring>
#include <mutex>
#include <vector>

#include "kernels.cuh"

#ifndef UGD_DADD_PRED
#define UGD_DADD_PRED 0                   // 1: conditional stage-sum add written as a predicated DADD -- ptxas turns it into DADD + 2 FSEL (4 instructions); 0: FSET.BF mask + DFMA (3)
#endif

namespace ugd {

__constant__ DNodeL c_dn[2][kMaxDenseNodes];                  // in-tile nodes, one table per tile layout (grid step 1 / 2)
__constant__ int4 c_coef[kMaxDenseNodes];                     // their coefficients as 12 int8 (only nodes outside the four patterns read them)
__constant__ int c_stage_first_node[kMaxDenseStages + 1];   // first in-tile node of stage s (s <= dense_stages)
__constant__ float c_stage_thr[kMaxDenseStages];
__constant__ double c_stage_base[kMaxDenseStages];
__constant__ int4 c_stage_nc_first[kMaxDenseStages][3];      // exact stump evaluator: first node of each evaluator class (4 patterns, then 6 / 8 / 9 / 12 corners), end
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

// Tile height in windows.  The wide (128-column) tiles of the grid-step-1 levels are also 32 rows high: 128 candidate
// windows per bank class instead of 64 even the per-class survivor counts out further (rounds x nodes of the in-tile stages
// -7 % on the oracle's stage maps), every warp gets twice the rounds per stage (fewer warps idle at the end-of-stage
// barriers) and the halo rows of the S patch are shared by twice as many window rows (52 tile rows for 32 instead of 36 for 16).
// (64-row tiles on 16 warps, two blocks per SM, 256 candidates per class: 7.68 ms per 64 frames against 7.19 -- the tile no
// longer fits 16-bit byte offsets, and 16 warps wait at every end-of-stage barrier.)
static int tile_rows(int si, int win_w, int win_h, bool tilted) {
    if (tile_cols(si, win_w, win_h, tilted) != 2 * kTileCols || getenv("UGD_SCAN_R16")) return kTileRows;
    const int tc = (2 * kTileCols - 1) + win_w + 1, tr = (2 * kTileRows - 1) + win_h + 1;
    return (tc + 7) * tr * 4 <= 36 * 1024 ? 2 * kTileRows : kTileRows;
}

ScanGeom make_scan_geom(int win_w, int win_h, bool tilted) {
    ScanGeom g{};
    for (int si = 0; si < 2; ++si) {
        const int step = si + 1;
        g.cols[si] = tile_cols(si, win_w, win_h, tilted);
        g.rows[si] = tile_rows(si, win_w, win_h, tilted);
        g.TC[si] = (g.cols[si] - 1) * step + win_w + 1;
        g.TR[si] = (g.rows[si] - 1) * step + win_h + 1;
        if (si == 0) {
            int tp = g.TC[0];
            while (!good_shift(tp & 31)) ++tp;
            g.TP[0] = tp; g.TCH[0] = 0; g.SH[0] = tp & 31;
            g.OD[0] = 0; g.plane_ints[0] = (g.TR[0] * tp + 31) / 32 * 32;
        } else {
            // grid step 2: even and odd columns live in two sub-planes ([TR][TCH] each, the odd one OD ints behind the even
            // one), so that x-adjacent windows (x = 0, 2, 4, ...) sit in consecutive banks -- and so that each sub-plane is
            // exactly what one TMA box with element stride 2 delivers
            int tch = ((g.TC[1] + 1) / 2 + 3) & ~3;              // a multiple of 4: every sub-plane row is a whole number of 16-byte units
            while (!good_shift((2 * tch) & 31)) tch += 4;
            g.TCH[1] = tch; g.TP[1] = tch; g.SH[1] = (2 * tch) & 31;
            g.OD[1] = (g.TR[1] * tch + 31) / 32 * 32; g.plane_ints[1] = 2 * g.OD[1];
        }
    }
    g.tile_ints = std::max(g.plane_ints[0], g.plane_ints[1]);
    return g;
}

static inline int32_t tile_off(const ScanGeom& g, int si, int x, int y) {     // byte offset
    return 4 * (si ? y * g.TCH[1] + (x & 1) * g.OD[1] + (x >> 1) : y * g.TP[0] + x);
}
int32_t scan_tile_off(const ScanGeom& g, int si, int x, int y) { return tile_off(g, si, x, y); }
// Chunked form (calls with few frames): one scan block per (row band, tile) -- first = row | tile index << 16.
void build_scan_work_chunked(const ScanGeom& g, const std::vector<LevelInfo>& levels, std::vector<WorkItem>& wk) {
    wk.clear();
    for (int l = 0; l < (int)levels.size(); ++l) {
        const int si = levels[l].step - 1;
        if (levels[l].nyg > 0xffff || (levels[l].nxg + g.cols[si] - 1) / g.cols[si] > 0x7fff) { wk.clear(); return; }   // does not fit the packing: no chunked mode
        for (int r = 0; r < levels[l].nyg; r += g.rows[si])
            for (int t = 0; t * g.cols[si] < levels[l].nxg; ++t) wk.push_back({l, r | t << 16});
    }
}
// One scan block per g.rows[step - 1] grid rows of a level (the plan's generic table assumes kScanRows everywhere).
void build_scan_work(const ScanGeom& g, const std::vector<LevelInfo>& levels, std::vector<WorkItem>& wk) {
    wk.clear();
    for (int l = 0; l < (int)levels.size(); ++l)
        for (int r = 0; r < levels[l].nyg; r += g.rows[levels[l].step - 1]) wk.push_back({l, r});
}
// TMA box of a tile: step 1 = (pitch, rows, 1); step 2 = (sub-plane columns, rows, 1), fetched once per sub-plane.
// cuTensorMapEncodeTiled takes at most 256 elements per box dimension.
bool scan_tma_box(const ScanGeom& g, int si, uint32_t (&box)[3], uint32_t (&estride)[3]) {
    box[0] = (uint32_t)(si ? g.TCH[1] : g.TP[0]); box[1] = (uint32_t)g.TR[si]; box[2] = 1;
    estride[0] = 1; estride[1] = 1; estride[2] = 1;
    return box[0] <= 256 && box[1] <= 256 && !getenv("UGD_NO_TMA");
}
// The tilted-integral tile (cascades with tilted features only) sits right behind the S tile, same geometry.
int32_t scan_plane_bytes(const ScanGeom& g, int si) { return 4 * g.plane_ints[si]; }

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
    std::vector<DNode> dn((size_t)std::max(nn, 1), DNode{});
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
            if (mode == 1) d.delta = ((double)h.left_val - (double)h.right_val) * (UGD_DADD_PRED ? 1.0 : 128.0);   // see add_if_lt
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
    ncf.assign((size_t)std::max(dense_stages, 1) * 12, 0);
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
                ncf[s * 12 + k] = n;
                while (n < z && keyed[n - a].first == k) { dn[n] = keyed[n - a].second; ++n; }
            }
            ncf[s * 12 + 8] = n;
        }
    // device form: one 64-byte node per tile layout (offsets, then {thr, meta, delta} as one 16-byte word) + int8 coefficients
    for (int si = 0; si < 2; ++si) {
        img.dn[si].assign(dn.size(), DNodeL{});
        for (size_t n = 0; n < dn.size(); ++n) {
            DNodeL& o = img.dn[si][n];
            memcpy(o.off, dn[n].off[si], sizeof o.off);
            o.thr = dn[n].thr; o.meta = dn[n].meta; o.delta = dn[n].delta;
        }
    }
    img.coef.assign(dn.size(), make_int4(0, 0, 0, 0));
    for (size_t n = 0; n < dn.size(); ++n) {
        int8_t c8[16] = {0};
        for (int j = 0; j < 12; ++j) c8[j] = (int8_t)dn[n].coef[j];
        memcpy(&img.coef[n], c8, 16);
    }
    img.first_node.assign(c.stage_first_node.begin(), c.stage_first_node.begin() + dense_stages + 1);
    img.thr.assign(c.stage_thr.begin(), c.stage_thr.begin() + dense_stages);
    img.geom = g;
    return cudaSuccess;
}

// Sum of the right leaves of every stage: the exact stump evaluators add (left - right) where a node goes left.
void stage_right_sums(const Cascade& c, std::vector<double>& qbase) {
    qbase.assign((size_t)c.nstages(), 0.0);
    for (int st = 0; st < c.nstages(); ++st)
        for (int n = c.stage_first_node[st]; n < c.stage_first_node[st + 1]; ++n) qbase[st] += (double)c.nodes[n].right_val;
}

// Asynchronous on `st`: the image belongs to the detector and outlives the copies.
cudaError_t upload_scan_constants(const ScanConstImage& img, cudaStream_t st) {
    cudaError_t e;
    for (int si = 0; si < 2; ++si)
        if ((e = cudaMemcpyToSymbolAsync(c_dn, img.dn[si].data(), sizeof(DNodeL) * img.dn[si].size(), sizeof(DNodeL) * kMaxDenseNodes * (size_t)si, cudaMemcpyHostToDevice, st))) return e;
    if ((e = cudaMemcpyToSymbolAsync(c_coef, img.coef.data(), sizeof(int4) * img.coef.size(), 0, cudaMemcpyHostToDevice, st))) return e;
    if ((e = cudaMemcpyToSymbolAsync(c_stage_first_node, img.first_node.data(), sizeof(int) * img.first_node.size(), 0, cudaMemcpyHostToDevice, st))) return e;
    if ((e = cudaMemcpyToSymbolAsync(c_stage_thr, img.thr.data(), sizeof(float) * img.thr.size(), 0, cudaMemcpyHostToDevice, st))) return e;
    if ((e = cudaMemcpyToSymbolAsync(c_stage_base, img.base.data(), sizeof(double) * img.base.size(), 0, cudaMemcpyHostToDevice, st))) return e;
    if ((e = cudaMemcpyToSymbolAsync(c_stage_nc_first, img.ncf.data(), sizeof(int) * img.ncf.size(), 0, cudaMemcpyHostToDevice, st))) return e;
    return cudaMemcpyToSymbolAsync(c_geom, &img.geom, sizeof img.geom, 0, cudaMemcpyHostToDevice, st);
}

__device__ __forceinline__ int near_thr(double sum, double thr, double eps) { return fabs(sum - thr) <= eps; }

// ---- in-tile evaluators --------------------------------------------------------------------------------------------
// `t` points at a window's origin inside the shared tile; node data is warp-uniform and comes from the constant bank
// through the uniform datapath (LDCU.128: four offsets, or {thr, meta, delta}, per instruction); the node pointer is
// walked, never recomputed.  K windows per lane share every node fetch.
// -DUGD_CHECKED=1 (python _build.py --checked; tools/r02_checked.sh): every read of the scan tiles is checked against the
// block's dynamic shared-memory window and every read of the squared-integral plane against the level's plane, and a miss
// traps -- the bounds half of what compute-sanitizer's memcheck would say about this kernel (the tool is closed on the
// GPU pool this was developed on).
#ifndef UGD_CHECKED
#define UGD_CHECKED 0
#endif
__device__ __forceinline__ void chk_tile(uint32_t saddr) {
#if UGD_CHECKED
    extern __shared__ __align__(16) unsigned char s_dyn_chk[];
    uint32_t size;
    asm("mov.u32 %0, %%dynamic_smem_size;" : "=r"(size));
    const uint32_t base = (uint32_t)__cvta_generic_to_shared(s_dyn_chk);
    if (saddr < base || saddr + 4 > base + size || (saddr & 3u)) __trap();
#endif
}
__device__ __forceinline__ int tile_at(const char* t, int off) {
    chk_tile((uint32_t)__cvta_generic_to_shared(t + off));
    return *reinterpret_cast<const int*>(t + off);
}
__device__ __forceinline__ int coef_of(const int4& c, int j) {
    const int w = j < 4 ? c.x : j < 8 ? c.y : c.z;
    return (int)(int8_t)(w >> (8 * (j & 3)));
}

// feature response of one node (any corner count) for one window, coefficients from the table
__device__ __forceinline__ int node_resp(const DNodeL& d, const int4 cf, const char* __restrict__ t) {
    const int4* o4 = reinterpret_cast<const int4*>(d.off);
    const int4 o0 = o4[0], o1 = o4[1];
    int a = coef_of(cf, 0) * tile_at(t, o0.x) + coef_of(cf, 1) * tile_at(t, o0.y) + coef_of(cf, 2) * tile_at(t, o0.z);
    int c = coef_of(cf, 3) * tile_at(t, o0.w) + coef_of(cf, 4) * tile_at(t, o1.x) + coef_of(cf, 5) * tile_at(t, o1.y);
    const int nc = d.meta & 255u;
    if (nc > 6) {
        a += coef_of(cf, 6) * tile_at(t, o1.z); c += coef_of(cf, 7) * tile_at(t, o1.w);
        if (nc > 8) {
            const int4 o2 = o4[2];
            a += coef_of(cf, 8) * tile_at(t, o2.x);
            if (nc > 9) { c += coef_of(cf, 9) * tile_at(t, o2.y); a += coef_of(cf, 10) * tile_at(t, o2.z); c += coef_of(cf, 11) * tile_at(t, o2.w); }
        }
    }
    return a + c;
}

// Generic evaluator (any tree shape, plain fp64 sum of leaves in file order), one window per lane.
__device__ __forceinline__ double tile_stage_generic(int s, const DNodeL* __restrict__ dn, const char* __restrict__ t, float vnf) {
    const int n0 = c_stage_first_node[s], n1 = c_stage_first_node[s + 1];
    double sum = 0.0;
    float leaf = 0.f;
    int at = 0, k = 0;
    for (int n = n0; n < n1; ++n) {
        const DNodeL& d = dn[n];
        const uint32_t meta = d.meta;
        const bool lt = __fmul_rn(__int2float_rn(node_resp(d, c_coef[n], t)), vnf) < d.thr;
        if (at == k) {
            const int go = lt ? ((meta >> 8) & 255u) : ((meta >> 16) & 255u);
            leaf = lt ? d.lval : d.rval;
            at = go ? go : 255;
        }
        if (meta >> 24) { sum += (double)leaf; at = 0; k = 0; } else ++k;
    }
    return sum;
}

// sum += delta where val < thr, in two instructions: FSETP + a predicated DADD (rounded once, like the reference's
// `sum += leaf`; the compiler's own choice for the conditional add is DADD + two FSEL behind the FSETP).  delta128 is the
// leaf difference times 128 (kept that way for the mask form below, exact): UGD_DADD_PRED selects the predicated form,
// which needs no zeroed low word for the mask.
__device__ __forceinline__ void add_if_lt(double& sum, float val, float thr, double delta128) {
#if UGD_DADD_PRED
    asm("{\n\t.reg .pred p;\n\tsetp.lt.f32 p, %1, %2;\n\t@p add.rn.f64 %0, %0, %3;\n\t}" : "+d"(sum) : "f"(val), "f"(thr), "d"(delta128));
#else
    // FSET.BF writes 1.0f / 0.0f; those bits read as the HIGH word of a double (low word 0) are 2^-7 / 0.0, so one DFMA
    // with delta * 128 adds delta or nothing
    asm("{\n\t.reg .f32 m;\n\t.reg .b32 z;\n\t.reg .f64 md;\n\tset.lt.f32.f32 m, %1, %2;\n\tmov.b32 z, 0;\n\t"
        "mov.b64 md, {z, m};\n\tfma.rn.f64 %0, md, %3, %0;\n\t}" : "+d"(sum) : "f"(val), "f"(thr), "d"(delta128));
#endif
}
// sum += m * delta128 with m = 2^-7 or 0 given as the high word of a double
__device__ __forceinline__ void add_masked(double& sum, unsigned m_hi, double delta128) {
    asm("{\n\t.reg .b32 z;\n\t.reg .f64 md;\n\tmov.b32 z, 0;\n\tmov.b64 md, {z, %1};\n\tfma.rn.f64 %0, md, %2, %0;\n\t}" : "+d"(sum) : "r"(m_hi), "d"(delta128));
}

// the NC tile offsets of a node, fetched 16 / 8 / 4 bytes at a time
template <int NC>
__device__ __forceinline__ void load_offsets(const DNodeL* d, int (&o)[12]) {
    const int4* o4 = reinterpret_cast<const int4*>(d->off);
#pragma unroll
    for (int q = 0; q < NC / 4; ++q) { const int4 v = o4[q]; o[4 * q] = v.x; o[4 * q + 1] = v.y; o[4 * q + 2] = v.z; o[4 * q + 3] = v.w; }
    if (NC % 4 >= 2) { const int2 v = *reinterpret_cast<const int2*>(d->off + (NC / 4) * 4); o[(NC / 4) * 4] = v.x; o[(NC / 4) * 4 + 1] = v.y; }
    if (NC % 2) o[NC - 1] = d->off[NC - 1];
}

// Exact stump evaluator, nodes [d, e) of one coefficient pattern (PAT >= 0: compile-time coefficients, +-1 terms fold
// into three-input adds) or of one corner count with coefficients from the table (PAT < 0).
template <int PAT, int NC, int K>
__device__ __forceinline__ void stage_run(const DNodeL* __restrict__ d, const DNodeL* __restrict__ e, const int4* __restrict__ cfp,
                                          const char* const (&t)[K], const float (&vnf)[K], double (&sum)[K]) {
#pragma unroll 1
    for (; d != e; ++d) {
        int o[12];
        load_offsets<NC>(d, o);
        int4 cf = make_int4(0, 0, 0, 0);
        if (PAT < 0) { cf = *cfp; ++cfp; }
        int acc[K];
#pragma unroll
        for (int k = 0; k < K; ++k) acc[k] = 0;
#pragma unroll
        for (int j = 0; j < NC; ++j) {                  // corner by corner: the offset is fetched once for the K windows
            const int cj = PAT >= 0 ? pat_coef(PAT < 0 ? 0 : PAT, j) : coef_of(cf, j);
#pragma unroll
            for (int k = 0; k < K; ++k) acc[k] += cj * tile_at(t[k], o[j]);
        }
        const int4 tail = *reinterpret_cast<const int4*>(&d->thr);          // {thr, meta, delta}: one 16-byte fetch
        const float thr = __int_as_float(tail.x);
        const double delta = __hiloint2double(tail.w, tail.z);
#pragma unroll
        for (int k = 0; k < K; ++k) add_if_lt(sum[k], __fmul_rn(__int2float_rn(acc[k]), vnf[k]), thr, delta);
    }
}

template <int K>
__device__ __forceinline__ void tile_stage_k(int s, const DNodeL* __restrict__ dn, const char* const (&t)[K], const float (&vnf)[K], double (&sum)[K]) {
    const int4 fa = c_stage_nc_first[s][0], fb = c_stage_nc_first[s][1];
    const int f8 = c_stage_nc_first[s][2].x;
#pragma unroll
    for (int k = 0; k < K; ++k) sum[k] = k == 0 ? c_stage_base[s] : sum[0];
    stage_run<0, 6, K>(dn + fa.x, dn + fa.y, nullptr, t, vnf, sum);
    stage_run<1, 8, K>(dn + fa.y, dn + fa.z, nullptr, t, vnf, sum);
    if (fa.z != fb.x) {                                 // centre-surround patterns: rare
        stage_run<2, 9, K>(dn + fa.z, dn + fa.w, nullptr, t, vnf, sum);
        stage_run<3, 9, K>(dn + fa.w, dn + fb.x, nullptr, t, vnf, sum);
    }
    if (fb.x != f8) {                                   // nodes outside the four patterns: coefficients from the table
        stage_run<-1, 6, K>(dn + fb.x, dn + fb.y, c_coef + fb.x, t, vnf, sum);
        stage_run<-1, 8, K>(dn + fb.y, dn + fb.z, c_coef + fb.y, t, vnf, sum);
        stage_run<-1, 9, K>(dn + fb.z, dn + fb.w, c_coef + fb.z, t, vnf, sum);
        stage_run<-1, 12, K>(dn + fb.w, dn + f8, c_coef + fb.w, t, vnf, sum);
    }
}

// Exact evaluator for two-node trees (scan_mode 2), K windows per lane.  Corner counts differ from node to
// node, so the extra corners sit behind warp-uniform branches (paid once per K windows).
template <int K>
__device__ __forceinline__ void node_resp_k(const DNodeL& d, const int4 cf, const char* const (&t)[K], int (&acc)[K]) {
    const int32_t* o = d.off;
    const int nc = d.meta & 255u;
    const unsigned pat = (d.meta >> 26) & 3u;
#pragma unroll
    for (int k = 0; k < K; ++k) acc[k] = 0;
#define PCORNER(P, j)                                                                               \
    {                                                                                               \
        const int oj = o[j];                                                                        \
        _Pragma("unroll") for (int k = 0; k < K; ++k) acc[k] += pat_coef(P, j) * tile_at(t[k], oj); \
    }
    if (pat == 1) { PCORNER(0, 0) PCORNER(0, 1) PCORNER(0, 2) PCORNER(0, 3) PCORNER(0, 4) PCORNER(0, 5) return; }
    if (pat == 2) { PCORNER(1, 0) PCORNER(1, 1) PCORNER(1, 2) PCORNER(1, 3) PCORNER(1, 4) PCORNER(1, 5) PCORNER(1, 6) PCORNER(1, 7) return; }
#undef PCORNER
#define CORNER(j)                                                                                   \
    {                                                                                               \
        const int oj = o[j], cj = coef_of(cf, j);                                                   \
        _Pragma("unroll") for (int k = 0; k < K; ++k) acc[k] += cj * tile_at(t[k], oj);             \
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

// (Pattern classes were tried here too -- pairs ordered by (root, child) pattern, nine loops: alt2 13.7 ms against 11.7
// for this single loop; its stages hold 3 to 34 trees, too few per class to pay for nine loop set-ups.)
template <int K>
__device__ __forceinline__ void tile_trees_k(int s, const DNodeL* __restrict__ dn, const char* const (&t)[K], const float (&vnf)[K], double (&sum)[K]) {
#pragma unroll
    for (int k = 0; k < K; ++k) sum[k] = c_stage_base[s];
    const int n0 = c_stage_first_node[s], n1 = c_stage_first_node[s + 1];
    const int4* cf = c_coef + n0;
#pragma unroll 1
    for (const DNodeL* root = dn + n0; root != dn + n1; root += 2, cf += 2) {
        const DNodeL& kid = root[1];
        int ar[K], ak[K];
        node_resp_k<K>(*root, cf[0], t, ar);
        node_resp_k<K>(kid, cf[1], t, ak);
        const unsigned flip = (root->meta >> 25) & 1u ? 0x3f800000u : 0u;     // the root's leaf is on its right: taken when NOT lt
        const float thr_r = root->thr, thr_k = kid.thr;
        const double d_r = root->delta, d_k = kid.delta;
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

template <int MODE, int K>
__device__ __forceinline__ void stage_eval(int s, const DNodeL* __restrict__ dn, const char* const (&t)[K], const float (&vnf)[K], double (&sum)[K]) {
    if (MODE == 1) tile_stage_k<K>(s, dn, t, vnf, sum);
    else tile_trees_k<K>(s, dn, t, vnf, sum);
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
    chk_tile(addr);
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

// the same from node words already in registers (the group evaluator fetches a node once for several windows)
template <bool WORDS>
__device__ __forceinline__ int lane_resp_abc(const uint4& A, const uint4& B, uint32_t c8, uint32_t pb2, uint32_t nc) {
    const uint32_t o01 = A.x + pb2, o23 = A.y + pb2, o45 = A.z + pb2;
    int a = sx8(B.z, 0) * lds32(lo16<WORDS>(o01)) + sx8(B.z, 1) * lds32(hi16<WORDS>(o01)) + sx8(B.z, 2) * lds32(lo16<WORDS>(o23));
    int c = sx8(B.z, 3) * lds32(hi16<WORDS>(o23)) + sx8(B.w, 0) * lds32(lo16<WORDS>(o45)) + sx8(B.w, 1) * lds32(hi16<WORDS>(o45));
    if (nc > 6) {
        const uint32_t o67 = A.w + pb2;
        a += sx8(B.w, 2) * lds32(lo16<WORDS>(o67));
        c += sx8(B.w, 3) * lds32(hi16<WORDS>(o67));
        if (nc > 8) {
            const uint32_t o89 = B.x + pb2;
            a += sx8(c8, 0) * lds32(lo16<WORDS>(o89));
            if (nc > 9) {
                const uint32_t oab = B.y + pb2;
                c += sx8(c8, 1) * lds32(hi16<WORDS>(o89));
                a += sx8(c8, 2) * lds32(lo16<WORDS>(oab));
                c += sx8(c8, 3) * lds32(hi16<WORDS>(oab));
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

// ---- tile staging ------------------------------------------------------------------------------------------------
// TMA (cp.async.bulk.tensor): one elected thread asks for the whole patch of an integral plane -- a 3-D box (columns,
// rows, frame) of a per-level tensor map built on the host -- and every thread waits on an mbarrier for the bytes.
// Out-of-range rows / columns arrive as zeros (only windows outside the grid ever read them).  On the grid-step-2 levels
// the integral pass already stores every row as [even columns][odd columns] (TMA cannot stride its innermost dimension),
// so the patch is two plain boxes, one per sub-plane of the tile.  No per-element address arithmetic, no LDGSTS: SASS
// shows UTMALDG.
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned long long* bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect(unsigned long long* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long* bar, uint32_t parity) {
    uint32_t done = 0;
    for (int spin = 0; !done; ++spin) {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(done) : "r"(smem_u32(bar)), "r"(parity) : "memory");
        if (spin > (1 << 24)) __trap();                // a copy that never lands must not hang the device
    }
}
__device__ __forceinline__ void tma_box3(int* dst, const void* tmap, int x, int y, int z, unsigned long long* bar) {
    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                 ::"r"(smem_u32(dst)), "l"(tmap), "r"(x), "r"(y), "r"(z), "r"(smem_u32(bar)) : "memory");
}
// one thread: asks for every plane of the tile whose first window is at level pixel (X0, Y0)
__device__ __forceinline__ void tma_stage(int* tile, const char* maps, int level, int planes, int plane_ints, int si, int OD, uint32_t bytes,
                                          int X0, int Y0, int frame, unsigned long long* bar) {
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");     // the tile's earlier readers (generic proxy) come first
    mbar_expect(bar, bytes);
    for (int pl = 0; pl < planes; ++pl) {
        const char* m = maps + (size_t)((level * 2 + pl) * 2) * 128;
        int* dst = tile + pl * plane_ints;
        // grid step 2: the planes are stored de-interleaved (DevBufs::deint2), map 0 = even columns, map 1 = odd columns
        tma_box3(dst, m, si ? X0 >> 1 : X0, Y0, frame, bar);
        if (si) tma_box3(dst + OD, m + 128, X0 >> 1, Y0, frame, bar);
    }
}

// Fallback (tiles a TMA box cannot describe -- more than 256 columns or rows, i.e. windows wider than 128 pixels -- or a
// driver without cuTensorMapEncodeTiled): plain loads and shared stores with the source index clamped per element.
// Copies the patch of one integral plane under the tile whose first window is (xi0, r0) (grid units).
template <int NW>
__device__ __forceinline__ void stage_tile(int* tile, const int32_t* __restrict__ plane, int P, int sw, int sh, int X0, int Y0,
                                           int TR, int TC, int TP, int OD, int si, int w, int lane) {
    for (int r = w; r < TR; r += NW) {
        const int32_t* src = plane + (size_t)min(Y0 + r, sh) * P;
        int* dst = tile + r * TP;
        for (int c = lane; c < TC; c += 32) dst[si ? (c & 1) * OD + (c >> 1) : c] = __ldg(src + min(X0 + c, sw));
    }
}

// Windows per lane that share a node fetch in the exact evaluators, and resident blocks per SM the register
// budget is cut for.  Measured on B200 (alt, 720p, ms per 64 frames): K=4 / 4 blocks (64 regs) 10.7, K=4 / 5 blocks 11.2,
// K=3 / 5 blocks 11.8, K=2 / 5 blocks 10.2, K=2 / 6 blocks (40 regs) 9.9, K=2 / 8 blocks (32 regs, spills) 10.8,
// K=1 / 6 blocks 11.3: two windows per fetch already buy most of the amortisation, and the 12 extra warps per SM
// fill the issue slots the stage barriers leave empty.
// Warps per block (NW): 16 warps on the same 16 x 128 tile (three blocks per SM, same 48 resident warps) measured 9.5 %
// slower than 8 (9.78 vs 8.93 ms): half as many rounds per warp leaves fewer window pairs, and twice as many warps
// wait at every end-of-stage barrier.
#ifndef UGD_PAIR_MIN_ROWS
#define UGD_PAIR_MIN_ROWS 2                          // rows per warp from which stage 0 takes the rows two at a time (see phase 2 of k_scan)
#endif
#ifndef UGD_SCAN_MINBLOCKS
#define UGD_SCAN_MINBLOCKS 6
#endif
#ifndef UGD_SCAN_MINBLOCKS_TALL
#define UGD_SCAN_MINBLOCKS_TALL 4                       // 32-row tiles: 50 KB of shared memory per block, four blocks per SM
#endif

// The survivors of one bank class: bit rl * (C / 32) + h of the mask is the window in tile row rl, column segment h
// (32 bits for 64-column tiles, 64 bits as two words for 128-column tiles).
template <int NMW>
struct Live {
    unsigned w[NMW];
    __device__ __forceinline__ int count() const {
        int n = 0;
#pragma unroll
        for (int i = 0; i < NMW; ++i) n += __popc(w[i]);
        return n;
    }
    __device__ __forceinline__ void skip(int n) {      // drop the n lowest live windows
#pragma unroll
        for (int i = 0; i < NMW; ++i) {
            const int p = __popc(w[i]);
            if (n >= p) { w[i] = 0; n -= p; }
            else { for (; n > 0; --n) w[i] &= w[i] - 1; }
        }
    }
    __device__ __forceinline__ bool pop(int& bit) {    // take the lowest live window; bit = 0 when none is left
        bool act = false;
        bit = 0;
#pragma unroll
        for (int i = 0; i < NMW; ++i)
            if (!act && w[i]) { bit = __ffs((int)w[i]) - 1 + 32 * i; w[i] &= w[i] - 1; act = true; }
        return act;
    }
};

template <int MODE, bool WORDS, int C, int R = kTileRows, int NW = kScanWarps, bool CHUNK = false>
__global__ void __launch_bounds__(NW * 32, CHUNK ? 4 : MODE ? (R == kTileRows ? UGD_SCAN_MINBLOCKS : UGD_SCAN_MINBLOCKS_TALL) * kScanWarps / NW : 1536 / (NW * 32))
k_scan(DevBufs b, const WorkItem* __restrict__ wk, int dbl) {
    constexpr int SEG = C / 32, LSEG = SEG == 2 ? 1 : 2, NMW = R * C / 1024;   // column segments per row, survivor mask words per class
    static_assert((C == 64 || C == 128) && (R == 16 || R == 32) && R % NW == 0, "32, 64 or 128 windows per bank class");
    extern __shared__ __align__(16) unsigned char s_dyn[];
    // TMA wants its shared destination on 128 bytes: the launch asks for 128 bytes more and the tile buffer starts at the next boundary
    int* const s_tiles = reinterpret_cast<int*>(s_dyn + ((128u - (smem_u32(s_dyn) & 127u)) & 127u));
    __shared__ __align__(8) unsigned long long s_mbar;
    __shared__ float s_vnf[R * C];
    __shared__ unsigned s_alive[3][NMW][32];           // rotating survivor masks, NMW words per bank class
    __shared__ int s_pop;
    __shared__ unsigned int s_ctr[5];                  // evaluated, lowvar, near, weak, queued

    // grid = (frames, work items): consecutive blocks are the SAME row band of different frames, so the long blocks (the wide
    // levels lead the table) of every frame start first and the launch ends on short ones (UGD_SCAN_FRAME_MAJOR: the old order)
    const WorkItem it = wk[b.scan_frame_major ? blockIdx.x : blockIdx.y];
    const int frame = b.scan_frame_major ? blockIdx.y : blockIdx.x, tid = threadIdx.x, lane = tid & 31;
    const int w = __shfl_sync(0xffffffffu, tid >> 5, 0);     // warp index, provably warp-uniform
    const LevelInfo& Lg = b.levels[it.level];
    const int sw = Lg.sw, sh = Lg.sh, P = Lg.P, step = Lg.step, nxg = Lg.nxg, nyg = Lg.nyg, nxg_pad = Lg.nxg_pad;
    const float scale = Lg.scale;
    const int rect_w = Lg.rect_w, rect_h = Lg.rect_h;
    const uint32_t pos_off = Lg.pos_off;
    const int si = step - 1;
    const DNodeL* __restrict__ dn = c_dn[si];
    const int TP = c_geom.TP[si], OD = c_geom.OD[si], TC = c_geom.TC[si], SH = c_geom.SH[si];
    const int TR = (R - 1) * step + b.win_h + 1;
    const int plane_ints = si ? c_geom.plane_ints[1] : (TR * TP + 31) / 32 * 32;
    // chunked mode (calls with few frames, see launch_scan): the block takes ONE tile of its row band, not the whole row
    constexpr bool chunked = CHUNK;                    // (its own instantiations: the backward pass below costs the row-walking variants registers)
    const int r0 = chunked ? (it.first & 0xffff) : it.first;
    const int32_t* S = b.S + (size_t)frame * b.int_stride + Lg.int_off;
    const int32_t* Q = b.SQ + (size_t)frame * b.int_stride + Lg.int_off;
    const int ow = b.win_w, oh = b.win_h, D = b.dense_stages, nstages = b.nstages;
    const double area = (double)((ow - 2) * (oh - 2));
    const unsigned below = lane ? (0xffffffffu >> (32 - lane)) : 0u;
    int16_t* map = (b.stage_map && frame == 0) ? b.stage_map + pos_off : nullptr;
    if (tid < 5) s_ctr[tid] = 0;
#if UGD_CHECKED
    // self-test hook of the checked build (tools/r02_checked.sh): UGD_SCAN_KMIN=12345 asks for one read a megabyte past the tile
    if (b.scan_kmin == 12345 && blockIdx.x == 0 && blockIdx.y == 0 && tid == 0) (void)tile_at(reinterpret_cast<const char*>(s_tiles), 1 << 20);
#endif
    unsigned run_par = 0;                              // per owned row: parity of the stage-0 failure run ending just before the next segment
    // per-lane tallies, reduced once at the end of the block: evaluated | low variance << 16,  near a threshold | passed the
    // variance test << 16 (a lane sees at most a few hundred windows of a block: 16 bits each are plenty)
    unsigned cA = 0, cB = 0;
    unsigned weak = 0;                                 // lane 0 of each warp: weak classifiers of stages >= 1
    const int nw0 = c_stage_first_node[1] - c_stage_first_node[0];
    const float thr0 = c_stage_thr[0];
    const double thr0d = (double)thr0, eps0 = 1e-5 * fabs((double)thr0);

    // The tile buffer is double (dbl = 1) when shared memory allows: the next tile's patch is then copied
    // asynchronously while this tile is being evaluated, and its latency disappears behind the stages.
    const int planes = b.T ? 2 : 1, buf_ints = planes * plane_ints;
    const int32_t* Tp = b.T ? b.T + (size_t)frame * b.int_stride + Lg.int_off : nullptr;
    const char* tmaps = reinterpret_cast<const char*>(b.tmaps);     // null: this tile shape has no TMA box (LDGSTS fallback)
    const uint32_t tma_bytes = (uint32_t)(planes * TR * (si ? 2 * TP : TP) * 4);
    uint32_t mphase = 0;
    if (tmaps && tid == 0) mbar_init(&s_mbar, 1);
    int* s_tile = s_tiles;
    __syncthreads();
    const int x_begin = chunked ? (it.first >> 16) * C : 0, x_end = chunked ? min(nxg, x_begin + C) : nxg;
    if (chunked && x_begin > 0) {
        // The x-skip rule needs, per row, the parity of the run of stage-0 failures that ends just left of the chunk.  Walk
        // back 32 positions at a time (their tile is staged like any other; only stage 0 runs, nothing is recorded) until every
        // row has met a position that does not fail: 32 failures in a row leave the parity as it is.
        constexpr int RPW0 = R / NW;
        unsigned unresolved = (1u << RPW0) - 1u;
        for (int xb = x_begin - 32;; xb -= 32) {
            __syncthreads();                           // the tile buffer is free
            if (tmaps) {
                if (tid == 0) tma_stage(s_tile, tmaps, it.level, planes, plane_ints, si, OD, tma_bytes, xb * step, r0 * step, frame, &s_mbar);
                mbar_wait(&s_mbar, mphase);
                mphase ^= 1u;
            } else {
                stage_tile<NW>(s_tile, S, P, sw, sh, xb * step, r0 * step, TR, TC, TP, OD, si, w, lane);
                if (Tp) stage_tile<NW>(s_tile + plane_ints, Tp, P, sw, sh, xb * step, r0 * step, TR, TC, TP, OD, si, w, lane);
            }
            __syncthreads();
            const bool dq = b.deint2 && si;
            const int half = P >> 1, dxb = ow - 1;
            const int oa = dq ? half : 1, ob = dq ? (dxb & 1) * half + (dxb >> 1) : dxb;
#pragma unroll 1
            for (int k = 0; k < RPW0; ++k) {
                if (!(unresolved >> k & 1)) continue;  // warp-uniform
                const int rl = w + k * NW, row = r0 + rl;
                const bool row_valid = row < nyg;
                const int y = row_valid ? row * step : 0;
                const int x = (xb + lane) * step, xq = dq ? x >> 1 : x;
                const char* t = reinterpret_cast<const char*>(s_tile + (rl * step) * TP) + 4 * lane;
                const int32_t* no = c_geom.norm_off[si];
                const int vs = tile_at(t, no[0]) - tile_at(t, no[1]) - tile_at(t, no[2]) + tile_at(t, no[3]);
                const uint32_t vq = (uint32_t)Q[(y + 1) * P + oa + xq] - (uint32_t)Q[(y + 1) * P + ob + xq] -
                                    (uint32_t)Q[(y + oh - 1) * P + oa + xq] + (uint32_t)Q[(y + oh - 1) * P + ob + xq];
                float vnf = 1.f;
                bool ok = false;
                double nf = area * (double)vq - (double)vs * (double)vs;
                if (nf > 0.0) {
                    nf = sqrt(nf);
                    vnf = (float)(1.0 / nf);
                    ok = area * (double)vnf < 1e-1;
                }
                if (!ok) vnf = 1.f;
                double sum;
                if (MODE) {
                    const char* t1[1] = {t}; const float v1[1] = {vnf}; double s1[1];
                    stage_eval<MODE ? MODE : 1, 1>(0, dn, t1, v1, s1);
                    sum = s1[0];
                } else sum = tile_stage_generic(0, dn, t, vnf);
                const unsigned F = __ballot_sync(0xffffffffu, row_valid && ok && sum < thr0d);
                if (F != 0xffffffffu || xb == 0) {
                    if (F != 0xffffffffu) run_par ^= (unsigned)(__clz((int)~F) & 1) << k;
                    unresolved &= ~(1u << k);
                }
            }
            if (!__syncthreads_or(unresolved != 0u)) break;
        }
        __syncthreads();
    }
    if (dbl) {
        if (tmaps) { if (tid == 0) tma_stage(s_tile, tmaps, it.level, planes, plane_ints, si, OD, tma_bytes, x_begin * step, r0 * step, frame, &s_mbar); }
        else {
            stage_tile<NW>(s_tile, S, P, sw, sh, x_begin * step, r0 * step, TR, TC, TP, OD, si, w, lane);
            if (Tp) stage_tile<NW>(s_tile + plane_ints, Tp, P, sw, sh, x_begin * step, r0 * step, TR, TC, TP, OD, si, w, lane);
        }
    }
    for (int xi0 = x_begin; xi0 < x_end; xi0 += C) {
        __syncthreads();                               // previous tile fully consumed
        for (int i = tid; i < 96 * NMW; i += NW * 32) (&s_alive[0][0][0])[i] = 0;
        if (tid == 0) s_pop = 0;                       // the late-stage queue's class counter (phase 4)
        // ---- 1. the S tile (and, for cascades with tilted features, the T tile behind it) ----
        if (tmaps) {
            if (!dbl && tid == 0) tma_stage(s_tile, tmaps, it.level, planes, plane_ints, si, OD, tma_bytes, xi0 * step, r0 * step, frame, &s_mbar);
            mbar_wait(&s_mbar, mphase);
            mphase ^= 1u;
        } else {
            if (!dbl) {
                stage_tile<NW>(s_tile, S, P, sw, sh, xi0 * step, r0 * step, TR, TC, TP, OD, si, w, lane);
                if (Tp) stage_tile<NW>(s_tile + plane_ints, Tp, P, sw, sh, xi0 * step, r0 * step, TR, TC, TP, OD, si, w, lane);
            }
        }
        __syncthreads();
        if (dbl && xi0 + C < x_end) {                  // prefetch the next tile into the other buffer
            int* nxt = s_tile == s_tiles ? s_tiles + buf_ints : s_tiles;
            if (tmaps) { if (tid == 0) tma_stage(nxt, tmaps, it.level, planes, plane_ints, si, OD, tma_bytes, (xi0 + C) * step, r0 * step, frame, &s_mbar); }
            else {
                stage_tile<NW>(nxt, S, P, sw, sh, (xi0 + C) * step, r0 * step, TR, TC, TP, OD, si, w, lane);
                if (Tp) stage_tile<NW>(nxt + plane_ints, Tp, P, sw, sh, (xi0 + C) * step, r0 * step, TR, TC, TP, OD, si, w, lane);
            }
        }
        // ---- 2. variance normalisation + stage 0 + x-skip: a warp owns R / NW rows of the tile and walks each in
        //         segments of 32 windows; everything that depends on the row only is warp-uniform.  With four rows per
        //         warp (32-row tiles) the rows are taken two at a time: a lane holds the windows (row a, column c) and
        //         (row b, column c), and stage 0 runs on both at once -- they share every node fetch, like the pairs of
        //         the later stages ----
        constexpr int RPW = R / NW, KP = (MODE != 0 && RPW % 2 == 0 && RPW >= UGD_PAIR_MIN_ROWS) ? 2 : 1;
#pragma unroll 1
        for (int k = 0; k < RPW; k += KP) {
            int rl[KP], row[KP], y[KP], run_in[KP];
            bool row_valid[KP];
            const int32_t *qa[KP], *qb[KP], *qc[KP], *qd[KP];
            const char* trow[KP];
            // the four corners of the squared-integral sum over the window's inner (ow-2) x (oh-2) box, as four row pointers
            // (de-interleaved rows, grid step 2: column x + dx of an even x lives at half (dx & 1), index x / 2 + dx / 2, so
            // adjacent windows read adjacent words)
            const bool dq = b.deint2 && si;
            const int half = P >> 1, dxb = ow - 1;
            const int oa = dq ? half : 1, ob = dq ? (dxb & 1) * half + (dxb >> 1) : dxb;
#pragma unroll
            for (int j = 0; j < KP; ++j) {
                rl[j] = w + (k + j) * NW; row[j] = r0 + rl[j];
                row_valid[j] = row[j] < nyg;
                y[j] = row_valid[j] ? row[j] * step : 0;
                qa[j] = Q + ((y[j] + 1) * P + oa);
                qb[j] = Q + ((y[j] + 1) * P + ob);
                qc[j] = Q + ((y[j] + oh - 1) * P + oa);
                qd[j] = Q + ((y[j] + oh - 1) * P + ob);
                trow[j] = reinterpret_cast<const char*>(s_tile + (rl[j] * step) * TP);
                run_in[j] = (run_par >> (k + j)) & 1;  // only the parity of the run length matters
            }
#pragma unroll 1
            for (int h = 0; h < SEG; ++h) {
                if (xi0 + 32 * h >= nxg) break;        // warp-uniform: segment past the end of the row
                const int c = 32 * h + lane, xi = xi0 + c;
                const bool in_row = xi < nxg;
                const int x = in_row ? xi * step : 0;
                const int xq = dq ? x >> 1 : x;          // index along the SQ row pointers
                const char* t[KP];
                float vnf[KP];
                bool ok[KP];
#pragma unroll
                for (int j = 0; j < KP; ++j) {
                    t[j] = trow[j] + 4 * c;
                    vnf[j] = 1.f;
                    ok[j] = false;
                    const int32_t* no = c_geom.norm_off[si];
                    const int vs = tile_at(t[j], no[0]) - tile_at(t[j], no[1]) - tile_at(t[j], no[2]) + tile_at(t[j], no[3]);
#if UGD_CHECKED
                    {
                        const long long plane = (long long)(sh + 1) * P;
                        const long long ia = qa[j] + xq - Q, ib = qb[j] + xq - Q, ic = qc[j] + xq - Q, id = qd[j] + xq - Q;
                        if (ia < 0 || ib < 0 || ic < 0 || id < 0 || ia >= plane || ib >= plane || ic >= plane || id >= plane) __trap();
                    }
#endif
                    const uint32_t vq = (uint32_t)qa[j][xq] - (uint32_t)qb[j][xq] - (uint32_t)qc[j][xq] + (uint32_t)qd[j][xq];
                    double nf = area * (double)vq - (double)vs * (double)vs;
                    if (nf > 0.0) {
                        nf = sqrt(nf);
                        vnf[j] = (float)(1.0 / nf);
                        ok[j] = area * (double)vnf[j] < 1e-1;
                    }
                    if (!ok[j]) vnf[j] = 1.f;
                }
                double sum[KP];
                if (MODE) stage_eval<MODE ? MODE : 1, KP>(0, dn, t, vnf, sum);
                else sum[0] = tile_stage_generic(0, dn, t[0], vnf[0]);
#pragma unroll
                for (int j = 0; j < KP; ++j) {
                    const bool valid = row_valid[j] && in_row;
                    const bool pass0 = !(sum[j] < thr0d);
                    // x-skip rule in closed form: position i is evaluated iff the run of stage-0 failures
                    // ending at i-1 has even length (a failure is only seen at evaluated positions)
                    const bool f0 = valid && ok[j] && !pass0;
                    const unsigned F = __ballot_sync(0xffffffffu, f0);
                    const unsigned inv = ~F & below;
                    const int run = inv ? (lane - 1 - (31 - __clz((int)inv))) : (lane + run_in[j]);
                    run_in[j] = ((F == 0xffffffffu) ? run_in[j] + 32 : __clz((int)~F)) & 1;
                    const bool evaluated = valid && !(run & 1);
                    const bool varok = evaluated && ok[j];
                    const bool alive = varok && pass0;
                    cA += (evaluated ? 1u : 0u) + ((evaluated && !ok[j]) ? 0x10000u : 0u);
                    cB += ((varok && near_thr(sum[j], thr0d, eps0)) ? 1u : 0u) + (varok ? 0x10000u : 0u);
                    if (map) {
                        if (valid && !alive) map[row[j] * nxg_pad + xi] = !evaluated ? kMapSkip : !ok[j] ? kMapLowVar : (int16_t)0;
                    }
                    s_vnf[rl[j] * C + c] = vnf[j];
                    if (alive) {
                        if (nstages == 1) {                // one-stage cascade: accepted right here
                            emit_raw(b, frame, scale, rect_w, rect_h, x, y[j], sum[j]);
                            if (map) map[row[j] * nxg_pad + xi] = 1;
                        } else {
                            atomicOr(&s_alive[0][(rl[j] * SEG + h) >> 5][(c + SH * rl[j]) & 31], 1u << ((rl[j] * SEG + h) & 31));
                        }
                    }
                }
            }
#pragma unroll
            for (int j = 0; j < KP; ++j) run_par = (run_par & ~(1u << (k + j))) | ((unsigned)run_in[j] << (k + j));
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
            const float thr = c_stage_thr[s];
            const double thrd = (double)thr, eps = 1e-5 * fabs((double)thr);
            const bool last = s + 1 == nstages;         // block-uniform: the stage that accepts
            if (tid == 0) weak += (unsigned)total * (unsigned)(c_stage_first_node[s + 1] - c_stage_first_node[s]);   // every live window runs the whole stage
            // the rounds are dealt to the warps in runs of `per` (at least scan_kmin, so that late stages with fewer
            // rounds than warps still pair windows up); a warp takes two consecutive rounds at once: the two windows
            // a lane holds share every node fetch
            Live<NMW> mm = m;
            const int per = max((rounds + NW - 1) / NW, b.scan_kmin);
            const int g_beg = min(w * per, rounds), g_end = min(g_beg + per, rounds);
            mm.skip(g_beg);
            for (int g = g_beg; g < g_end; g += MODE ? 2 : 1) {
                const bool two = MODE != 0 && g + 1 < g_end;                   // warp-uniform
                int bit0, bit1 = 0;
                const bool act0 = mm.pop(bit0);
                bool act1 = false;
                if (two) act1 = mm.pop(bit1);
                const int rl0 = bit0 >> LSEG, c0 = ((lane - SH * rl0) & 31) + ((bit0 & (SEG - 1)) << 5);
                const int rl1 = bit1 >> LSEG, c1 = ((lane - SH * rl1) & 31) + ((bit1 & (SEG - 1)) << 5);
                const char* tw0 = reinterpret_cast<const char*>(s_tile + (rl0 * step) * TP + c0);
                const char* tw1 = reinterpret_cast<const char*>(s_tile + (rl1 * step) * TP + c1);
                double sum0, sum1 = 0.0;
                if (MODE == 0) {
                    sum0 = tile_stage_generic(s, dn, tw0, s_vnf[rl0 * C + c0]);
                } else if (two) {
                    const char* t2[2] = {tw0, tw1}; const float v2[2] = {s_vnf[rl0 * C + c0], s_vnf[rl1 * C + c1]}; double s2[2];
                    stage_eval<MODE ? MODE : 1, 2>(s, dn, t2, v2, s2);
                    sum0 = s2[0]; sum1 = s2[1];
                } else {
                    const char* t1[1] = {tw0}; const float v1[1] = {s_vnf[rl0 * C + c0]}; double s1[1];
                    stage_eval<MODE ? MODE : 1, 1>(s, dn, t1, v1, s1);
                    sum0 = s1[0];
                }
                const bool pass0 = act0 && !(sum0 < thrd), pass1 = act1 && !(sum1 < thrd);
                cB += ((act0 && near_thr(sum0, thrd, eps)) ? 1u : 0u) + ((act1 && near_thr(sum1, thrd, eps)) ? 1u : 0u);
                if (map) {
                    if (act0 && !pass0) map[(r0 + rl0) * nxg_pad + xi0 + c0] = (int16_t)s;
                    if (act1 && !pass1) map[(r0 + rl1) * nxg_pad + xi0 + c1] = (int16_t)s;
                }
                if (last) {                             // accepted: nothing left to do for these windows
                    if (pass0) { emit_raw(b, frame, scale, rect_w, rect_h, (xi0 + c0) * step, (r0 + rl0) * step, sum0); if (map) map[(r0 + rl0) * nxg_pad + xi0 + c0] = (int16_t)nstages; }
                    if (pass1) { emit_raw(b, frame, scale, rect_w, rect_h, (xi0 + c1) * step, (r0 + rl1) * step, sum1); if (map) map[(r0 + rl1) * nxg_pad + xi0 + c1] = (int16_t)nstages; }
                } else {
                    if (NMW == 1) {                       // one mask word per class: both windows in one atomic
                        const unsigned pb = (pass0 ? 1u << bit0 : 0u) | (pass1 ? 1u << bit1 : 0u);
                        if (pb) atomicOr(&s_alive[nxt][0][lane], pb);
                    } else {
                        if (pass0) atomicOr(&s_alive[nxt][bit0 >> 5][lane], 1u << (bit0 & 31));
                        if (pass1) atomicOr(&s_alive[nxt][bit1 >> 5][lane], 1u << (bit1 & 31));
                    }
                }
            }
            __syncthreads();
            cur = nxt; ++s;
        }
        // ---- 4. the few survivors of the in-tile stages: block-local work queue.  A warp takes one window at a time from a
        //         shared counter and its 32 lanes evaluate 32 weak classifiers of a stage against the tile in shared memory
        //         (node data from global memory through L1); a warp reduction forms the stage sum ----
        if (total > 0 && s < nstages) {
            const uint4* __restrict__ nA = b.nodeTA[si];
            const uint4* __restrict__ nB = b.nodeTB[si];
            if (MODE == 1) {
            // The survivor list (window ids) lives in the two mask buffers that are dead by now -- everything of s_alive but
            // the block of `cur` -- so it costs no shared memory of its own; it holds CAP entries, and the rare tile with
            // more survivors than that (a tile full of faces) is drained in passes of CAP.
            uint16_t* const lst = reinterpret_cast<uint16_t*>(&s_alive[0][0][0]);
            constexpr int BLK = NMW * 64, CAP = 2 * BLK;                       // uint16 entries per mask buffer / in the list
            const int lskip = cur * BLK;                                       // list slot e sits at e, or behind cur's block
            for (int lo = 0; lo < total; lo += CAP) {
            const int hi = min(total, lo + CAP);
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
                    if (pos >= lo && pos < hi) {
                        const int rl = bit >> LSEG, c = ((lane - SH * rl) & 31) + ((bit & (SEG - 1)) << 5);
                        const int e = pos - lo;
                        lst[e < lskip ? e : e + BLK] = (uint16_t)(rl * C + c);
                    }
                    ++pos;
                }
                if (lane == 0) { s_pop = lo; if (lo == 0) s_ctr[4] += (unsigned)total; }
            }
            __syncthreads();
                constexpr int KQ = 1;                  // (windows per pop that share the node fetches: 2, 3, 4 measured 3-8 % slower)
                for (;;) {
                    int i0 = 0;
                    if (lane == 0) i0 = atomicAdd(&s_pop, KQ);
                    i0 = __shfl_sync(0xffffffffu, i0, 0);
                    if (i0 >= hi) break;
                    uint32_t pb2[KQ];
                    float vn[KQ];
                    int wid[KQ];
                    unsigned alive = 0;                 // bit k: window k of the group is still in the cascade (warp-uniform)
#pragma unroll
                    for (int k = 0; k < KQ; ++k) {
                        { const int e = min(i0 + k, hi - 1) - lo; wid[k] = lst[e < lskip ? e : e + BLK]; }
                        vn[k] = s_vnf[wid[k]];
                        const uint32_t pbase = (uint32_t)__cvta_generic_to_shared(s_tile + ((wid[k] / C) * step) * TP + (wid[k] & (C - 1)));
                        const uint32_t pw = WORDS ? pbase >> 2 : pbase;
                        pb2[k] = pw | pw << 16;
                        if (i0 + k < hi) alive |= 1u << k;
                    }
                    for (int st = s; st < nstages && alive; ++st) {
                        const int n0 = b.stage_first_node[st], nt = b.stage_ntrees[st];
                        double part[KQ];
#pragma unroll
                        for (int k = 0; k < KQ; ++k) part[k] = 0.0;
                        for (int t = lane; t < nt; t += 32) {
                            const int n = n0 + t;
                            const uint4 A = __ldg(nA + n), B = __ldg(nB + n), Cq = __ldg(b.nodeC + n);
                            // slots to walk (6, 8, 9 or 12) from the coefficient bytes already loaded -- an unused slot has coefficient 0,
                            // a padded one inside the walked range is harmless -- instead of one more load from nodeM
                            const uint32_t nc = (Cq.x >> 8) ? 12u : Cq.x ? 9u : (B.w >> 16) ? 8u : 6u;
                            const float nthr = __uint_as_float(Cq.y);
                            const double delta = __hiloint2double((int)Cq.w, (int)Cq.z);      // (left - right) * 128, see add_if_lt
#pragma unroll
                            for (int k = 0; k < KQ; ++k) {
                                if (!(alive >> k & 1)) continue;                                // warp-uniform
                                const int r = lane_resp_abc<WORDS>(A, B, Cq.x, pb2[k], nc);
                                add_if_lt(part[k], __fmul_rn(__int2float_rn(r), vn[k]), nthr, delta);
                            }
                        }
                        const float thr = b.stage_thr[st];
                        const double thrd = (double)thr, eps = 1e-5 * fabs((double)thr), base = b.qbase[st];
                        if (lane == 0) weak += (unsigned)nt * (unsigned)__popc(alive);
#pragma unroll
                        for (int k = 0; k < KQ; ++k) {
                            if (!(alive >> k & 1)) continue;
                            const double sum = warp_sum(part[k]) + base;
                            if (lane == 0 && near_thr(sum, thrd, eps)) ++cB;
                            const bool rejected = sum < thrd;
                            if (rejected || st + 1 == nstages) {
                                alive &= ~(1u << k);
                                if (lane == 0) {
                                    const int rl = wid[k] / C, c = wid[k] & (C - 1);
                                    if (!rejected) emit_raw(b, frame, scale, rect_w, rect_h, (xi0 + c) * step, (r0 + rl) * step, sum);
                                    if (map) map[(r0 + rl) * nxg_pad + xi0 + c] = (int16_t)(rejected ? st : nstages);
                                }
                            }
                        }
                    }
                }
            if (hi < total) __syncthreads();           // the list is rebuilt for the next pass
            }
            } else {
            // Tree cascades: no list and no barrier in front of the queue -- queue entry i is the (i - excl)-th live window
            // of the one class whose range [excl, incl) of the running sum of the per-class survivor counts holds i; every
            // warp finds it in the masks by itself (s_pop was zeroed at the top of the tile)
            if (tid == 0) s_ctr[4] += (unsigned)total;
            auto nth_window = [&](int i) -> int {      // window id (tile row * C + column) of queue entry i, warp-uniform
                Live<NMW> mm;
#pragma unroll
                for (int j = 0; j < NMW; ++j) mm.w[j] = s_alive[cur][j][lane];
                const int cnt = mm.count();
                int incl = cnt;
#pragma unroll
                for (int d = 1; d < 32; d <<= 1) {
                    const int v = __shfl_up_sync(0xffffffffu, incl, d);
                    if (lane >= d) incl += v;
                }
                const unsigned own = __ballot_sync(0xffffffffu, i >= incl - cnt && i < incl);
                const int src = __ffs((int)own) - 1;
                int id = 0;
                if (lane == src) {
                    mm.skip(i - (incl - cnt));
                    int bit;
                    mm.pop(bit);
                    const int rl = bit >> LSEG;
                    id = rl * C + ((lane - SH * rl) & 31) + ((bit & (SEG - 1)) << 5);
                }
                return __shfl_sync(0xffffffffu, id, src);
            };
            for (;;) {
                int i = 0;
                if (lane == 0) i = atomicAdd(&s_pop, 1);
                i = __shfl_sync(0xffffffffu, i, 0);
                if (i >= total) break;
                const int id = nth_window(i), rl = id / C, c = id & (C - 1);
                const float vnf = s_vnf[id];
                const uint32_t pbase = (uint32_t)__cvta_generic_to_shared(s_tile + (rl * step) * TP + c);
                const uint32_t pw = WORDS ? pbase >> 2 : pbase;
                const uint32_t pb2 = pw | pw << 16;
                int st = s;
                double sum = 0.0;
                for (; st < nstages; ++st) {
                    const int nt = b.stage_ntrees[st];
                    double part = 0.0, seq = 0.0;
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
                    if (MODE == 0 && b.seq_sum) {       // (the exact evaluators, MODE 1 and 2, are only chosen for cascades with exact sums)
                        // the stage sums of this cascade are not exactly representable (ugd_cascade_info.sum_exact == 0):
                        // add the leaves one by one in file order, exactly as the oracle's loop does
                        for (int k0 = 0; k0 < nt; k0 += 32) {
                            const float leaf = k0 + lane < nt ? tree_leaf(k0 + lane) : 0.f;
                            const int mcnt = min(32, nt - k0);
                            for (int i2 = 0; i2 < mcnt; ++i2) seq += (double)__shfl_sync(0xffffffffu, leaf, i2);
                        }
                    } else {
                        for (int t = lane; t < nt; t += 32) part += (double)tree_leaf(t);
                    }
                    visited = __reduce_add_sync(0xffffffffu, visited);
                    if (lane == 0) weak += visited;
                    sum = (MODE == 0 && b.seq_sum) ? seq : warp_sum(part);
                    const float thr = b.stage_thr[st];
                    if (lane == 0 && near_thr(sum, (double)thr, 1e-5 * fabs((double)thr))) ++cB;
                    if (sum < (double)thr) break;
                }
                if (lane == 0) {
                    if (st == nstages) emit_raw(b, frame, scale, rect_w, rect_h, (xi0 + c) * step, (r0 + rl) * step, sum);
                    if (map) map[(r0 + rl) * nxg_pad + xi0 + c] = (int16_t)st;
                }
            }
            }
        }
        if (dbl) s_tile = s_tile == s_tiles ? s_tiles + buf_ints : s_tiles;
    }
    // ---- per-lane tallies -> block counters -> per-frame counters ----
    {
        const unsigned ev = __reduce_add_sync(0xffffffffu, cA & 0xffffu), lv = __reduce_add_sync(0xffffffffu, cA >> 16);
        const unsigned nr = __reduce_add_sync(0xffffffffu, cB & 0xffffu), vk = __reduce_add_sync(0xffffffffu, cB >> 16);
        if (lane == 0) {
            if (ev) atomicAdd(&s_ctr[0], ev);
            if (lv) atomicAdd(&s_ctr[1], lv);
            if (nr) atomicAdd(&s_ctr[2], nr);
            weak += vk * (unsigned)nw0;                // stage 0: every window that passed the variance test
            if (weak) atomicAdd(&s_ctr[3], weak);
        }
    }
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
    const size_t one = (size_t)g.plane_ints[si] * sizeof(int) * (tilted ? 2 : 1);
    *dbl = 2 * one + 7 * 1024 <= 37 * 1024 ? 1 : 0;          // 6 blocks x 37 KB < 228 KB
    if (getenv("UGD_SCAN_SINGLE")) *dbl = 0;
    return one * (*dbl ? 2 : 1);                             // (+ the alignment pad of the variant, see scan_align_pad)
}

// The work table is ordered by level and the grid-step-2 levels (scale < 2) come first, so the launch is
// split in two: each half asks only for the shared memory its own tile layout needs.
// TMA wants the tile buffer on a 128-byte boundary; the dynamic region starts right behind the kernel's static shared
// arrays, so the pad is (128 - static size mod 128) mod 128 -- asked for exactly, because the grid-step-2 tile fills its
// sixth of the SM's shared memory to within a few dozen bytes.
template <class K>
static size_t scan_align_pad(K kernel) {
    static const size_t pad = [&] {                    // (one static per kernel variant)
        cudaFuncAttributes a{};
        return cudaFuncGetAttributes(&a, kernel) != cudaSuccess ? (size_t)128 : (128 - a.sharedSizeBytes % 128) % 128;
    }();
    return pad;
}

// With a side stream the two launches run CONCURRENTLY (they touch disjoint levels): the grid-step-1 launch goes first on
// `st`, the grid-step-2 launch next to it on `side`, so the tail of one fills with the blocks of the other.
void launch_scan(const DevBufs& b0, const WorkItem* wk, int nblocks, int nblocks_step2, cudaStream_t st, cudaStream_t side,
                 cudaEvent_t ev_fork, cudaEvent_t ev_join) {
    if (!nblocks) return;
    DevBufs b = b0;
    const ScanGeom g = make_scan_geom(b.win_w, b.win_h, b.T != nullptr);
    const bool both = nblocks_step2 > 0 && nblocks_step2 < nblocks;
    const bool fork = side && both;
    if (fork) { cudaEventRecord(ev_fork, st); cudaStreamWaitEvent(side, ev_fork, 0); }
    for (int pass = 0; pass < 2; ++pass) {
        const int half = fork ? 1 - pass : pass;               // concurrent: the large grid-step-1 launch is issued first
        cudaStream_t launch_st = fork && half == 0 ? side : st;
        const int first = half ? nblocks_step2 : 0, n = half ? nblocks - nblocks_step2 : nblocks_step2;
        if (!n) continue;
        int dbl = 0;
        const size_t smem = scan_smem_bytes(g, half ? 0 : 1, b.T != nullptr, &dbl);
        b.tmaps = half || b0.deint2 ? b0.tmaps : nullptr;     // the grid-step-2 launch has TMA boxes only over de-interleaved rows
        static const bool frame_major = getenv("UGD_SCAN_FRAME_MAJOR") != nullptr || false;
        b.scan_frame_major = (frame_major || n > 65535) ? 1 : 0;
        const dim3 grid = b.scan_frame_major ? dim3(n, b.nframes) : dim3(b.nframes, n);
        const bool wide = g.cols[half ? 0 : 1] == 2 * kTileCols;          // only ever with byte offsets (one plane, small window)
        const bool tall = g.rows[half ? 0 : 1] == 2 * kTileRows;          // only ever with wide tiles
        const int variant = tall ? 9 + b.fast_scan : wide ? 6 + b.fast_scan : b.fast_scan * 2 + (b.word_off ? 1 : 0);
#define UGD_LAUNCH_SCAN(M, W, CC, RR)                                                                                                  \
    do {                                                                                                                               \
        if (b.scan_chunked)                                                                                                            \
            k_scan<M, W, CC, RR, kScanWarps, true><<<grid, kScanWarps * 32, smem + scan_align_pad(k_scan<M, W, CC, RR, kScanWarps, true>), launch_st>>>(b, wk + first, dbl); \
        else                                                                                                                           \
            k_scan<M, W, CC, RR><<<grid, kScanWarps * 32, smem + scan_align_pad(k_scan<M, W, CC, RR>), launch_st>>>(b, wk + first, dbl);  \
    } while (0)
        switch (variant) {
            case 0: UGD_LAUNCH_SCAN(0, false, 64, 16); break;
            case 1: UGD_LAUNCH_SCAN(0, true, 64, 16); break;
            case 2: UGD_LAUNCH_SCAN(1, false, 64, 16); break;
            case 3: UGD_LAUNCH_SCAN(1, true, 64, 16); break;
            case 4: UGD_LAUNCH_SCAN(2, false, 64, 16); break;
            case 5: UGD_LAUNCH_SCAN(2, true, 64, 16); break;
            case 6: UGD_LAUNCH_SCAN(0, false, 128, 16); break;
            case 7: UGD_LAUNCH_SCAN(1, false, 128, 16); break;
            case 8: UGD_LAUNCH_SCAN(2, false, 128, 16); break;
            case 9: UGD_LAUNCH_SCAN(0, false, 128, 32); break;
            case 10: UGD_LAUNCH_SCAN(1, false, 128, 32); break;
            default: UGD_LAUNCH_SCAN(2, false, 128, 32); break;
        }
#undef UGD_LAUNCH_SCAN
    }
    if (fork) { cudaEventRecord(ev_join, side); cudaStreamWaitEvent(st, ev_join, 0); }
}

cudaError_t init_kernel_attributes(int device, int win_w, int win_h, bool has_tilted) {
    const ScanGeom g = make_scan_geom(win_w, win_h, has_tilted);
    int dbl;
    const int scan_smem = (int)std::max(scan_smem_bytes(g, 0, has_tilted, &dbl), scan_smem_bytes(g, 1, has_tilted, &dbl)) + 128;   // + alignment pad
    std::lock_guard<std::mutex> lk(g_attr_mu);
    int& granted = g_scan_smem[device & 63];
    if (scan_smem > granted) {
#define UGD_BOTH(M, W, CC, RR) (const void*)k_scan<M, W, CC, RR>, (const void*)k_scan<M, W, CC, RR, kScanWarps, true>
        const void* variants[24] = {UGD_BOTH(0, false, 64, 16), UGD_BOTH(0, true, 64, 16), UGD_BOTH(1, false, 64, 16), UGD_BOTH(1, true, 64, 16),
                                    UGD_BOTH(2, false, 64, 16), UGD_BOTH(2, true, 64, 16), UGD_BOTH(0, false, 128, 16), UGD_BOTH(1, false, 128, 16),
                                    UGD_BOTH(2, false, 128, 16), UGD_BOTH(0, false, 128, 32), UGD_BOTH(1, false, 128, 32), UGD_BOTH(2, false, 128, 32)};
#undef UGD_BOTH
        for (const void* f : variants) {
            cudaError_t e = cudaFuncSetAttribute(f, cudaFuncAttributeMaxDynamicSharedMemorySize, scan_smem);
            if (e != cudaSuccess) return e;
        }
        granted = scan_smem;
    }
    return cudaSuccess;
}


}  // namespace ugd
