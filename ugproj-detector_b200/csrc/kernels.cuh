// Device-side view of a detector's buffers and the launch wrappers of every kernel on the path.
// All kernels are hand-written for sm_100a; none has a CPU fallback.
#pragma once
#include <cuda_runtime.h>

#include <cstdint>
#include <vector>

#include "../../include/ugdet.h"
#include "cascade.h"
#include "plan.h"

namespace ugd {

constexpr int kMaxDenseNodes = 380;       // nodes of the in-tile stages, held in __constant__ (2 x 64 B + 16 B each)
constexpr int kMaxDenseStages = 48;

// In-tile node: merged corner list.  BYTE offsets into the shared-memory tile are precomputed for the
// two tile layouts (index 0: grid step 1; index 1: grid step 2, even/odd columns de-interleaved).
// Every field is a full 32-bit word so the uniform datapath can fetch it with LDCU.128 and use it
// without unpacking.
struct alignas(16) DNode {
    int32_t off[2][12];
    int32_t coef[12];
    float   thr;
    uint32_t meta;                        // nc | lnode << 8 | rnode << 16 | last << 24
    union {
        struct { float lval, rval; };     // generic evaluator: the two leaves
        double delta;                     // exact evaluators: leaf difference (stumps: times 128, see add_if_lt)
    };
};
static_assert(sizeof(DNode) == 160, "DNode layout");          // host-side working form (both layouts, int32 coefficients)

// Device form of an in-tile node for ONE tile layout: 64 bytes, every group fetched with one LDCU.128 through the
// uniform datapath -- offsets 0-3 / 4-7 / 8-11, then {thr, meta, delta} together.
struct alignas(16) DNodeL {
    int32_t off[12];                      // tile BYTE offsets of the merged corners (sorted by coefficient for pattern nodes)
    float   thr;
    uint32_t meta;                        // nc | lnode << 8 | rnode << 16 | last << 24 | flags (see build_scan_constants)
    union {
        struct { float lval, rval; };     // generic evaluator: the two leaves
        double delta;                     // exact evaluators: leaf difference times 128 (see add_if_lt)
    };
};
static_assert(sizeof(DNodeL) == 64, "DNodeL layout");

constexpr int kTileRows = 16;             // scan tile: kTileRows x kTileCols windows = 32 per bank class
constexpr int kTileCols = 64;             // (grid-step-1 levels: 128 columns, 64 per class, when shared memory allows)
constexpr int kScanWarps = 8;              // warps per scan block
struct ScanGeom {                          // per layout (step 1 / step 2)
    int cols[2];                           // tile width in windows (kTileCols, or twice that: see tile_cols)
    int rows[2];                           // tile height in windows = grid rows per scan block (kTileRows, or twice that: see tile_rows)
    int TP[2], TCH[2], TC[2], TR[2];       // tile row pitch (ints), columns per sub-plane (step 2: == pitch), columns, rows
    int OD[2];                             // step 2: ints from the even-column sub-plane to the odd-column one (0 for step 1)
    int plane_ints[2];                     // ints of one integral plane's tile (multiple of 32: the next plane / buffer stays 128-byte aligned)
    int SH[2];                             // bank shift between consecutive window rows: (step * TP) mod 32
    int32_t norm_off[2][4];                // tile BYTE offsets of the 4 variance-normalisation corners
    int tile_ints;                         // ints of shared memory for the tile
};
constexpr int kCtrPerFrame = 8;           // per-frame device counters (see enum below)
enum { CTR_EVALUATED = 0, CTR_LOWVAR, CTR_QUEUED, CTR_RAW, CTR_GROUPED, CTR_NEAR, CTR_RAW_OVERFLOW, CTR_WEAK };

// Stage-map codes (match the oracle's): -2 skipped, -1 low variance, k rejected at stage k, nstages accepted.
constexpr int16_t kMapSkip = -2, kMapLowVar = -1;

struct DevBufs {
    // geometry
    int w, h, nlevels, nframes;
    int win_w, win_h, nstages, dense_stages, stumps_only;
    int fast_scan;                                    // scan_mode(cascade): 0 generic, 1 exact stumps, 2 exact two-node trees
    const LevelInfo* levels;
    const uint32_t* xtab; const uint32_t* ytab;
    // cascade (per-lane evaluator of the late stages, SoA in global memory; the dense stages live in __constant__)
    // node n: A = 8 x u16 byte offsets from the window origin (corners 0-7); B = {off8|off9<<16, off10|off11<<16, coef0-3, coef4-7};
    //              C = {coef8-11, thr, left leaf, right leaf} -- exact stump cascades (fast_scan == 1): {coef8-11, thr, (left - right) * 128 as a double};
    //              M = nc | lnode<<8 | rnode<<16
    const uint4* nodeC; const uint32_t* nodeM;
    const uint4* nodeTA[2]; const uint4* nodeTB[2];   // A/B with byte offsets into the scan tile (grid step 1 / 2 layouts)
    const double* qbase;                              // exact stump cascades: per stage, the sum of the right leaves (nodeC then holds thr and (left - right) * 128)
    int seq_sum;                                      // stage sums are not exact in fp64: the late stages add their leaves in file order
    int word_off;                                     // the A/B offsets are in 32-bit words (two tiles / large windows: bytes do not fit 16 bits)
    const int32_t* stage_first_tree; const int32_t* stage_ntrees; const float* stage_thr;
    const int32_t* tree_first_node; const int32_t* stage_first_node;
    // frames
    const uint8_t* bgr; size_t bgr_stride;            // input BGR frames (device)
    const uint8_t* area2_src; size_t area2_row_stride, area2_frame_stride;   // non-null: full-size BGR frames, decimated 2x2 (INTER_AREA) inside k_preprocess; bgr is then unused
    uint8_t* gray_raw; uint8_t* gray_eq; uint8_t* mask; size_t px_stride;   // per-frame planes
    const uint8_t* src_gray; size_t src_stride; int src_pitch;             // what the pyramid is built from
    uint32_t* hist;                                   // [nframes][256]
    uint8_t skin_lo[3], skin_hi[3];
    int use_skin;
    // pyramid + integrals
    uint8_t* pyr; size_t pyr_stride;
    int32_t* S; int32_t* SQ; size_t int_stride;
    int deint2;                                       // S / SQ rows of the grid-step-2 levels are stored [even columns][odd columns] (see store_row16, k_scan)
    const void* tmaps;                                // TMA tensor maps of the S (and T) planes of this buffer set: [level][plane][half] x 128 B; null = LDGSTS staging
    int32_t* T;                                       // tilted integral, same layout as S; null unless the cascade has tilted features
    uint32_t* carryS; uint32_t* carryQ; size_t carry_stride;
    // window scan
    size_t pos_stride;                                // padded grid positions per frame
    int scan_kmin;                                    // a warp takes at least this many rounds of a stage at once (they share the node fetches)
    int scan_frame_major;                             // k_scan grid is (work items, frames) instead of (frames, work items)
    int scan_chunked;                                 // the work items are (row band, tile) pairs: first = row | tile << 16 (calls with few frames; see k_scan)
    int tile_min;                                     // hand a tile's survivors to the per-warp evaluator once fewer than this remain
    ugd_rect* raw; double* rawsum; int rawcap; int* rawcount;
    int16_t* stage_map;                               // frame 0 only, padded layout; may be null
    // grouping / output
    int* grp_scratch; size_t grp_stride;              // per frame: 11*rawcap ints (labels, class sums, counts, kept rects, class list)
    ugd_rect* out; int outcap; int* outcount;
    int min_neighbors;
    int no_clip;                                      // OpenCV-2.4 legacy mode: grouped rects are returned as they are
    int gray14;                                       // OpenCV-2.4 legacy mode: BGR -> gray with the 14-bit coefficients
    unsigned long long* counters;                     // [nframes][kCtrPerFrame]
};

// Builds the in-tile node table for stages [0, dense_stages) and uploads it with the tile geometry
// to __constant__ (call before launching k_scan with a new cascade).
ScanGeom make_scan_geom(int win_w, int win_h, bool tilted);
void build_scan_work(const ScanGeom& g, const std::vector<LevelInfo>& levels, std::vector<WorkItem>& wk);
void build_scan_work_chunked(const ScanGeom& g, const std::vector<LevelInfo>& levels, std::vector<WorkItem>& wk);   // block -> (level, first grid row | tile << 16)   // block -> (level, first grid row)
bool scan_tma_box(const ScanGeom& g, int si, uint32_t (&box)[3], uint32_t (&estride)[3]);   // the TMA box of a tile, false if it has none
int32_t scan_tile_off(const ScanGeom& g, int si, int x, int y);   // byte offset of S(x, y) from a window's origin in the tile
int32_t scan_plane_bytes(const ScanGeom& g, int si);              // bytes from the S tile to the tilted-integral tile
bool scan_word_offsets(const Cascade& c);                         // per-lane tables in words instead of bytes
int scan_mode(const Cascade& c);
// Host image of the __constant__ data of one (cascade, dense_stages) pair; built once per detector.
struct ScanConstImage {
    std::vector<DNodeL> dn[2];                        // per tile layout
    std::vector<int4> coef;                           // 12 int8 coefficients per node
    std::vector<int> first_node;                      // [dense_stages + 1]
    std::vector<float> thr;                           // [dense_stages]
    std::vector<double> base;                         // [dense_stages]
    std::vector<int> ncf;                             // [dense_stages][12] (9 used)
    ScanGeom geom;
};
void stage_right_sums(const Cascade& c, std::vector<double>& qbase);   // per stage: sum of the right leaves
cudaError_t build_scan_constants(const Cascade& c, int dense_stages, ScanConstImage& img);
cudaError_t upload_scan_constants(const ScanConstImage& img, cudaStream_t st);   // asynchronous: img must outlive the copies

// ---- INTER_AREA decimation to the scan size (k_area.cu) ----
struct AreaTap { int32_t si; float a; };              // source index, weight
struct AreaPlanHost {
    int sw = 0, sh = 0, dw = 0, dh = 0, ix = 1, iy = 1;
    bool fast = false;                                // both scale factors integral: integer box sums
    std::vector<int32_t> xfirst, yfirst;              // CSR: taps of destination index i are [first[i], first[i+1])
    std::vector<AreaTap> xtaps, ytaps;
};
struct AreaArgs {
    const uint8_t* src; size_t src_row_stride, src_frame_stride;
    uint8_t* dst; size_t dst_frame_stride;
    int dw, dh, ix, iy, fast;
    const int32_t* xfirst; const int32_t* yfirst; const AreaTap* xtaps; const AreaTap* ytaps;
};
bool build_area_plan(int sw, int sh, int dw, int dh, AreaPlanHost& p);
void launch_area(const AreaArgs& a, int channels, int nframes, cudaStream_t st);
struct CropItem { int32_t frame, x, y, w, h, pad_; uint64_t offset; };      // one face patch: source rect, byte offset in the packed output
void launch_crop(const uint8_t* frames, size_t frame_stride, size_t row_stride, int channels, const CropItem* items, int nitems,
                 uint8_t* out, cudaStream_t st);

// ---- OpenCV-2.4 compatibility mode (k_legacy.cu) ----
struct LegacyLevel {
    double factor;                                    // the scale as the double running product (rect coordinates use it)
    int32_t sw, sh, ww, wh, ystep, gx, gy;            // level size, window size in the frame, grid step, grid positions
    uint32_t img_off, int_off, xtab_off, ytab_off, map_off, pad_;
};
struct LegacyPlan {
    int w = 0, h = 0;
    std::vector<LegacyLevel> levels;
    std::vector<uint32_t> tab;                        // per level: sw x entries then sh y entries: (index + 1) | coefficient 1 << 16
    std::vector<int16_t> c0;                          // coefficient 0 (11 bits)
    size_t img_bytes = 0, int_entries = 0, map_entries = 0;
    int64_t grid_windows = 0, level_pixels = 0;
};
struct LegacyNode { int32_t rect[3][4]; float w[3]; float thr; int32_t left, right; float lval, rval; int32_t has3; };
struct LegacyCascadeHost { std::vector<LegacyNode> nodes; std::vector<float> thr24; std::vector<int32_t> child; };
struct LegacyArgs {
    const LegacyLevel* levels; const uint32_t* tab; const int16_t* c0;
    const uint8_t* src; size_t src_stride; int src_pitch, w, h;
    uint8_t* img; size_t img_stride;
    int32_t* S; double* SQ; size_t int_stride;
    int win_w, win_h, nstages;
    const LegacyNode* nodes; const int32_t* stage_first_tree; const int32_t* stage_ntrees; const int32_t* tree_first_node;
    const float* thr24; const int32_t* parent; const int32_t* next; const int32_t* child;
    ugd_rect* raw; double* rawsum; int rawcap; int* rawcount;
    int16_t* stage_map; unsigned long long* counters;
};
void build_legacy_plan(int w, int h, int win_w, int win_h, double scale_factor, int min_w, int min_h, int max_w, int max_h, LegacyPlan& p);
void build_legacy_cascade(const Cascade& c, LegacyCascadeHost& out);
void launch_legacy(const LegacyArgs& a, const LegacyPlan& p, int nframes, cudaStream_t st);

void launch_preprocess(const DevBufs& b, cudaStream_t st);
bool preprocess_can_fuse_area2(const uint8_t* src, size_t row_stride, size_t frame_stride, int scan_w);
void launch_equalize(const DevBufs& b, cudaStream_t st);
void launch_resize_colsum(const DevBufs& b, const WorkItem* wk, int nblocks, cudaStream_t st);   // pyramid + per-strip column sums
void launch_carry(const DevBufs& b, const WorkItem* wk, int nblocks, cudaStream_t st);
void launch_integral_bands(const DevBufs& b, const WorkItem* wk, int nblocks, cudaStream_t st);
void launch_tilted(const DevBufs& b, int max_level_w, cudaStream_t st);
cudaError_t init_tilted_attributes(int device, int max_w);
void launch_scan(const DevBufs& b, const WorkItem* wk, int nblocks, int nblocks_step2, cudaStream_t st, cudaStream_t side = nullptr,
                 cudaEvent_t ev_fork = nullptr, cudaEvent_t ev_join = nullptr);
void launch_group(const DevBufs& b, cudaStream_t st);
void launch_pack_lists(const uint8_t* recv, size_t msg_bytes, int world, int slot_frames, int cap, int skip_rank, int* packed,
                       size_t rect_off_bytes, cudaStream_t st);
cudaError_t init_kernel_attributes(int device, int win_w, int win_h, bool has_tilted);

}  // namespace ugd
