// Pyramid plan for one frame size: scale list, level geometry, buffer offsets, resize coefficient
// tables and the block->work tables every kernel is launched with.  Built once per (w, h) on the
// host, uploaded once, reused for every frame of that size.
// Semantics: SURVEY.md A.1 (scale list), A.2 (level size, resize coefficients), A.3 (grid).
#pragma once
#include <cstdint>
#include <vector>

namespace ugd {

constexpr int kStripRows = 32;       // integral strips: 32 output rows each
constexpr int kIntBand = 512;        // band integral: output columns per warp
constexpr int kIntSegRows = 64;      // band integral: pixel rows per warp (a multiple of kStripRows)
constexpr int kMaxLevels = 256;
constexpr int kScanRows = 16;        // grid rows per scan block (a scan tile is 64 x 16 windows, 8 warps)

// Device-visible description of one pyramid level (all offsets are per frame).
struct LevelInfo {
    int32_t sw, sh;            // level image size
    int32_t img_pitch;         // bytes per level-image row (multiple of 16, padding written as 0)
    int32_t P;                 // ints per integral row (multiple of 32); planes have sh+1 rows
    int32_t nx, ny;            // window origins: sw+1-win_w, sh+1-win_h
    int32_t step;              // grid step: scale >= 2 ? 1 : 2
    int32_t nxg, nyg;          // grid positions per row, grid rows
    int32_t nxg_pad;           // nxg rounded up to 32 (one warp segment never straddles rows)
    int32_t nstrips;           // ceil((sh+1)/32)
    float   scale;
    int32_t rect_w, rect_h;    // rint(win_w*scale), rint(win_h*scale)
    uint32_t img_off;          // bytes, into a frame's pyramid buffer
    uint32_t int_off;          // ints, into a frame's S (and SQ, and tilted T) buffer
    uint32_t pos_off;          // entries, into a frame's per-position arrays (padded layout)
    uint32_t carry_off;        // ints, into a frame's strip-carry arrays (nstrips x img_pitch)
    uint32_t xtab_off, ytab_off;  // entries, into the resize tables
    uint32_t map_off;          // entries, into the scan-order (unpadded) stage map
    uint32_t pad_;
};

struct WorkItem { int32_t level; int32_t first; };   // one CUDA block's share of a kernel

struct Plan {
    int w = 0, h = 0;
    std::vector<LevelInfo> levels;
    std::vector<uint32_t> xtab, ytab;           // entry = i0 | c1 << 16  (c1 in 0..256)
    std::vector<WorkItem> wk_resize, wk_carry, wk_bands, wk_scan;   // wk_resize: (strip | column chunk << 16) of k_resize_colsum
    // per-frame sizes
    int max_level_w = 0;                  // widest pyramid level (sizes the tilted-integral kernel's shared memory)
    size_t pyr_bytes = 0, int_ints = 0, pos_entries = 0, carry_ints = 0, map_entries = 0;
    int64_t level_pixels = 0, grid_windows = 0;
};

// Returns false when no level survives the min/max filter (the plan is then empty: zero detections).
void build_plan(int w, int h, int win_w, int win_h, double scale_factor, int min_w, int min_h, int max_w,
                int max_h, Plan& out);

}  // namespace ugd
