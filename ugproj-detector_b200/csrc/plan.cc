// Host-side plan builder (see plan.h).  Plain C++; all floating-point steps are written so that no
// contraction can change them (the file is compiled with -ffp-contract=off).
#include "plan.h"

#include <algorithm>
#include <cmath>

namespace ugd {
namespace {

inline int rint_d(double v) { return (int)nearbyint(v); }     // cvRound(double): half to even
inline int rint_f(float v) { return (int)nearbyintf(v); }     // cvRound(float)
inline uint32_t round_up(uint32_t v, uint32_t m) { return (v + m - 1) / m * m; }

// SURVEY A.2 coefficient rule for one axis: src length n -> dst length m.
void axis_table(int n, int m, uint32_t* tab) {
    double inv = (double)m / (double)n;
    double scale = 1.0 / inv;
    for (int d = 0; d < m; ++d) {
        double fx = ((double)d + 0.5) * scale - 0.5;
        int i = (int)floor(fx);
        double a = fx - (double)i;
        uint32_t i0, c1;
        if (i < 0) { i0 = 0; c1 = 0; }
        else if (i >= n - 1) {                        // last sample alone: written as (n-2, 256) so that i0 + 1 stays inside (k_resize_colsum)
            if (n >= 2) { i0 = (uint32_t)(n - 2); c1 = 256; } else { i0 = 0; c1 = 0; }
        }
        else { i0 = (uint32_t)i; c1 = (uint32_t)rint_d(a * 256.0); }
        tab[d] = i0 | c1 << 16;
    }
}

void add_work(std::vector<WorkItem>& v, int level, int64_t items, int per_block) {
    for (int64_t f = 0; f < items; f += per_block) v.push_back({level, (int32_t)f});
}

}  // namespace

void build_plan(int w, int h, int win_w, int win_h, double scale_factor, int min_w, int min_h, int max_w,
                int max_h, Plan& p) {
    p = Plan();
    p.w = w; p.h = h;
    if (max_w <= 0 || max_h <= 0) { max_w = w; max_h = h; }
    // A.1: running product in double, stored as float32; two filters.
    std::vector<float> scales;
    for (double f = 1.0;; f *= scale_factor) {
        if (rint_d(win_w * f) > w || rint_d(win_h * f) > h) break;
        float s = (float)f;
        int ww = rint_f((float)win_w * s), wh = rint_f((float)win_h * s);
        if (ww > max_w || wh > max_h) break;
        if (ww < min_w || wh < min_h) continue;
        scales.push_back(s);
        if ((int)scales.size() >= kMaxLevels) break;
    }
    uint32_t img_off = 0, int_off = 0, pos_off = 0, carry_off = 0, xt = 0, yt = 0, map_off = 0;
    // Rows the oracle really visits: cv2 cuts the rows of every level into the same number of stripes,
    // ns = ceil(working width of the first level / 32), each ceil(floor(ny / step) / ns) * step rows high, and
    // walks y < min(ns * stripe, ny) -- so a level with odd ny at step 2 loses its last grid row whenever the
    // stripes tile floor(ny / 2) rows exactly (verified against cv2 4.13 with a pass-everything cascade).
    int ns = 1;
    if (!scales.empty()) {
        const int nx0 = rint_f((float)w / scales[0]) + 1 - win_w;
        ns = std::max((nx0 + 31) / 32, 1);
    }
    for (float s : scales) {
        LevelInfo L{};
        L.scale = s;
        L.sw = rint_f((float)w / s);                      // A.2: float32 division, half-even
        L.sh = rint_f((float)h / s);
        L.nx = L.sw + 1 - win_w; L.ny = L.sh + 1 - win_h;
        if (L.nx <= 0 || L.ny <= 0) continue;
        L.img_pitch = (int32_t)round_up((uint32_t)L.sw, 16);
        L.P = (int32_t)round_up((uint32_t)L.sw + 1, 32);
        L.step = s >= 2.f ? 1 : 2;                        // A.3
        L.nxg = (L.nx + L.step - 1) / L.step;
        {
            const int stripe = std::max((L.ny / L.step + ns - 1) / ns, 1) * L.step;
            const int64_t lim = std::min<int64_t>((int64_t)ns * stripe, L.ny);
            L.nyg = (int32_t)((lim + L.step - 1) / L.step);
        }
        L.nxg_pad = (int32_t)round_up((uint32_t)L.nxg, 32);
        L.nstrips = (L.sh + 1 + kStripRows - 1) / kStripRows;
        L.rect_w = rint_f((float)win_w * s);
        L.rect_h = rint_f((float)win_h * s);
        L.img_off = img_off; L.int_off = int_off; L.pos_off = pos_off; L.carry_off = carry_off;
        L.xtab_off = xt; L.ytab_off = yt; L.map_off = map_off;
        int level = (int)p.levels.size();
        img_off += round_up((uint32_t)L.img_pitch * (uint32_t)L.sh, 256);
        int_off += round_up((uint32_t)L.P * (uint32_t)(L.sh + 1), 64);
        uint32_t npos = (uint32_t)L.nxg_pad * (uint32_t)L.nyg;
        pos_off += round_up(npos, 256);
        carry_off += (uint32_t)L.nstrips * (uint32_t)L.img_pitch;
        map_off += (uint32_t)L.nxg * (uint32_t)L.nyg;
        // resize tables (x table padded to the image pitch so uint4 loads stay in bounds)
        p.xtab.resize(xt + L.img_pitch, 0u);
        axis_table(w, L.sw, p.xtab.data() + xt);
        xt += (uint32_t)L.img_pitch;
        p.ytab.resize(yt + round_up((uint32_t)L.sh, 4), 0u);
        axis_table(h, L.sh, p.ytab.data() + yt);
        yt += round_up((uint32_t)L.sh, 4);
        // work tables
        int quads = L.img_pitch / 4;
        for (int strip = 0; strip < L.nstrips; ++strip)                                // one block per (strip of 32 rows, 256 columns); an empty last strip writes zero sums
            for (int ch = 0; ch * 64 < quads; ++ch) p.wk_resize.push_back({level, strip | ch << 16});
        add_work(p.wk_carry, level, quads, 256);
        for (int seg = 0; seg * kIntSegRows < L.sh; ++seg)                         // one warp per (segment of 64 rows, band of 512 integral columns)
            for (int bnd = 0; bnd * kIntBand < L.sw + 1; ++bnd) p.wk_bands.push_back({level, bnd | seg << 8});
        add_work(p.wk_scan, level, L.nyg, kScanRows);      // one block per kScanRows grid rows
        if (L.sw > p.max_level_w) p.max_level_w = L.sw;
        p.level_pixels += (int64_t)L.sw * L.sh;
        p.grid_windows += (int64_t)L.nxg * L.nyg;
        p.levels.push_back(L);
    }
    p.pyr_bytes = img_off; p.int_ints = int_off; p.pos_entries = pos_off; p.carry_ints = carry_off;
    p.map_entries = map_off;
}

}  // namespace ugd
