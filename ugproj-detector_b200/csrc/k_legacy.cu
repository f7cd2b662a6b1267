// OpenCV-2.4 compatibility mode (ugd_params.legacy_24): what cv::CascadeClassifier::detectMultiScale did
// in the library generation the reference was written against (old-format cascade -> cvHaarDetectObjects with
// CV_HAAR_SCALE_IMAGE).  It differs from the 4.x evaluator of k_cascade.cu in every numeric detail -- cvResize's 11-bit
// INTER_LINEAR, level sizes rounded in double, a window range one position shorter, ystep = factor > 2 ? 1 : 2 with
// no x-skip, a double squared integral, mean / variance over the inset window, float weights pre-scaled by 1/area,
// stage threshold bias 1e-4 -- and it walks the STAGE TREE of the cascade file (<parent>/<next>: pass -> first child,
// fail -> `next` of the nearest ancestor-or-self that has one, reject if none), which is what makes
// haarcascade_frontalface_alt_tree.xml a tree.  No clipping after grouping.
//
// This mode exists for behavioural compatibility, not for speed: one thread per pixel / row / column / window, node
// data from global memory.  Parity is UNPINNED (no OpenCV 2.4 in the image): tests compare it with oracle/haar_legacy.c.
#include <algorithm>
#include <cmath>

#include "kernels.cuh"

namespace ugd {

namespace {

inline int cv_round_d(double v) { return (int)nearbyint(v); }
inline int cv_floor_f(float v) { int i = (int)v; return i - (v < (float)i); }
inline int sat_short(float v) { int r = (int)nearbyintf(v); return r < -32768 ? -32768 : r > 32767 ? 32767 : r; }

// one axis of cvResize(INTER_LINEAR): entry = source index | coefficient of the second sample (11 bits) << 16;
// `x_rules`: the x axis clamps at both ends (and multiplies by 2048 beyond xmax, which the same formula gives with f = 0)
void axis24(int ssize, int dsize, bool x_rules, uint32_t* tab, int16_t* c0) {
    const double inv = (double)dsize / ssize, scale = 1.0 / inv;
    for (int d = 0; d < dsize; ++d) {
        float f = (float)((d + 0.5) * scale - 0.5);
        int s = cv_floor_f(f);
        f -= (float)s;
        if (x_rules) {
            if (s < 0) { f = 0.f; s = 0; }
            if (s >= ssize - 1) { f = 0.f; s = ssize - 1; }
        }
        const int a0 = sat_short((1.f - f) * 2048.f), a1 = sat_short(f * 2048.f);
        // rows are clipped when read (y axis); a second sample beyond the last column carries coefficient 0 (x axis)
        tab[d] = (uint32_t)(s + 1) | (uint32_t)a1 << 16;          // index stored + 1 so that -1 fits
        c0[d] = (int16_t)a0;
    }
}

}  // namespace

void build_legacy_plan(int w, int h, int win_w, int win_h, double scale_factor, int min_w, int min_h, int max_w, int max_h,
                       LegacyPlan& p) {
    p = LegacyPlan();
    p.w = w; p.h = h;
    if (max_w <= 0 || max_h <= 0) { max_w = w; max_h = h; }
    uint32_t img_off = 0, int_off = 0, tab = 0, map_off = 0;
    for (double f = 1.0;; f *= scale_factor) {
        LegacyLevel L{};
        L.ww = cv_round_d(win_w * f); L.wh = cv_round_d(win_h * f);
        L.sw = cv_round_d(w / f); L.sh = cv_round_d(h / f);
        if (L.sw - win_w + 1 <= 0 || L.sh - win_h + 1 <= 0) break;
        if (L.ww > max_w || L.wh > max_h) break;
        if (L.ww < min_w || L.wh < min_h) continue;
        L.factor = f;
        L.ystep = f > 2.0 ? 1 : 2;
        L.gx = std::max((L.sw - win_w + L.ystep - 1) / L.ystep, 0);      // x in [0, sw - win_w) step ystep
        L.gy = std::max((L.sh - win_h + L.ystep - 1) / L.ystep, 0);
        L.img_off = img_off; L.int_off = int_off; L.xtab_off = tab; L.ytab_off = tab + (uint32_t)L.sw; L.map_off = map_off;
        img_off += (uint32_t)((L.sw * L.sh + 255) / 256 * 256);
        int_off += (uint32_t)(((L.sw + 1) * (L.sh + 1) + 63) / 64 * 64);
        map_off += (uint32_t)(L.gx * L.gy);
        p.tab.resize(tab + (size_t)L.sw + L.sh);
        p.c0.resize(tab + (size_t)L.sw + L.sh);
        axis24(w, L.sw, true, p.tab.data() + tab, p.c0.data() + tab);
        axis24(h, L.sh, false, p.tab.data() + tab + L.sw, p.c0.data() + tab + L.sw);
        tab += (uint32_t)(L.sw + L.sh);
        p.grid_windows += (int64_t)L.gx * L.gy;
        p.level_pixels += (int64_t)L.sw * L.sh;
        p.levels.push_back(L);
        if ((int)p.levels.size() >= kMaxLevels) break;
    }
    p.img_bytes = img_off; p.int_entries = int_off; p.map_entries = map_off;
}

// The hidden cascade of 2.4: float weights scaled by 1/area with rect 0 re-derived, thresholds biased by 1e-4, the
// first-child table of the stage tree.
void build_legacy_cascade(const Cascade& c, LegacyCascadeHost& out) {
    out = LegacyCascadeHost();
    const double ws = 1.0 / ((double)(c.win_w - 2) * (c.win_h - 2));
    out.nodes.resize((size_t)c.nnodes());
    for (int n = 0; n < c.nnodes(); ++n) {
        const HostNode& h = c.nodes[n];
        LegacyNode& o = out.nodes[n];
        double sum0 = 0.0, area0 = 0.0;
        for (int k = 0; k < 3; ++k) {
            for (int j = 0; j < 4; ++j) o.rect[k][j] = h.rect[k][j];
            if (h.weight[k] == 0.f && k > 0) { o.w[k] = 0.f; continue; }
            o.w[k] = (float)(h.weight[k] * (h.tilted ? ws * 0.5 : ws));
            if (k == 0) area0 = (double)h.rect[k][2] * h.rect[k][3];
            else sum0 += (double)o.w[k] * h.rect[k][2] * h.rect[k][3];
        }
        o.w[0] = (float)(-sum0 / area0);
        o.has3 = h.weight[2] != 0.f;
        o.thr = h.thr; o.left = h.left; o.right = h.right; o.lval = h.left_val; o.rval = h.right_val;
    }
    const int ns = c.nstages();
    out.thr24.resize((size_t)ns); out.child.assign((size_t)ns, -1);
    for (int s = 0; s < ns; ++s) out.thr24[s] = c.stage_thr_raw[s] - 0.0001f;
    for (int s = 0; s < ns; ++s)
        if (c.stage_parent[s] >= 0 && out.child[c.stage_parent[s]] < 0) out.child[c.stage_parent[s]] = s;
}

// ---- kernels ---------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k24_resize(LegacyArgs a, int level) {
    const LegacyLevel& L = a.levels[level];
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i >= L.sw * L.sh) return;
    const int dy = i / L.sw, dx = i - dy * L.sw;
    const uint8_t* src = a.src + (size_t)blockIdx.y * a.src_stride;
    const uint32_t xe = a.tab[L.xtab_off + dx], ye = a.tab[L.ytab_off + dy];
    const int a0 = a.c0[L.xtab_off + dx], a1 = (int)(int16_t)(xe >> 16), b0 = a.c0[L.ytab_off + dy], b1 = (int)(int16_t)(ye >> 16);
    const int sx = (int)(xe & 0xffffu) - 1, sx1 = min(sx + 1, a.w - 1);
    const int sy = (int)(ye & 0xffffu) - 1, y0 = min(max(sy, 0), a.h - 1), y1 = min(max(sy + 1, 0), a.h - 1);
    const uint8_t* S0 = src + (size_t)y0 * a.src_pitch;
    const uint8_t* S1 = src + (size_t)y1 * a.src_pitch;
    const int d0 = S0[sx] * a0 + S0[sx1] * a1, d1 = S1[sx] * a0 + S1[sx1] * a1;
    const int v = (((b0 * (d0 >> 4)) >> 16) + ((b1 * (d1 >> 4)) >> 16) + 2) >> 2;
    a.img[(size_t)blockIdx.y * a.img_stride + L.img_off + i] = (uint8_t)min(max(v, 0), 255);
}

// integral, pass 1: prefix along every row (row 0 and column 0 of the planes are zero)
__global__ void __launch_bounds__(128) k24_rows(LegacyArgs a, int level) {
    const LegacyLevel& L = a.levels[level];
    const int y = blockIdx.x * 128 + threadIdx.x;                 // plane row 0 .. sh
    if (y > L.sh) return;
    const int P = L.sw + 1;
    int32_t* S = a.S + (size_t)blockIdx.y * a.int_stride + L.int_off + (size_t)y * P;
    double* Q = a.SQ + (size_t)blockIdx.y * a.int_stride + L.int_off + (size_t)y * P;
    S[0] = 0; Q[0] = 0.0;
    if (y == 0) { for (int x = 1; x <= L.sw; ++x) { S[x] = 0; Q[x] = 0.0; } return; }
    const uint8_t* img = a.img + (size_t)blockIdx.y * a.img_stride + L.img_off + (size_t)(y - 1) * L.sw;
    int32_t s = 0; double q = 0.0;
    for (int x = 0; x < L.sw; ++x) {
        const int v = img[x];
        s += v; q += (double)(v * v);                            // integers below 2^53: every partial sum is exact
        S[x + 1] = s; Q[x + 1] = q;
    }
}
// pass 2: prefix down every column
__global__ void __launch_bounds__(128) k24_cols(LegacyArgs a, int level) {
    const LegacyLevel& L = a.levels[level];
    const int x = blockIdx.x * 128 + threadIdx.x + 1;
    if (x > L.sw) return;
    const int P = L.sw + 1;
    int32_t* S = a.S + (size_t)blockIdx.y * a.int_stride + L.int_off + x;
    double* Q = a.SQ + (size_t)blockIdx.y * a.int_stride + L.int_off + x;
    int32_t s = 0; double q = 0.0;
    for (int y = 1; y <= L.sh; ++y) {
        s += S[(size_t)y * P]; q += Q[(size_t)y * P];
        S[(size_t)y * P] = s; Q[(size_t)y * P] = q;
    }
}

__device__ __forceinline__ int rect_sum24(const int32_t* __restrict__ S, int P, const int32_t* r) {
    const int32_t* p = S + r[1] * P + r[0];
    return p[0] - p[r[2]] - p[r[3] * P] + p[r[3] * P + r[2]];
}

// cvRunHaarClassifierCascadeSum, one thread per grid position
__global__ void __launch_bounds__(128) k24_scan(LegacyArgs a, int level) {
    const LegacyLevel& L = a.levels[level];
    const int i = blockIdx.x * 128 + threadIdx.x;
    if (i >= L.gx * L.gy) return;
    const int frame = blockIdx.y;
    const int gy = i / L.gx, gx = i - gy * L.gx;
    const int x = gx * L.ystep, y = gy * L.ystep, P = L.sw + 1;
    const int32_t* S = a.S + (size_t)frame * a.int_stride + L.int_off + (size_t)y * P + x;
    const double* Q = a.SQ + (size_t)frame * a.int_stride + L.int_off + (size_t)y * P + x;
    const int ew = a.win_w - 2, eh = a.win_h - 2;
    const double inv_area = __ddiv_rn(1.0, __dmul_rn((double)ew, (double)eh));
    const int32_t* p = S + P + 1;
    const double* q = Q + P + 1;
    const double mean = __dmul_rn((double)(p[0] - p[ew] - p[eh * P] + p[eh * P + ew]), inv_area);
    double vnf = __dadd_rn(__dadd_rn(__dadd_rn(q[0], -q[ew]), -q[eh * P]), q[eh * P + ew]);
    vnf = __dadd_rn(__dmul_rn(vnf, inv_area), -__dmul_rn(mean, mean));
    vnf = vnf >= 0.0 ? sqrt(vnf) : 1.0;
    int st = 0, last = 0;
    bool accepted = false;
    unsigned weak = 0;
    while (st >= 0) {
        double stage_sum = 0.0;
        const int t0 = a.stage_first_tree[st], t1 = t0 + a.stage_ntrees[st];
        for (int t = t0; t < t1; ++t) {
            const int base = a.tree_first_node[t];
            int idx = 0;
            float leaf;
            for (;;) {
                const LegacyNode& nd = a.nodes[base + idx];
                const double thr = __dmul_rn((double)nd.thr, vnf);
                double sum = (double)__fmul_rn((float)rect_sum24(S, P, nd.rect[0]), nd.w[0]);
                sum = __dadd_rn(sum, (double)__fmul_rn((float)rect_sum24(S, P, nd.rect[1]), nd.w[1]));
                if (nd.has3) sum = __dadd_rn(sum, (double)__fmul_rn((float)rect_sum24(S, P, nd.rect[2]), nd.w[2]));
                ++weak;
                const bool lt = sum < thr;
                const int go = lt ? nd.left : nd.right;
                leaf = lt ? nd.lval : nd.rval;
                if (!go) break;
                idx = go;
            }
            stage_sum = __dadd_rn(stage_sum, (double)leaf);
        }
        last = st;
        if (stage_sum >= (double)a.thr24[st]) {
            if (a.child[st] < 0) { accepted = true; break; }
            st = a.child[st];
        } else {
            while (st >= 0 && a.next[st] < 0) st = a.parent[st];
            if (st < 0) break;
            st = a.next[st];
        }
    }
    if (a.stage_map && frame == 0) a.stage_map[L.map_off + i] = (int16_t)(accepted ? a.nstages : last);
    atomicAdd(&a.counters[(size_t)frame * kCtrPerFrame + CTR_WEAK], (unsigned long long)weak);
    atomicAdd(&a.counters[(size_t)frame * kCtrPerFrame + CTR_EVALUATED], 1ull);
    if (accepted) {
        const int slot = atomicAdd(&a.rawcount[frame], 1);
        if (slot < a.rawcap) {
            ugd_rect r;
            r.x = __double2int_rn(__dmul_rn((double)x, L.factor));
            r.y = __double2int_rn(__dmul_rn((double)y, L.factor));
            r.width = L.ww; r.height = L.wh;
            a.raw[(size_t)frame * a.rawcap + slot] = r;
            a.rawsum[(size_t)frame * a.rawcap + slot] = 0.0;
        }
    }
}

void launch_legacy(const LegacyArgs& a, const LegacyPlan& p, int nframes, cudaStream_t st) {
    for (int l = 0; l < (int)p.levels.size(); ++l) {
        const LegacyLevel& L = p.levels[l];
        k24_resize<<<dim3((L.sw * L.sh + 255) / 256, nframes), 256, 0, st>>>(a, l);
        k24_rows<<<dim3((L.sh + 1 + 127) / 128, nframes), 128, 0, st>>>(a, l);
        k24_cols<<<dim3((L.sw + 127) / 128, nframes), 128, 0, st>>>(a, l);
        if (L.gx * L.gy > 0) k24_scan<<<dim3((L.gx * L.gy + 127) / 128, nframes), 128, 0, st>>>(a, l);
    }
}

}  // namespace ugd
