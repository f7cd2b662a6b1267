/* TEST INFRASTRUCTURE -- CPU restatement of the OpenCV 2.4.x detection path.
 *
 * The reference's authors built against OpenCV 2.4.x (/root/reference/src/tracker/associator.cc:3 includes
 * opencv2/nonfree/features2d.hpp, a 2.4-only header; CMakeLists.txt:13 pins nothing).  With an old-format cascade,
 * cv::CascadeClassifier::detectMultiScale (/root/reference/src/tracker/detector.cc:25-31) in 2.4 goes through
 * cvHaarDetectObjectsForROC with CV_HAAR_SCALE_IMAGE, which differs from the 4.x evaluator the main oracle restates:
 *   - cvResize(INTER_LINEAR): 11-bit coefficients, ((b0*(S0>>4))>>16 + (b1*(S1>>4))>>16 + 2)>>2 vertical pass;
 *   - level size cvRound(img/factor) in double; window range x < sw - win_w, y < sh - win_h (one position fewer);
 *   - ystep = factor > 2 ? 1 : 2, no x-skip after a stage-0 rejection;
 *   - sum int32, sqsum double; mean / variance over the window inset by 1 px, variance >= 0 ? sqrt : 1;
 *   - rect weights pre-scaled by 1/area (rect 0 re-derived so that the weighted areas cancel), kept as float,
 *     feature = sum of float(int rect sum * float weight) in double, compared with threshold * std;
 *   - stage threshold bias 1e-4;
 *   - the STAGE TREE of <parent>/<next> (haarcascade_frontalface_alt_tree.xml:1303-1304, 1921-1922): a passed stage
 *     goes to its first child, a failed one to the `next` of its nearest ancestor-or-self that has one, reject if none,
 *     accept when a passed stage has no child;
 *   - groupRectangles(max(minNeighbors, 1), 0.2) when minNeighbors != 0; NO clipping to the image.
 *
 * PARITY UNPINNED: no OpenCV 2.4 exists in this image and the reference holds no vectors for it, so this file is a
 * restatement of the published 2.4 algorithm from the builder's reading of it; the GPU legacy mode is checked against
 * this file only (self-consistency).  It is test infrastructure: only tests/ may call it.
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include "haar_oracle.h"

static int cv_round_d(double v) { return (int)nearbyint(v); }
static int cv_floor_f(float v) { int i = (int)v; return i - (v < (float)i); }
static short sat_short(float v) { int r = (int)nearbyintf(v); return (short)(r < -32768 ? -32768 : r > 32767 ? 32767 : r); }

/* cvResize(src, dst, CV_INTER_LINEAR) for 8UC1, fixed point (INTER_RESIZE_COEF_BITS = 11). */
void orc24_axis(int ssize, int dsize, int32_t* ofs, int16_t* coef /* [dsize][2] */, int* dmax) {
    const double inv = (double)dsize / ssize, scale = 1.0 / inv;
    *dmax = dsize;
    for (int d = 0; d < dsize; ++d) {
        float f = (float)((d + 0.5) * scale - 0.5);
        int s = cv_floor_f(f);
        f -= (float)s;
        /* x axis rules (the y axis keeps f and clips the rows when reading, see axis_y) */
        if (s < 0) { f = 0.f; s = 0; }
        if (s + 1 >= ssize) { if (d < *dmax) *dmax = d; if (s >= ssize - 1) { f = 0.f; s = ssize - 1; } }
        ofs[d] = s;
        coef[2 * d] = sat_short((1.f - f) * 2048.f);
        coef[2 * d + 1] = sat_short(f * 2048.f);
    }
}
static void axis_y(int ssize, int dsize, int32_t* ofs, int16_t* coef) {
    const double inv = (double)dsize / ssize, scale = 1.0 / inv;
    for (int d = 0; d < dsize; ++d) {
        float f = (float)((d + 0.5) * scale - 0.5);
        int s = cv_floor_f(f);
        f -= (float)s;
        ofs[d] = s;
        coef[2 * d] = sat_short((1.f - f) * 2048.f);
        coef[2 * d + 1] = sat_short(f * 2048.f);
    }
}

void orc24_resize_linear(const uint8_t* src, int sw, int sh, size_t sstride, uint8_t* dst, int dw, int dh) {
    int32_t* xofs = malloc(sizeof(int32_t) * (size_t)dw); int16_t* xa = malloc(sizeof(int16_t) * 2 * (size_t)dw);
    int32_t* yofs = malloc(sizeof(int32_t) * (size_t)dh); int16_t* yb = malloc(sizeof(int16_t) * 2 * (size_t)dh);
    int xmax;
    orc24_axis(sw, dw, xofs, xa, &xmax);
    axis_y(sh, dh, yofs, yb);
    for (int dy = 0; dy < dh; ++dy) {
        int y0 = yofs[dy], y1 = yofs[dy] + 1;
        y0 = y0 < 0 ? 0 : y0 > sh - 1 ? sh - 1 : y0;
        y1 = y1 < 0 ? 0 : y1 > sh - 1 ? sh - 1 : y1;
        const uint8_t* S0 = src + (size_t)y0 * sstride;
        const uint8_t* S1 = src + (size_t)y1 * sstride;
        const int b0 = yb[2 * dy], b1 = yb[2 * dy + 1];
        for (int dx = 0; dx < dw; ++dx) {
            const int sx = xofs[dx];
            int d0, d1;
            if (dx < xmax) {
                d0 = S0[sx] * xa[2 * dx] + S0[sx + 1] * xa[2 * dx + 1];
                d1 = S1[sx] * xa[2 * dx] + S1[sx + 1] * xa[2 * dx + 1];
            } else { d0 = S0[sx] * 2048; d1 = S1[sx] * 2048; }
            const int v = (((b0 * (d0 >> 4)) >> 16) + ((b1 * (d1 >> 4)) >> 16) + 2) >> 2;
            dst[(size_t)dy * dw + dx] = (uint8_t)(v < 0 ? 0 : v > 255 ? 255 : v);
        }
    }
    free(xofs); free(xa); free(yofs); free(yb);
}

void orc24_integral(const uint8_t* img, int w, int h, int32_t* S, double* SQ) {
    const int P = w + 1;
    for (int x = 0; x <= w; ++x) { S[x] = 0; SQ[x] = 0.0; }
    for (int y = 0; y < h; ++y) {
        int32_t s = 0; double q = 0.0;
        S[(size_t)(y + 1) * P] = 0; SQ[(size_t)(y + 1) * P] = 0.0;
        for (int x = 0; x < w; ++x) {
            const int v = img[(size_t)y * w + x];
            s += v; q += (double)v * v;
            S[(size_t)(y + 1) * P + x + 1] = S[(size_t)y * P + x + 1] + s;
            SQ[(size_t)(y + 1) * P + x + 1] = SQ[(size_t)y * P + x + 1] + q;
        }
    }
}

/* BGR -> gray with the 2.4 coefficients (14 bits: 1868, 9617, 4899). */
void orc24_bgr2gray(const uint8_t* bgr, int n, uint8_t* gray) {
    for (int i = 0; i < n; ++i) gray[i] = (uint8_t)((bgr[3 * i] * 1868 + bgr[3 * i + 1] * 9617 + bgr[3 * i + 2] * 4899 + 8192) >> 14);
}

/* Scale list of cvHaarDetectObjectsForROC (CV_HAAR_SCALE_IMAGE): factors as doubles. */
int orc24_scale_list(int W, int H, int ow, int oh, double scale_factor, int min_w, int min_h, int max_w, int max_h,
                     double* factors, int cap) {
    if (max_w <= 0 || max_h <= 0) { max_w = W; max_h = H; }
    int n = 0;
    for (double f = 1.0;; f *= scale_factor) {
        const int ww = cv_round_d(ow * f), wh = cv_round_d(oh * f);
        const int sw = cv_round_d(W / f), sh = cv_round_d(H / f);
        if (sw - ow + 1 <= 0 || sh - oh + 1 <= 0) break;
        if (ww > max_w || wh > max_h) break;
        if (ww < min_w || wh < min_h) continue;
        if (n < cap) factors[n] = f;
        if (++n >= cap) break;
    }
    return n;
}

/* The hidden cascade's float weights (cvSetImagesForHaarClassifierCascade at scale 1). */
void orc24_weights(const orc_cascade* c, float* w /* [nnodes][3] */) {
    const double ws = 1.0 / ((double)(c->win_w - 2) * (c->win_h - 2));
    for (int n = 0; n < c->nnodes; ++n) {
        double sum0 = 0.0, area0 = 0.0;
        for (int k = 0; k < 3; ++k) {
            const int32_t* r = c->node_rects + (size_t)n * 12 + 4 * k;
            const float wk = c->node_weights[3 * n + k];
            if (wk == 0.f && k > 0) { w[3 * n + k] = 0.f; continue; }
            const double corr = c->node_tilted[n] ? ws * 0.5 : ws;
            w[3 * n + k] = (float)(wk * corr);
            if (k == 0) area0 = (double)r[2] * r[3];
            else sum0 += (double)w[3 * n + k] * r[2] * r[3];
        }
        w[3 * n] = (float)(-sum0 / area0);
    }
}

static int rect_sum(const int32_t* S, int P, int x, int y, const int32_t* r) {
    const int32_t* p = S + (size_t)(y + r[1]) * P + x + r[0];
    return p[0] - p[r[2]] - p[(size_t)r[3] * P] + p[(size_t)r[3] * P + r[2]];
}

/* cvRunHaarClassifierCascadeSum: 1 = accepted, else 0; *path (optional) receives the last stage evaluated. */
int orc24_run_at(const orc_cascade* c, const float* w24, const float* thr24, const int32_t* parent, const int32_t* next,
                 const int32_t* child, const int32_t* S, const double* SQ, int P, int x, int y, int* last_stage, int64_t* weak) {
    const int ew = c->win_w - 2, eh = c->win_h - 2;
    const double inv_area = 1.0 / ((double)ew * eh);
    const int32_t* p = S + (size_t)(y + 1) * P + x + 1;
    const double* q = SQ + (size_t)(y + 1) * P + x + 1;
    const double mean = (p[0] - p[ew] - p[(size_t)eh * P] + p[(size_t)eh * P + ew]) * inv_area;
    double vnf = q[0] - q[ew] - q[(size_t)eh * P] + q[(size_t)eh * P + ew];
    vnf = vnf * inv_area - mean * mean;
    vnf = vnf >= 0.0 ? sqrt(vnf) : 1.0;
    int st = 0;
    while (st >= 0) {
        double stage_sum = 0.0;
        for (int t = c->stage_first_tree[st]; t < c->stage_first_tree[st] + c->stage_ntrees[st]; ++t) {
            const int base = c->tree_first_node[t];
            int idx = 0;
            float leaf;
            for (;;) {
                const int n = base + idx;
                const double thr = (double)c->node_thr[n] * vnf;
                double sum = (float)rect_sum(S, P, x, y, c->node_rects + (size_t)n * 12) * w24[3 * n];
                sum += (float)rect_sum(S, P, x, y, c->node_rects + (size_t)n * 12 + 4) * w24[3 * n + 1];
                if (c->node_weights[3 * n + 2] != 0.f) sum += (float)rect_sum(S, P, x, y, c->node_rects + (size_t)n * 12 + 8) * w24[3 * n + 2];
                if (weak) ++*weak;
                const int lt = sum < thr;
                const int go = lt ? c->node_left[n] : c->node_right[n];
                leaf = lt ? c->node_left_val[n] : c->node_right_val[n];
                if (!go) break;
                idx = go;
            }
            stage_sum += (double)leaf;
        }
        if (last_stage) *last_stage = st;
        if (stage_sum >= (double)thr24[st]) {
            if (child[st] < 0) return 1;
            st = child[st];
        } else {
            while (st >= 0 && next[st] < 0) st = parent[st];
            if (st < 0) return 0;
            st = next[st];
        }
    }
    return 0;
}

/* cvHaarDetectObjects(..., CV_HAAR_SCALE_IMAGE) up to (not including) the grouping.  thr_raw = f32(xml) stage
 * thresholds; parent / next as in the XML (-1 = none).  Returns the number of candidates (unclipped).
 * stage_map (optional): last stage evaluated per grid position, or nstages for accepted windows. */
int orc24_detect_raw(const orc_cascade* c, const float* thr_raw, const int32_t* parent, const int32_t* next,
                     const uint8_t* gray, int W, int H, size_t stride, double scale_factor, int min_w, int min_h,
                     int max_w, int max_h, orc_rect* out, int cap, orc_counters* ctr, int16_t* stage_map, size_t map_cap) {
    double factors[ORC_MAX_LEVELS];
    const int nl = orc24_scale_list(W, H, c->win_w, c->win_h, scale_factor, min_w, min_h, max_w, max_h, factors, ORC_MAX_LEVELS);
    float* w24 = malloc(sizeof(float) * 3 * (size_t)c->nnodes);
    float* thr24 = malloc(sizeof(float) * (size_t)c->nstages);
    int32_t* child = malloc(sizeof(int32_t) * (size_t)c->nstages);
    orc24_weights(c, w24);
    for (int s = 0; s < c->nstages; ++s) { thr24[s] = thr_raw[s] - 0.0001f; child[s] = -1; }
    for (int s = 0; s < c->nstages; ++s)
        if (parent[s] >= 0 && child[parent[s]] < 0) child[parent[s]] = s;
    uint8_t* tight = malloc((size_t)W * H);
    for (int y = 0; y < H; ++y) memcpy(tight + (size_t)y * W, gray + (size_t)y * stride, (size_t)W);
    uint8_t* img = malloc((size_t)W * H);
    int32_t* S = malloc(sizeof(int32_t) * (size_t)(W + 1) * (H + 1));
    double* SQ = malloc(sizeof(double) * (size_t)(W + 1) * (H + 1));
    int n = 0;
    size_t mpos = 0;
    if (ctr) memset(ctr, 0, sizeof *ctr);
    for (int l = 0; l < nl; ++l) {
        const double f = factors[l];
        const int sw = cv_round_d(W / f), sh = cv_round_d(H / f);
        const int ww = cv_round_d(c->win_w * f), wh = cv_round_d(c->win_h * f);
        orc24_resize_linear(tight, W, H, (size_t)W, img, sw, sh);
        orc24_integral(img, sw, sh, S, SQ);
        const int ystep = f > 2.0 ? 1 : 2, P = sw + 1;
        const int xe = sw - c->win_w, ye = sh - c->win_h;           /* exclusive: one position fewer than the 4.x grid */
        if (ctr) { ctr->levels++; ctr->level_pixels += (int64_t)sw * sh; }
        for (int y = 0; y < ye; y += ystep)
            for (int x = 0; x < xe; x += ystep) {
                int last = 0;
                const int ok = orc24_run_at(c, w24, thr24, parent, next, child, S, SQ, P, x, y, &last, ctr ? &ctr->weak_evals : NULL);
                if (ctr) { ctr->grid_windows++; ctr->windows_evaluated++; }
                if (stage_map && mpos < map_cap) stage_map[mpos] = (int16_t)(ok ? c->nstages : last);
                ++mpos;
                if (ok) {
                    if (n < cap) { out[n].x = cv_round_d(x * f); out[n].y = cv_round_d(y * f); out[n].w = ww; out[n].h = wh; }
                    ++n;
                }
            }
    }
    if (ctr) ctr->raw = n;
    free(w24); free(thr24); free(child); free(tight); free(img); free(S); free(SQ);
    return n;
}

/* == cv::CascadeClassifier::detectMultiScale of 2.4 with an old-format cascade: candidates, grouping, no clip. */
int orc24_detect_multiscale(const orc_cascade* c, const float* thr_raw, const int32_t* parent, const int32_t* next,
                            const uint8_t* gray, int W, int H, size_t stride, double scale_factor, int min_neighbors,
                            int min_w, int min_h, int max_w, int max_h, orc_rect* out, int cap, orc_counters* ctr) {
    const int rcap = 1 << 18;
    orc_rect* raw = malloc(sizeof(orc_rect) * (size_t)rcap);
    int n = orc24_detect_raw(c, thr_raw, parent, next, gray, W, H, stride, scale_factor, min_w, min_h, max_w, max_h, raw, rcap, ctr, NULL, 0);
    if (n > rcap) n = rcap;
    int m;
    if (min_neighbors != 0) {
        orc_rect* g = malloc(sizeof(orc_rect) * (size_t)(n > 0 ? n : 1));
        int* cnt = malloc(sizeof(int) * (size_t)(n > 0 ? n : 1));
        m = orc_group_rectangles(raw, n, min_neighbors > 1 ? min_neighbors : 1, 0.2, g, cnt);
        for (int i = 0; i < m && i < cap; ++i) out[i] = g[i];
        free(g); free(cnt);
    } else {
        m = n;
        for (int i = 0; i < m && i < cap; ++i) out[i] = raw[i];
    }
    if (ctr) ctr->grouped = m;
    free(raw);
    return m;
}
