/* TEST INFRASTRUCTURE -- interface of the CPU oracle (see haar_oracle.c header comment). */
#ifndef HAAR_ORACLE_H
#define HAAR_ORACLE_H
#include <stddef.h>
#include <stdint.h>

#define ORC_MAX_LEVELS 256
#define ORC_LOWVAR (-1)   /* window failed the variance test, never evaluated        */
#define ORC_SKIP   (-2)   /* grid position skipped by the x-skip rule after a stage-0 rejection   */

typedef struct { int32_t x, y, w, h; } orc_rect;

/* Flat cascade; arrays are owned by the caller (numpy).  Layout == oracle/cascade_xml.py FlatCascade. */
typedef struct {
    int32_t win_w, win_h, nstages, ntrees, nnodes;
    const int32_t* stage_first_tree;
    const int32_t* stage_ntrees;
    const float*   stage_thr;        /* f32(xml) - 1e-5f */
    const int32_t* tree_first_node;
    const int32_t* tree_nnodes;
    const int32_t* node_rects;       /* [nnodes][3][4] x,y,w,h */
    const float*   node_weights;     /* [nnodes][3], 0 = absent */
    const float*   node_thr;
    const int32_t* node_left;        /* >0 child index within tree, 0 = leaf */
    const int32_t* node_right;
    const float*   node_left_val;
    const float*   node_right_val;
    const int32_t* node_tilted;      /* 1 = rects are 45-degree tilted: read the tilted integral */
    int32_t has_tilted;              /* any node tilted -> the tilted plane is built per level   */
} orc_cascade;

typedef struct {
    int64_t levels, level_pixels, grid_windows, windows_evaluated, lowvar, weak_evals, raw, grouped;
} orc_counters;

void orc_bgr2gray(const uint8_t* bgr, int n, uint8_t* gray);
void orc_bgr2ycrcb(const uint8_t* bgr, int n, uint8_t* ycrcb);
void orc_inrange3(const uint8_t* src, int n, const uint8_t lo[3], const uint8_t hi[3], uint8_t* mask);
void orc_equalize_hist(const uint8_t* src, int n, uint8_t* dst);
int  orc_resize_area(const uint8_t* src, int sw, int sh, int cn, size_t sstride, uint8_t* dst, int dw, int dh);
void orc_resize_exact(const uint8_t* src, int sw, int sh, size_t sstride, uint8_t* dst, int dw, int dh);
void orc_integral(const uint8_t* img, int w, int h, int32_t* S, int32_t* SQ);
void orc_integral_tilted(const uint8_t* img, int w, int h, int32_t* T);
int  orc_scale_list(int W, int H, int ow, int oh, double scale_factor, int min_w, int min_h, int max_w,
                    int max_h, float* scales, int cap);
void orc_level_size(int W, int H, float s, int* sw, int* sh);
int  orc_stripe_count(int W, int H, int ow, float first_scale);
int  orc_grid_rows(int ny, int step, int ns);
int  orc_run_at(const orc_cascade* c, const int32_t* S, const int32_t* SQ, const int32_t* T, int pitch, int x, int y,
                double* last_sum, int64_t* weak_evals);
int  orc_detect_raw(const orc_cascade* c, const uint8_t* gray, int W, int H, size_t stride, double scale_factor,
                    int min_w, int min_h, int max_w, int max_h, orc_rect* out, double* out_sum, int cap,
                    orc_counters* ctr, int16_t* stage_map, size_t map_cap);
int  orc_group_rectangles(const orc_rect* in, int n, int g, double eps, orc_rect* out, int* out_n_members);
int  orc_clip_rects(orc_rect* r, int n, int W, int H);
int  orc_detect_multiscale(const orc_cascade* c, const uint8_t* gray, int W, int H, size_t stride,
                           double scale_factor, int min_neighbors, int min_w, int min_h, int max_w, int max_h,
                           orc_rect* out, int cap, orc_counters* ctr);
int  orc_skin_filter(const uint8_t* mask, int W, const orc_rect* in, int n, orc_rect* out);
int  orc_detect_faces(const orc_cascade* c, const uint8_t* bgr, int W, int H, double scale_factor,
                      int min_neighbors, int min_w, int min_h, int max_w, int max_h, const uint8_t lo[3],
                      const uint8_t hi[3], orc_rect* out, int cap, orc_counters* ctr);
#endif
