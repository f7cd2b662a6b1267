// Device-side view of a detector's buffers and the launch wrappers of every kernel on the path.
// All kernels are hand-written for sm_100a; none has a CPU fallback.
#pragma once
#include <cuda_runtime.h>

#include <cstdint>

#include "../../include/ugdet.h"
#include "cascade.h"
#include "plan.h"

namespace ugd {

constexpr int kMaxDenseNodes = 1024;      // nodes of the densely evaluated stages, held in __constant__
constexpr int kCtrPerFrame = 8;           // per-frame device counters (see enum below)
enum { CTR_EVALUATED = 0, CTR_LOWVAR, CTR_QUEUED, CTR_RAW, CTR_GROUPED, CTR_NEAR, CTR_RAW_OVERFLOW, CTR_WEAK };

// Stage-map codes (match the oracle's): -2 skipped, -1 low variance, k rejected at stage k, nstages accepted.
constexpr int16_t kMapSkip = -2, kMapLowVar = -1;

struct DevBufs {
    // geometry
    int w, h, nlevels, nframes;
    int win_w, win_h, nstages, dense_stages, stumps_only;
    const LevelInfo* levels;
    const uint32_t* xtab; const uint32_t* ytab;
    // cascade (tail evaluator, SoA in global memory; dense stages live in __constant__)
    const uint4* nodeA; const uint4* nodeB; const uint2* nodeC;
    const int32_t* stage_first_tree; const int32_t* stage_ntrees; const float* stage_thr;
    const int32_t* tree_first_node; const int32_t* stage_first_node;
    // frames
    const uint8_t* bgr; size_t bgr_stride;            // input BGR frames (device)
    uint8_t* gray_raw; uint8_t* gray_eq; uint8_t* mask; size_t px_stride;   // per-frame planes
    const uint8_t* src_gray; size_t src_stride; int src_pitch;             // what the pyramid is built from
    uint32_t* hist;                                   // [nframes][256]
    uint8_t skin_lo[3], skin_hi[3];
    int use_skin;
    // pyramid + integrals
    uint8_t* pyr; size_t pyr_stride;
    int32_t* S; int32_t* SQ; size_t int_stride;
    uint32_t* carryS; uint32_t* carryQ; size_t carry_stride;
    // window scan
    float* posv; size_t pos_stride;                   // >0 vnf (alive after stage 0), 0 low variance, -1 failed stage 0
    uint2* queue; int* qcount;                        // per frame: pos_stride entries, count
    ugd_rect* raw; double* rawsum; int rawcap; int* rawcount;
    int16_t* stage_map;                               // frame 0 only, padded layout; may be null
    // grouping / output
    int* grp_scratch; size_t grp_stride;              // per frame: 6*rawcap ints
    ugd_rect* out; int outcap; int* outcount;
    int min_neighbors;
    unsigned long long* counters;                     // [nframes][kCtrPerFrame]
};

// dense-stage nodes -> __constant__ (call before launching stage0/dense with a new cascade)
cudaError_t upload_dense_nodes(const PackedNode* nodes, int n, cudaStream_t st);

// first_node: n entries (stage s covers nodes [first_node[s], first_node[s+1])), thr: n-1 entries
cudaError_t upload_dense_stage_table(const int* first_node, const float* thr, int n, cudaStream_t st);

void launch_preprocess(const DevBufs& b, cudaStream_t st);
void launch_equalize(const DevBufs& b, cudaStream_t st);
void launch_resize(const DevBufs& b, const WorkItem* wk, int nblocks, cudaStream_t st);
void launch_colsum(const DevBufs& b, const WorkItem* wk, int nblocks, cudaStream_t st);
void launch_carry(const DevBufs& b, const WorkItem* wk, int nblocks, cudaStream_t st);
void launch_integral(const DevBufs& b, const WorkItem* wk, int nblocks, cudaStream_t st);
void launch_stage0(const DevBufs& b, const WorkItem* wk, int nblocks, cudaStream_t st);
void launch_dense(const DevBufs& b, const WorkItem* wk, int nblocks, cudaStream_t st);
void launch_tail(const DevBufs& b, int blocks_per_frame, cudaStream_t st);
void launch_group(const DevBufs& b, cudaStream_t st);
cudaError_t init_kernel_attributes(int win_w, int win_h);

}  // namespace ugd
