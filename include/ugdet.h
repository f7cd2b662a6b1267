/*
 * ugdet.h -- C ABI of the B200-native Haar-cascade face-detection engine (libugdet.so).
 *
 * This is the drop-in boundary for ONE hot path of kyukyukyu/ugproj-detector:
 *     ugproj::FaceDetector::detectFaces            (reference src/tracker/detector.cc:10-43,
 *                                                   class at src/tracker/detector.h:10-19)
 * and the slice of OpenCV it drives (cvtColor, inRange, equalizeHist,
 * CascadeClassifier::load / detectMultiScale, mean).  The reference has no FFI layer; its seam is
 * that C++ class.  Everything below is plain C (pointers + sizes, no OpenCV / torch types) so the
 * reference's C++ host code can bind it directly -- see INTEGRATION.md for the 20-line
 * FaceDetector shim a maintainer adds.
 *
 * Conventions (mirror the reference: int return, 0 = OK, non-zero = failure,
 * src/file_io.cc:7-15, src/tracker/main.cc:19-41):
 *   - every function returns UGD_OK (0) or a UGD_ERR_* code; nothing throws across the ABI;
 *   - the caller owns every input and output buffer;
 *   - a ugd_cascade is immutable after load and may be shared by many detectors;
 *   - a ugd_detector is single-threaded (one per host thread / CUDA stream);
 *   - there is NO CPU fallback: without a CUDA device ugd_create fails with UGD_ERR_CUDA.
 */
#ifndef UGDET_H
#define UGDET_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define UGD_OK               0
#define UGD_ERR_ARG          1   /* bad argument (null pointer, size out of range)                  */
#define UGD_ERR_IO           2   /* file cannot be opened / read                                    */
#define UGD_ERR_PARSE        3   /* malformed XML / INI                                             */
#define UGD_ERR_UNSUPPORTED  4   /* cascade uses something the engine does not implement (LBP/HOG
                                    features, non-integer rect weights, > 255 tree nodes ...)       */
#define UGD_ERR_CUDA         5   /* CUDA runtime error; text via ugd_last_error()                   */
#define UGD_ERR_CAPACITY     6   /* frame larger than the detector was created for, or an internal
                                    per-frame candidate buffer overflowed                           */
#define UGD_ERR_NOMEM        7
#define UGD_ERR_TRUNCATED    8   /* some frame produced more rectangles than `cap`: out holds the first cap
                                    of each such frame, counts[] / *n hold the TRUE counts -- call again
                                    with cap >= the largest count (the reference's std::vector is unbounded) */
#define UGD_ERR_COMM         9   /* NCCL missing or failed; text via ugd_last_error()               */

/* == cv::Rect (x, y, width, height as int32) -- detector.h:15 std::vector<cv::Rect>& rects */
typedef struct { int32_t x, y, width, height; } ugd_rect;

/* Runtime parameters of the path.  Defaults = the reference's literals / cfg defaults. */
typedef struct {
    double  scale_factor;     /* detection.scale, default 1.1            (src/structure.cc:83-85)   */
    int32_t min_neighbors;    /* literal 2                               (src/tracker/detector.cc:29) */
    int32_t min_w, min_h;     /* literal Size(30,30)                     (detector.cc:31)            */
    int32_t max_w, max_h;     /* 0,0 = no maxSize (the reference passes none)                        */
    uint8_t skin_lower[3];    /* Y,Cr,Cb  default 0,133,77               (structure.cc:75-78)        */
    uint8_t skin_upper[3];    /* Y,Cr,Cb  default 255,173,127            (structure.cc:79-82)        */
    uint8_t _pad[2];
    int32_t dense_stages;     /* engine tuning: cascade stages evaluated inside the scan tiles before
                                 survivors move to the block-local work queue; 0 = engine default.
                                 Does not change results.                                            */
    int32_t max_raw_per_frame;/* capacity of the per-frame candidate list fed to groupRectangles;
                                 0 = engine default (8192)                                           */
    int32_t legacy_24;        /* 1 = OpenCV-2.4 compatibility mode: the evaluator of the
                                 library generation the reference was written against -- cvResize's 11-bit
                                 INTER_LINEAR, double squared integral, float weights scaled by 1/area,
                                 stage bias 1e-4, ystep = factor > 2 ? 1 : 2 without x-skip, no final clip,
                                 and the cascade file's STAGE TREE (<parent>/<next>) walked as a tree
                                 (haarcascade_frontalface_alt_tree.xml).  Slower path, same C ABI.
                                 Parity unpinned: no OpenCV 2.4 exists to check it against.           */
} ugd_params;

typedef struct ugd_cascade  ugd_cascade;
typedef struct ugd_detector ugd_detector;

typedef struct {
    int32_t win_w, win_h;     /* detection window of the cascade (20x20 / 24x24)                     */
    int32_t nstages, ntrees, nnodes;
    int32_t max_tree_nodes;   /* 1 = stump cascade                                                   */
    int32_t old_format;       /* 1 = type_id "opencv-haar-classifier" (the reference data XMLs)         */
    int32_t sum_exact;        /* 1 = every stage sum is exactly representable in fp64, so the
                                 engine's parallel summation order cannot change a decision         */
    int32_t has_tilted;       /* 1 = some features are 45-degree tilted: the engine also builds the
                                 tilted integral of every pyramid level            */
} ugd_cascade_info;

/* Work counters of the most recent detect call (summed over its frames). */
typedef struct {
    int64_t frames;
    int64_t levels;             /* pyramid levels per frame                                          */
    int64_t level_pixels;       /* sum of level pixels per frame                                     */
    int64_t grid_windows;       /* grid positions (all frames)                                       */
    int64_t windows_evaluated;  /* grid positions minus x-skipped ones == oracle runAt calls         */
    int64_t lowvar;             /* windows rejected by the variance test                             */
    int64_t queued;             /* windows that survived the dense stages and entered a work queue */
    int64_t raw;                /* accepted windows before grouping                                  */
    int64_t grouped;            /* rectangles returned                                               */
    int64_t near_threshold;     /* windows whose stage sum lay within 1e-5 relative of a stage
                                   threshold (north-star counter; informational)                     */
    int64_t weak_evals;         /* weak-classifier nodes evaluated at evaluated windows (== the oracle's
                                   count for stump cascades; the stage-0 pre-pass also touches
                                   x-skipped positions, which are not counted)                       */
} ugd_counters;

/* Names of the engine's kernels, in launch order, for ugd_get_kernel_times(). */
enum { UGD_K_PREPROCESS = 0,    /* BGR -> gray + skin mask + histogram                (a2-a4)        */
       UGD_K_EQUALIZE,          /* LUT from histogram, apply, AND mask                (a5-a6)        */
       UGD_K_RESIZE,            /* all pyramid levels, INTER_LINEAR_EXACT             (a7)           */
       UGD_K_COLSUM,            /* integral phase 1: per-strip column sums            (a8)           */
       UGD_K_CARRY,             /* integral phase 2: scan of strip sums               (a8)           */
       UGD_K_INTEGRAL,          /* integral phase 3: S and SQ planes                  (a8)           */
       UGD_K_TILTED,            /* tilted integral plane; launched only for cascades with tilted
                                   features                                           (8(f) N4)      */
       UGD_K_SCAN,              /* tiled window scan: variance norm, x-skip, in-tile stages on
                                   per-bank-class survivor masks, block-local queue   (a9)           */
       UGD_K_GROUP,             /* groupRectangles + clip + skin post-filter          (a10-a12)      */
       UGD_K_AREA,              /* INTER_AREA decimation to the scan size (ugd_detect_batch_scaled)  (8(f) N1) */
       UGD_K_COUNT };

/* ---- parameters / configuration -------------------------------------------------------------- */
void ugd_params_default(ugd_params* p);

/* Reads the [detection] section (and [scan] frame_size) of a reference conf/<name>.cfg with the
 * reference's keys and defaults (structure.cc:62-85); a relative cascade path is resolved against
 * the cfg's directory (structure.cc:199-204).  scan_w/scan_h receive -1,-1 when absent
 * (structure.cc:64-65).  Replaces Boost.program_options for this section only. */
int ugd_config_load(const char* cfg_path, ugd_params* p, char* cascade_path, size_t cascade_path_cap,
                    int32_t* scan_w, int32_t* scan_h);

/* ---- cascade model: replaces cv::CascadeClassifier::load (src/file_io.cc:11) ------------------ */
int  ugd_cascade_load_xml(const char* path, ugd_cascade** out);
int  ugd_cascade_get_info(const ugd_cascade* c, ugd_cascade_info* info);
/* Copies the parsed model out as flat arrays (any pointer may be NULL).  Sizes: stage_* [nstages],
 * tree_* [ntrees], node_rects [nnodes*12] (3 x (x,y,w,h)), node_weights [nnodes*3], others [nnodes].
 * stage_thr already holds f32(xml) - 1e-5f; node_tilted is 1 for nodes whose rects are 45-degree tilted.
 * Used by the loader parity tests. */
int  ugd_cascade_dump(const ugd_cascade* c, int32_t* stage_first_tree, int32_t* stage_ntrees, float* stage_thr,
                      int32_t* tree_first_node, int32_t* tree_nnodes, int32_t* node_rects, float* node_weights,
                      float* node_thr, int32_t* node_left, int32_t* node_right, float* node_left_val,
                      float* node_right_val, int32_t* node_tilted);
void ugd_cascade_free(ugd_cascade* c);

/* ---- detector: replaces ugproj::FaceDetector (detector.h:10-19) ------------------------------- */
/* max_w/max_h: largest frame that will be passed; max_batch: frames processed per launch group. */
int  ugd_create(const ugd_cascade* c, const ugd_params* p, int cuda_device, int max_w, int max_h, int max_batch,
                ugd_detector** out);
void ugd_destroy(ugd_detector* d);
/* Run on a caller-provided cudaStream_t (e.g. torch's current stream); NULL = the detector's own. */
int  ugd_set_stream(ugd_detector* d, void* cuda_stream);

/* == FaceDetector::detectFaces(frame, rects): BGR 8UC3 host frame -> rect list.  Writes at most
 * `cap` rects to out and the true count to *n (the C++ shim appends them to the caller's vector,
 * as detector.cc:41 does). */
int  ugd_detect_bgr(ugd_detector* d, const uint8_t* bgr, int w, int h, size_t stride_bytes, ugd_rect* out,
                    int cap, int* n);
/* == cascade.detectMultiScale(gray, ...) on a ready 8UC1 image (no equalizeHist, no skin logic). */
int  ugd_detect_gray(ugd_detector* d, const uint8_t* gray, int w, int h, size_t stride_bytes, ugd_rect* out,
                     int cap, int* n);
/* Many frames per call.  frames: nframes tightly packed frames (w*h*(3 if is_bgr else 1) bytes each),
 * in host memory (is_device=0; copied inside the call) or device memory (is_device=1).
 * out: nframes*cap rects (host); counts: nframes (host).  Frames beyond max_batch are processed in
 * successive groups of max_batch. */
int  ugd_detect_batch(ugd_detector* d, const void* frames, int is_device, int is_bgr, int nframes, int w, int h,
                      ugd_rect* out, int cap, int* counts);

/* ---- observability / parity hooks ------------------------------------------------------------- */
/* Prefetching form of ugd_detect_batch: submit enqueues the copies, kernels and result copies of the
 * whole batch and returns at once, so the host tracker (face_tracker.cc:140-210, sequential, stateful) can consume
 * batch k while the GPU detects batch k+1; wait blocks until the submitted batch is done and delivers its lists in
 * frame order (out: nframes * cap rects, counts: nframes).  One batch in flight per detector; the frame memory passed
 * to submit must stay valid and unchanged until wait returns. */
int  ugd_submit_batch(ugd_detector* d, const void* frames, int is_device, int is_bgr, int nframes, int w, int h, int cap);
int  ugd_wait_batch(ugd_detector* d, ugd_rect* out, int* counts);

/* The step right before detectFaces in the tracker: cv::resize(orig, scan.frame_size, INTER_AREA)
 * (src/tracker/face_tracker.cc:112-114, cfg key scan.frame_size, src/structure.cc:63-67), fused in front of
 * the detector: full-size BGR frames in, rects in SCAN-frame coordinates out (what track_frame works in).
 * scan_w/scan_h <= 0 means "no resize" (structure.cc:64-65).  Decimation only. */
int  ugd_detect_batch_scaled(ugd_detector* d, const void* frames, int is_device, int nframes, int src_w, int src_h,
                             int scan_w, int scan_h, ugd_rect* out, int cap, int* counts);
/* The face patches the tracker keeps per detection -- `cv::Mat img_face(curr_frame, *it); Face f(curr_index, *it, img_face);`
 * (src/tracker/face_tracker.cc:296-302; Face's copy constructor copies the pixels, src/structure.h:137-146) -- cut out of
 * DEVICE-resident frames, for hosts whose frames never visit host memory (decoded on the GPU).
 * dev_frames: nframes tight frames of w*h*channels bytes in device memory; rects / counts: as returned by ugd_detect_batch
 * (cap slots per frame; rects must lie inside the frame, which the detect calls guarantee).  out (host) receives the patches
 * back to back, each row-major and tight (width*height*channels bytes), frame by frame in list order; offsets[f*cap+i] is the
 * byte offset of patch i of frame f; *total the bytes needed.  out_cap < *total returns UGD_ERR_CAPACITY with *total set and
 * nothing copied (call again with room). */
int  ugd_crop_faces(ugd_detector* d, const void* dev_frames, int nframes, int w, int h, int channels, const ugd_rect* rects,
                    const int* counts, int cap, uint8_t* out, size_t out_cap, size_t* offsets, size_t* total);
/* The resize alone (host in, host out; 1 or 3 channels), for parity tests against cv2.resize(INTER_AREA). */
int  ugd_resize_area(ugd_detector* d, const uint8_t* src, int src_w, int src_h, int channels, size_t stride_bytes,
                     uint8_t* dst, int dst_w, int dst_h);

/* ---- multi-GPU: one process per GPU, frames sharded, lists gathered to the root ---------------------
 * detectFaces has no inter-frame state (the tracker does: face_tracker.cc:140-149, and it stays on the root's host), so
 * rank r of N takes the contiguous frame range ugd_shard_range(F, r, N) and the ONE exchange step of the path is the
 * gather of the per-frame lists to the root.  The engine does it itself with NCCL send/recv over NVLink, straight from
 * the device-side result slots, on the detector's stream (libnccl.so.2 is bound at run time with dlopen; without it
 * these calls return UGD_ERR_COMM and everything else keeps working).
 *   ugd_comm_unique_id : ncclGetUniqueId -- the root calls it and hands the 128 bytes to every rank by whatever
 *                        means the host has (a file, MPI, torch.distributed ...);
 *   ugd_comm_init      : ncclCommInitRank on the detector's device (collective: every rank calls it);
 *   ugd_comm_attach    : use an ncclComm_t the host already owns instead (not destroyed by ugd_destroy);
 *   ugd_detect_batch_sharded : ugd_detect_batch on this rank's `nframes` frames (nframes <= slot_frames, the same
 *                        slot_frames and cap on every rank), then the gather.  On the root, out / counts hold
 *                        world x slot_frames frames: frame f of rank r at index r * slot_frames + f (out: cap rects
 *                        each), counts -1 for the slots a rank did not use.  On the other ranks out / counts receive
 *                        that rank's own nframes lists, as from ugd_detect_batch.  Collective: every rank calls it. */
#define UGD_COMM_ID_BYTES 128
int  ugd_comm_unique_id(uint8_t* id /* UGD_COMM_ID_BYTES */);
int  ugd_comm_init(ugd_detector* d, const uint8_t* id, int rank, int world);
int  ugd_comm_attach(ugd_detector* d, void* nccl_comm, int rank, int world);
int  ugd_comm_info(const ugd_detector* d, int* rank, int* world, int* nccl_version);
void ugd_shard_range(int nframes, int rank, int world, int* lo, int* hi);
int  ugd_detect_batch_sharded(ugd_detector* d, const void* frames, int is_device, int is_bgr, int nframes, int w, int h,
                              int slot_frames, int root, ugd_rect* out, int cap, int* counts);

int  ugd_get_counters(const ugd_detector* d, ugd_counters* c);
/* Per-kernel device time (ms, CUDA events on the detector's stream) of the most recent detect call;
 * recorded only after ugd_set_profiling(d, 1).  times_ms: UGD_K_COUNT entries. */
int  ugd_set_profiling(ugd_detector* d, int on);
int  ugd_get_kernel_times(const ugd_detector* d, float* times_ms, int64_t* launches);
const char* ugd_kernel_name(int k);

/* Pyramid geometry for a frame size: number of levels, and per level the float32 scale, level
 * width/height and grid step.  Arrays may be NULL; cap = their length. */
int  ugd_plan_info(ugd_detector* d, int w, int h, int* nlevels, float* scales, int32_t* lw, int32_t* lh,
                   int32_t* step, int cap);
/* Same, without a detector or a GPU (host arithmetic only): geometry for cascade c under params p. */
int  ugd_plan_query(const ugd_cascade* c, const ugd_params* p, int w, int h, int* nlevels, float* scales,
                    int32_t* lw, int32_t* lh, int32_t* step, int cap, int64_t* grid_windows, int64_t* level_pixels);
/* Raw view of the most recent single-frame gray/bgr detect (frame 0 of a batch): every accepted
 * window before grouping (unclipped) and its final-stage sum -- the same observables as cv2's
 * detectMultiScale(minNeighbors=0) / detectMultiScale3(outputRejectLevels). */
int  ugd_get_raw(ugd_detector* d, ugd_rect* out, double* sums, int cap, int* n);
/* When enabled, the next detect calls also record, for frame 0, the fate of every grid position in
 * scan order (level, row, column): -2 x-skipped, -1 low variance, k = rejected at stage k,
 * nstages = accepted.  ugd_get_stage_map copies it out (cap entries; *n = grid positions). */
int  ugd_set_stage_map(ugd_detector* d, int on);
int  ugd_get_stage_map(ugd_detector* d, int16_t* map, size_t cap, size_t* n);
/* Intermediate images of frame 0 of the most recent call, for per-subsystem parity tests:
 *   ugd_get_preprocessed: equalised+masked gray (w*h) and skin mask (w*h) after a BGR call;
 *   ugd_get_level: level image (lw*lh, tight) and integral planes S, SQ ((lh+1)*(lw+1), tight);
 *   ugd_get_level_tilted: the tilted integral plane of that level (cascades with tilted features only). */
int  ugd_get_preprocessed(ugd_detector* d, uint8_t* gray_eq, uint8_t* mask);
int  ugd_get_level(ugd_detector* d, int level, uint8_t* img, int32_t* S, int32_t* SQ);
int  ugd_get_level_tilted(ugd_detector* d, int level, int32_t* T);

/* Measured roofline denominator of k_scan: sustained GB/s of a conflict-free stream of 32-bit shared-memory loads on
 * every SM of the device (one probe kernel, CUDA-event timed; takes ~20 ms).  bytes_per_clk_per_sm (may be NULL)
 * relates it to the nominal 128 B/clk at the device's rated clock. */
int  ugd_measure_lds_peak(int cuda_device, double* gbytes_per_s, double* bytes_per_clk_per_sm);

/* Integral pass without any allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid it).  d_img: 8-bit image in DEVICE memory (h rows of w bytes, stride_bytes apart); d_S / d_SQ: DEVICE
 * planes of (h+1) x (w+1) int32, tight, receiving the same values as ugd_get_level's S and SQ (sum and squared sum,
 * wrapping int32, zero first row and column).  The work is enqueued on cuda_stream (a cudaStream_t; NULL = the default
 * stream) of device cuda_device, which is left current; the call does not synchronise.
 *   - The scratch of the pass (per-strip column sums) is a file-scope static __device__ array of 4 MiB: every GPU that
 *     loads the library holds it until the process exits.  Limit: w <= UGD_NOALLOC_MAX_W, h <= UGD_NOALLOC_MAX_H;
 *     anything larger returns UGD_ERR_CAPACITY (and says so on stderr).
 *   - No cudaMalloc, cudaFree or allocating library routine is called: three kernel launches, plus, on the first call
 *     per device, one cudaGetSymbolAddress (make that first call outside a graph capture).
 *   - Every call shares the array, which keeps the last call's contents; each call overwrites all it reads, so no
 *     clear is needed, but CALLS THAT OVERLAP ON THE GPU RACE: keep them on one stream or order them with events. */
#define UGD_NOALLOC_MAX_W 4096
#define UGD_NOALLOC_MAX_H 4096
int  ugd_integral_noalloc(int cuda_device, const uint8_t* d_img, int w, int h, size_t stride_bytes, int32_t* d_S,
                          int32_t* d_SQ, void* cuda_stream);

const char* ugd_strerror(int code);
const char* ugd_last_error(const ugd_detector* d);   /* detail of the detector's last failure */
const char* ugd_version(void);

#ifdef __cplusplus
}
#endif
#endif /* UGDET_H */
