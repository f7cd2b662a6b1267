// C ABI of libugdet.so (include/ugdet.h): detector object, device buffers, launch sequence.
// Host orchestration only -- all arithmetic of the path happens in the kernels of k_*.cu.
#include <algorithm>
#include <array>
#include <atomic>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include <cuda.h>       // CUtensorMap and its enums only: cuTensorMapEncodeTiled is fetched through cudaGetDriverEntryPoint, libcuda is not linked
#include <dlfcn.h>
#include <nccl.h>      // types and prototypes only: libnccl.so.2 is bound at run time (dlopen), never at link time

#include "kernels.cuh"

using namespace ugd;

namespace {

struct DevPtr {                       // grow-only device allocation
    void* p = nullptr; size_t cap = 0;
    cudaError_t reserve(size_t bytes) {
        if (bytes <= cap) return cudaSuccess;
        if (p) { cudaFree(p); p = nullptr; cap = 0; }
        cudaError_t e = cudaMalloc(&p, bytes);
        if (e == cudaSuccess) cap = bytes;
        return e;
    }
    void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
    template <class T> T* as() const { return static_cast<T*>(p); }
};

struct DevPlan {                      // a Plan plus its device copies
    Plan host;
    DevPtr levels, xtab, ytab, wk_resize, wk_carry, wk_bands, wk_scan;
    DevPtr wk_scan_chunk; std::vector<WorkItem> host_wk_scan_chunk;   // one block per (row band, tile): calls with few frames
    // TMA tensor maps of the integral planes, one table per buffer set; rebuilt when the planes move or grow
    DevPtr tmaps[2];
    const void* tm_S[2] = {nullptr, nullptr}; const void* tm_T[2] = {nullptr, nullptr};
    int tm_frames[2] = {0, 0};
    bool tm_ok[2] = {false, false};
    void release() { levels.release(); xtab.release(); ytab.release(); wk_resize.release();
                     wk_carry.release(); wk_bands.release(); wk_scan.release(); wk_scan_chunk.release(); tmaps[0].release(); tmaps[1].release(); }
};

struct DevAreaPlan {                  // INTER_AREA tables for one (source size -> scan size) pair
    AreaPlanHost host;
    DevPtr xfirst, yfirst, xtaps, ytaps;
    void release() { xfirst.release(); yfirst.release(); xtaps.release(); ytaps.release(); }
};

// The dense-stage nodes live in ONE __constant__ image per device, shared by every detector on that device.  The bank
// records whose image is resident (cascade uid + dense_stages: detectors of the same cascade share it without a reload)
// and which detectors have scans in flight against it; a detector with a different image waits (on its own stream, no
// host block) for those scans before it uploads its own.  The mutex is held from the check through the k_scan launches
// and the last-use event, so no other detector can swap the image between a bind and the launches that rely on it.
struct ConstBank {
    uint64_t uid = 0; int dense = -1;
    std::vector<ugd_detector*> users;
};
std::mutex g_const_mu;
ConstBank g_bank[64];

inline size_t round_up(size_t v, size_t m) { return (v + m - 1) / m * m; }

}  // namespace

struct ugd_detector {
    const Cascade* casc = nullptr;
    uint64_t casc_uid = 0;
    ugd_params prm{};
    int device = 0, max_w = 0, max_h = 0, max_batch = 1;
    int dense_stages = 1, dense_nodes = 0, rawcap = 8192, tile_min = 32;
    cudaStream_t own_stream = nullptr, stream = nullptr, copy_stream = nullptr;
    cudaEvent_t ev_copied[2] = {nullptr, nullptr}, ev_consumed[2] = {nullptr, nullptr};   // host-input pipeline
    // ugd_detect_batch_scaled with a 2x2 decimation: the full-size frames of the group being enqueued (k_preprocess decimates
    // them itself, see launch_preprocess) and the event to record once that kernel -- their last reader -- is in the stream
    const uint8_t* area2_src = nullptr; size_t area2_row_stride = 0, area2_frame_stride = 0; cudaEvent_t area2_consumed = nullptr;
    DevPtr stage2;                                    // second input staging buffer (the first is bgr / gray_in)
    DevPtr crop_items, crop_out;                      // ugd_crop_faces: patch descriptors, packed patches
    // Host frames in PAGEABLE memory (what a cv::Mat is, face_tracker.cc:112-114): cudaMemcpyAsync would bounce them through
    // the driver's staging buffer on the calling thread, one chunk at a time, and the kernels of the next group could not be
    // enqueued meanwhile.  Multi-group calls instead copy each group into one of two pinned buffers of the detector with a
    // few host threads (memcpy at memory speed) and issue a truly asynchronous copy from there.
    uint8_t* pin[2] = {nullptr, nullptr}; size_t pin_bytes = 0;
    // group pipeline of multi-group calls: the image stage (preprocess .. integrals, HBM-bound) of group g+1 runs on
    // img_stream next to the scan stage (k_scan + grouping, issue-bound) of group g; S / SQ / T / mask exist twice
    cudaStream_t img_stream = nullptr;
    // per buffer set: the grid-step-2 scan launch runs on scan_side[set], next to the grid-step-1 launch; in multi-group calls
    // the scan stage of set 1 has its own stream (scan2), so consecutive groups overlap and no launch tail leaves SMs idle
    cudaStream_t scan_side[2] = {nullptr, nullptr}, scan2 = nullptr;
    cudaEvent_t ev_scan_fork[2] = {nullptr, nullptr}, ev_scan_join[2] = {nullptr, nullptr}, ev_set1 = nullptr;
    DevPtr raw2, rawsum2, rawcount2, grp2, out2, outcount2, counters2;
    cudaEvent_t ev_img_done[2] = {nullptr, nullptr}, ev_scan_done[2] = {nullptr, nullptr}, ev_call = nullptr;
    bool set_used[2] = {false, false};
    DevPtr S2, SQ2, T2, mask2;
    std::map<std::pair<int, int>, std::unique_ptr<DevPlan>> plans;
    std::map<std::vector<int>, std::unique_ptr<DevAreaPlan>> area_plans;
    DevPtr src_full, area_dst;                        // full-size input frames / standalone resize output
    // cascade on device
    DevPtr qbase;
    DevPtr nodeC, nodeM, nodeTA[2], nodeTB[2], st_first_tree, st_ntrees, st_thr, tree_first_node, st_first_node;
    // frame buffers
    DevPtr bgr, gray_raw, gray_eq, mask, hist, pyr, S, SQ, T, carryS, carryQ, raw, rawsum,
        rawcount, grp, out, outcount, counters, stage_map, gray_in;
    // host staging (pinned)
    ugd_rect* h_out = nullptr; int* h_counts = nullptr; unsigned long long* h_ctr = nullptr;
    size_t h_out_cap = 0, h_frames_cap = 0;
    // state of the last call
    DevBufs last{};
    const DevPlan* last_plan = nullptr;
    bool last_was_bgr = false;
    ugd_counters ctr{};
    struct { bool active = false; int nframes = 0, cap = 0; DevPlan* plan = nullptr; } pending;   // ugd_submit_batch ... ugd_wait_batch
    bool profiling = false, want_map = false;
    bool tma = false;                                 // k_scan stages its grid-step-1 tiles with TMA (the encoder exists and the box fits)
    bool deint2 = false;                              // S / SQ rows of the grid-step-2 levels are stored de-interleaved (TMA staging of both tile layouts)
    bool truncated = false;                           // some frame of the current call had more rects than cap
    float ktime[UGD_K_COUNT] = {0}; int64_t klaunch[UGD_K_COUNT] = {0};
    cudaEvent_t ev[UGD_K_COUNT + 1] = {nullptr};
    cudaEvent_t ev_area[2] = {nullptr, nullptr};
    // multi-GPU gather (ugd_comm_*, ugd_detect_batch_sharded)
    struct {
        ncclComm_t comm = nullptr; bool own = false; int rank = 0, world = 1;
        bool call = false; int slot_frames = 0;       // a sharded call is being enqueued: run_group also files its lists in `send`
        DevPtr send, recv, packed;
        int* h_packed = nullptr; size_t h_packed_cap = 0;
    } shard;
    // OpenCV-2.4 compatibility mode (ugd_params.legacy_24): its own plans, cascade tables and planes
    struct LegacyDev {
        LegacyPlan host;
        DevPtr levels, tab, c0;
    };
    std::map<std::pair<int, int>, std::unique_ptr<LegacyDev>> lg_plans;
    DevPtr lg_nodes, lg_thr24, lg_parent, lg_next, lg_child, lg_img, lg_S, lg_SQ;
    const LegacyDev* last_lg = nullptr;
    ScanConstImage const_img;                         // host image of this detector's __constant__ data
    cudaEvent_t ev_const_use[2] = {nullptr, nullptr}; // recorded after every k_scan launch of this detector (one per scan stream)
    bool const_used[2] = {false, false};
    cudaEvent_t ev_bank_upload = nullptr;             // this detector's image was (re)uploaded on one of its scan streams ...
    bool upload_wait[2] = {false, false};             // ... and the other one has to wait for that copy before it scans
    std::string err;
    int sm_count = 148;
};

namespace {

int fail_cuda(ugd_detector* d, cudaError_t e, const char* what) {
    d->err = std::string(what) + ": " + cudaGetErrorString(e);
    return UGD_ERR_CUDA;
}
#define CK(call)                                                        \
    do { cudaError_t e_ = (call); if (e_ != cudaSuccess) return fail_cuda(d, e_, #call); } while (0)

template <class T>
int upload(ugd_detector* d, DevPtr& dst, const std::vector<T>& v) {
    CK(dst.reserve(std::max<size_t>(v.size() * sizeof(T), 16)));
    if (!v.empty()) CK(cudaMemcpyAsync(dst.p, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice, d->stream));
    return UGD_OK;
}

int get_plan(ugd_detector* d, int w, int h, DevPlan** out) {
    auto key = std::make_pair(w, h);
    auto it = d->plans.find(key);
    if (it != d->plans.end()) { *out = it->second.get(); return UGD_OK; }
    std::unique_ptr<DevPlan> dp(new DevPlan);
    const Cascade& c = *d->casc;
    build_plan(w, h, c.win_w, c.win_h, d->prm.scale_factor, d->prm.min_w, d->prm.min_h, d->prm.max_w, d->prm.max_h, dp->host);
    build_scan_work(make_scan_geom(c.win_w, c.win_h, c.has_tilted), dp->host.levels, dp->host.wk_scan);   // tile height depends on the level's grid step
    build_scan_work_chunked(make_scan_geom(c.win_w, c.win_h, c.has_tilted), dp->host.levels, dp->host_wk_scan_chunk);
    int rc;
    if ((rc = upload(d, dp->levels, dp->host.levels))) return rc;
    if ((rc = upload(d, dp->xtab, dp->host.xtab))) return rc;
    if ((rc = upload(d, dp->ytab, dp->host.ytab))) return rc;
    if ((rc = upload(d, dp->wk_resize, dp->host.wk_resize))) return rc;
    if ((rc = upload(d, dp->wk_carry, dp->host.wk_carry))) return rc;
    if ((rc = upload(d, dp->wk_bands, dp->host.wk_bands))) return rc;
    if ((rc = upload(d, dp->wk_scan, dp->host.wk_scan))) return rc;
    if ((rc = upload(d, dp->wk_scan_chunk, dp->host_wk_scan_chunk))) return rc;
    if (c.has_tilted) CK(init_tilted_attributes(d->device, dp->host.max_level_w));
    CK(cudaStreamSynchronize(d->stream));      // host vectors stay alive in dp->host, but keep it simple
    *out = dp.get();
    d->plans[key] = std::move(dp);
    return UGD_OK;
}

// Corner slots of the per-lane evaluator (k_scan's block-local queue): the 32 lanes of a warp evaluate 32 consecutive
// nodes of a stage against ONE window, slot by slot, so the shared-memory banks of a slot's 32 loads are whatever the
// nodes' corner offsets happen to be (2.8 wavefronts per load on frontalface_alt).  A feature sum does not care in
// which order its corners are added, so every node's corners are dealt to the slots such that the 32 nodes of a
// chunk hit different banks where possible: greedy placement (cost = distinct addresses already in that bank of
// that slot), 2-opt inside the node, three refinement sweeps per chunk.  1.5 wavefronts per load on the same cascade.
// Measured effect on the step is small (k_scan 8.93 -> 8.89 ms per 64 frames: the kernel is issue-bound, the queue is
// 18 % of it); it is kept because it takes 8 % of the kernel's shared-memory wavefronts away for one pass over the table.
// Banks are those of the grid-step-1 tile layout (80 % of the scan time); the step-2 layout keeps what it gets.
void assign_lane_slots(const Cascade& c, const ScanGeom& geom, const std::vector<int>& order, std::vector<std::array<int8_t, 12>>& perm) {
    perm.assign(c.nnodes(), std::array<int8_t, 12>{0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11});
    if (c.max_tree_nodes != 1 || getenv("UGD_NO_SLOT_ORDER")) return;
    auto addr = [&](const CornerNode& cn, int k) { return (int)scan_tile_off(geom, 0, cn.cx[k], cn.cy[k]) + (cn.plane ? (int)scan_plane_bytes(geom, 0) : 0); };
    for (int s = 0; s < c.nstages(); ++s)
        for (int i0 = c.stage_first_node[s]; i0 < c.stage_first_node[s + 1]; i0 += 32) {
            const int i1 = std::min(i0 + 32, c.stage_first_node[s + 1]);
            std::vector<std::pair<int, int>> load[12][32];                       // slot, bank -> (address, nodes using it)
            auto cost = [&](int a, int slot) {
                const auto& v = load[slot][(a >> 2) & 31];
                for (const auto& e : v) if (e.first == a) return 0;              // same word: broadcast, no extra wavefront
                return (int)v.size();
            };
            auto put = [&](int a, int slot, int sign) {
                auto& v = load[slot][(a >> 2) & 31];
                for (size_t k = 0; k < v.size(); ++k)
                    if (v[k].first == a) { if ((v[k].second += sign) == 0) v.erase(v.begin() + (long)k); return; }
                if (sign > 0) v.push_back({a, 1});
            };
            auto place = [&](int i) {
                const CornerNode& cn = c.corners[order[i]];
                const int n = cn.nc <= 6 ? 6 : cn.nc <= 8 ? 8 : cn.nc <= 9 ? 9 : 12;     // slots the evaluator walks for this node
                int a[12]; for (int k = 0; k < n; ++k) a[k] = addr(cn, k);
                int assign[12]; bool used_slot[12] = {false}, done[12] = {false};
                for (int step = 0; step < n; ++step) {                               // hardest corner first
                    int bk = -1, bbest = -1, bslot = 0;
                    for (int k = 0; k < n; ++k) {
                        if (done[k]) continue;
                        int best = 1 << 30, slot = 0;
                        for (int sl = 0; sl < n; ++sl) if (!used_slot[sl]) { const int v = cost(a[k], sl); if (v < best) { best = v; slot = sl; } }
                        if (best > bbest) { bbest = best; bk = k; bslot = slot; }
                    }
                    assign[bslot] = bk; used_slot[bslot] = true; done[bk] = true;
                }
                for (bool improved = true; improved;) {
                    improved = false;
                    for (int x = 0; x < n; ++x)
                        for (int y = x + 1; y < n; ++y)
                            if (cost(a[assign[y]], x) + cost(a[assign[x]], y) < cost(a[assign[x]], x) + cost(a[assign[y]], y)) { std::swap(assign[x], assign[y]); improved = true; }
                }
                for (int sl = 0; sl < n; ++sl) { perm[i][sl] = (int8_t)assign[sl]; put(a[assign[sl]], sl, +1); }
            };
            auto remove = [&](int i) {
                const CornerNode& cn = c.corners[order[i]];
                const int n = cn.nc <= 6 ? 6 : cn.nc <= 8 ? 8 : cn.nc <= 9 ? 9 : 12;
                for (int sl = 0; sl < n; ++sl) put(addr(cn, perm[i][sl]), sl, -1);
            };
            for (int i = i0; i < i1; ++i) place(i);
            for (int sweep = 0; sweep < 3; ++sweep)
                for (int i = i0; i < i1; ++i) { remove(i); place(i); }
        }
}

int upload_cascade(ugd_detector* d) {
    const Cascade& c = *d->casc;
    // Per-lane evaluator arrays: merged corner lists with byte offsets into the scan tile (two layouts).
    // Inside a stage of a stump cascade the nodes are ordered by corner count so that the 32 nodes a
    // warp evaluates together take the same branches (stage sums are order-independent).
    std::vector<int> order(c.nnodes());
    for (int i = 0; i < c.nnodes(); ++i) order[i] = i;
    if (c.max_tree_nodes == 1 && c.sum_exact)          // (a cascade whose sums are not exact keeps file order: see DevBufs::seq_sum)
        for (int s = 0; s < c.nstages(); ++s)
            std::stable_sort(order.begin() + c.stage_first_node[s], order.begin() + c.stage_first_node[s + 1],
                             [&](int a, int b) { return c.corners[a].nc < c.corners[b].nc; });
    const ScanGeom geom = make_scan_geom(c.win_w, c.win_h, c.has_tilted);
    const bool words = scan_word_offsets(c);
    const bool exact_stumps = scan_mode(c) == 1;
    std::vector<uint4> B(c.nnodes()), C(c.nnodes()), TA[2], TB[2];
    for (int si = 0; si < 2; ++si) { TA[si].resize(c.nnodes()); TB[si].resize(c.nnodes()); }
    std::vector<uint32_t> M(c.nnodes());
    std::vector<std::array<int8_t, 12>> perm;
    assign_lane_slots(c, geom, order, perm);
    for (int i = 0; i < c.nnodes(); ++i) {
        CornerNode cn = c.corners[order[i]];
        {
            const CornerNode src = cn;                // corners dealt to the evaluator's slots (see assign_lane_slots)
            for (int j = 0; j < 12; ++j) { cn.cx[j] = src.cx[perm[i][j]]; cn.cy[j] = src.cy[perm[i][j]]; cn.coef[j] = src.coef[perm[i][j]]; }
        }
        const HostNode& h = c.nodes[order[i]];
        uint32_t cf[12];
        for (int j = 0; j < 12; ++j) {
            cf[j] = (uint32_t)(uint8_t)cn.coef[j];
        }
        B[i] = make_uint4(0u, 0u,
                          cf[0] | cf[1] << 8 | cf[2] << 16 | cf[3] << 24, cf[4] | cf[5] << 8 | cf[6] << 16 | cf[7] << 24);
        for (int si = 0; si < 2; ++si) {              // the same lists with offsets into the scan tile
            uint32_t to[12];
            for (int j = 0; j < 12; ++j) {
                to[j] = (uint32_t)(scan_tile_off(geom, si, cn.cx[j], cn.cy[j]) + (cn.plane ? scan_plane_bytes(geom, si) : 0));
                if (words) to[j] >>= 2;
                if (to[j] > 0xffffu) { d->err = "scan tile too large for the per-lane offset tables"; return UGD_ERR_UNSUPPORTED; }
            }
            TA[si][i] = make_uint4(to[0] | to[1] << 16, to[2] | to[3] << 16, to[4] | to[5] << 16, to[6] | to[7] << 16);
            TB[si][i] = make_uint4(to[8] | to[9] << 16, to[10] | to[11] << 16, B[i].z, B[i].w);
        }
        uint32_t thr, lv, rv;
        memcpy(&thr, &h.thr, 4); memcpy(&lv, &h.left_val, 4); memcpy(&rv, &h.right_val, 4);
        if (exact_stumps) {                           // the group evaluator adds (left - right) where a node goes left: see add_if_lt
            const double d128 = ((double)h.left_val - (double)h.right_val) * 128.0;
            uint64_t u; memcpy(&u, &d128, 8);
            lv = (uint32_t)u; rv = (uint32_t)(u >> 32);
        }
        C[i] = make_uint4(cf[8] | cf[9] << 8 | cf[10] << 16 | cf[11] << 24, thr, lv, rv);
        M[i] = (uint32_t)cn.nc | (uint32_t)h.left << 8 | (uint32_t)h.right << 16;
    }
    int rc;
    if ((rc = upload(d, d->nodeC, C))) return rc;
    if ((rc = upload(d, d->nodeM, M))) return rc;
    for (int si = 0; si < 2; ++si) {
        if ((rc = upload(d, d->nodeTA[si], TA[si]))) return rc;
        if ((rc = upload(d, d->nodeTB[si], TB[si]))) return rc;
    }
    if ((rc = upload(d, d->st_first_tree, c.stage_first_tree))) return rc;
    if ((rc = upload(d, d->st_ntrees, c.stage_ntrees))) return rc;
    if ((rc = upload(d, d->st_thr, c.stage_thr))) return rc;
    if ((rc = upload(d, d->tree_first_node, c.tree_first_node))) return rc;
    if ((rc = upload(d, d->st_first_node, c.stage_first_node))) return rc;
    std::vector<double> qbase;
    stage_right_sums(c, qbase);
    if ((rc = upload(d, d->qbase, qbase))) return rc;
    CK(cudaStreamSynchronize(d->stream));
    return UGD_OK;
}

// Makes this detector's image the resident one (see ConstBank), launches the scan and records the last-use event.
int scan_with_constants(ugd_detector* d, const DevBufs& b, const WorkItem* wk, int nblocks, int nblocks_step2, cudaStream_t st, int set) {
    std::lock_guard<std::mutex> lk(g_const_mu);
    ConstBank& bk = g_bank[d->device & 63];
    if (bk.uid != d->casc_uid || bk.dense != d->dense_stages) {
        for (ugd_detector* u : bk.users)
            if (u != d)
                for (int i = 0; i < 2; ++i)
                    if (u->const_used[i]) CK(cudaStreamWaitEvent(st, u->ev_const_use[i], 0));
        bk.users.clear();
        CK(upload_scan_constants(d->const_img, st));
        bk.uid = d->casc_uid; bk.dense = d->dense_stages;
        CK(cudaEventRecord(d->ev_bank_upload, st));
        d->upload_wait[1 - set] = true; d->upload_wait[set] = false;
    } else if (d->upload_wait[set]) {                  // the image arrived through this detector's other scan stream
        CK(cudaStreamWaitEvent(st, d->ev_bank_upload, 0));
        d->upload_wait[set] = false;
    }
    static const bool serial = getenv("UGD_SCAN_SERIAL") != nullptr;
    launch_scan(b, wk, nblocks, nblocks_step2, st, serial ? nullptr : d->scan_side[set], d->ev_scan_fork[set], d->ev_scan_join[set]);
    CK(cudaEventRecord(d->ev_const_use[set], st));
    d->const_used[set] = true;
    if (std::find(bk.users.begin(), bk.users.end(), d) == bk.users.end()) bk.users.push_back(d);
    return UGD_OK;
}

int ensure_buffers(ugd_detector* d, const Plan& p, int nframes, bool need_bgr, int outcap, int total_frames, bool two_sets = false) {
    const size_t F = (size_t)nframes;
    const size_t px = round_up((size_t)p.w * p.h, 256);
    if (need_bgr) {
        CK(d->bgr.reserve(F * round_up((size_t)p.w * p.h * 3, 256) + 256));
        CK(d->gray_raw.reserve(F * px + 256));
        CK(d->gray_eq.reserve(F * px + 256));
        CK(d->hist.reserve(F * 256 * sizeof(uint32_t)));
    }
    CK(d->mask.reserve(F * px + 256));
    CK(d->gray_in.reserve(F * px + 256));
    CK(d->pyr.reserve(F * p.pyr_bytes + 256));
    CK(d->S.reserve(F * p.int_ints * 4 + 256));
    CK(d->SQ.reserve(F * p.int_ints * 4 + 256));
    if (d->casc->has_tilted) CK(d->T.reserve(F * p.int_ints * 4 + 256));
    if (two_sets) {
        CK(d->mask2.reserve(F * px + 256));
        CK(d->S2.reserve(F * p.int_ints * 4 + 256));
        CK(d->SQ2.reserve(F * p.int_ints * 4 + 256));
        if (d->casc->has_tilted) CK(d->T2.reserve(F * p.int_ints * 4 + 256));
        CK(d->raw2.reserve(F * d->rawcap * sizeof(ugd_rect)));
        CK(d->rawsum2.reserve(F * d->rawcap * sizeof(double)));
        CK(d->rawcount2.reserve(F * 4));
        CK(d->grp2.reserve(F * 11 * (size_t)d->rawcap * 4));
        CK(d->out2.reserve(F * (size_t)outcap * sizeof(ugd_rect)));
        CK(d->outcount2.reserve(F * 4));
        CK(d->counters2.reserve(F * kCtrPerFrame * 8));
    }
    CK(d->carryS.reserve(F * p.carry_ints * 4 + 256));
    CK(d->carryQ.reserve(F * p.carry_ints * 4 + 256));
    CK(d->raw.reserve(F * d->rawcap * sizeof(ugd_rect)));
    CK(d->rawsum.reserve(F * d->rawcap * sizeof(double)));
    CK(d->rawcount.reserve(F * 4));
    CK(d->grp.reserve(F * 11 * (size_t)d->rawcap * 4));
    CK(d->out.reserve(F * (size_t)outcap * sizeof(ugd_rect)));
    CK(d->outcount.reserve(F * 4));
    CK(d->counters.reserve(F * kCtrPerFrame * 8));
    if (d->want_map) CK(d->stage_map.reserve(p.pos_entries * 2 + 256));
    // pinned result staging for the whole call (groups are finished on the host after the last one is enqueued)
    const size_t need = (size_t)total_frames * (size_t)outcap;
    if (need > d->h_out_cap || (size_t)total_frames > d->h_frames_cap || !d->h_out) {
        if (d->h_out) cudaFreeHost(d->h_out);
        if (d->h_counts) cudaFreeHost(d->h_counts);
        if (d->h_ctr) cudaFreeHost(d->h_ctr);
        d->h_out = nullptr; d->h_counts = nullptr; d->h_ctr = nullptr;
        const size_t fr = std::max<size_t>({(size_t)total_frames, d->h_frames_cap, (size_t)d->max_batch});
        const size_t rects = std::max<size_t>({need, d->h_out_cap, (size_t)1});
        CK(cudaMallocHost((void**)&d->h_out, rects * sizeof(ugd_rect)));
        CK(cudaMallocHost((void**)&d->h_counts, fr * sizeof(int)));
        CK(cudaMallocHost((void**)&d->h_ctr, fr * kCtrPerFrame * 8));
        d->h_out_cap = rects; d->h_frames_cap = fr;
    }
    return UGD_OK;
}

void mark(ugd_detector* d, int k) { if (d->profiling) cudaEventRecord(d->ev[k], d->stream); }

// Host side of a finished group: results and counters from the pinned staging area (frames f0 .. f0+n of the call).
struct PlanTotals { int64_t levels, level_pixels, grid_windows; };
PlanTotals plan_totals(const ugd_detector* d, const Plan& p) {
    if (d->prm.legacy_24 && d->last_lg) return {(int64_t)d->last_lg->host.levels.size(), d->last_lg->host.level_pixels, d->last_lg->host.grid_windows};
    return {(int64_t)p.levels.size(), p.level_pixels, p.grid_windows};
}

void finish_group(ugd_detector* d, const Plan& p0, int f0, int nframes, ugd_rect* out, int cap, int* counts, bool* overflow) {
    const PlanTotals p = plan_totals(d, p0);
    for (int f = 0; f < nframes; ++f) {
        const int n = d->h_counts[f0 + f];
        counts[f] = n;
        if (n > cap) d->truncated = true;
        memcpy(out + (size_t)f * cap, d->h_out + (size_t)(f0 + f) * cap, (size_t)std::min(n, cap) * sizeof(ugd_rect));
        const unsigned long long* k = d->h_ctr + (size_t)(f0 + f) * kCtrPerFrame;
        d->ctr.windows_evaluated += (int64_t)k[CTR_EVALUATED];
        d->ctr.lowvar += (int64_t)k[CTR_LOWVAR];
        d->ctr.queued += (int64_t)k[CTR_QUEUED];
        d->ctr.raw += (int64_t)k[CTR_RAW];
        d->ctr.grouped += (int64_t)k[CTR_GROUPED];
        d->ctr.near_threshold += (int64_t)k[CTR_NEAR];
        d->ctr.weak_evals += (int64_t)k[CTR_WEAK];
        if (k[CTR_RAW_OVERFLOW]) *overflow = true;
    }
    d->ctr.frames += nframes;
    d->ctr.grid_windows += p.grid_windows * nframes;
    d->ctr.levels = p.levels;
    d->ctr.level_pixels = p.level_pixels;
}

// ---- TMA tensor maps for k_scan's tile staging ---------------------------------------------------------------------
// Per level and integral plane (S, and T for tilted cascades): a 3-D int32 tensor (row of P ints, sh + 1 rows, frame) over
// this buffer set; the box is the scan tile (scan_tma_box).  Grid-step-2 levels get the same map twice -- the kernel asks
// for the even and the odd columns with two boxes of element stride 2.
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn encode_tiled_fn() {
    static EncodeTiledFn fn = [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess) p = nullptr;
        return reinterpret_cast<EncodeTiledFn>(p);
    }();
    return fn;
}

int ensure_tmaps(ugd_detector* d, DevPlan* dp, int set, const int32_t* S, const int32_t* T, int frames_cap) {
    if (dp->tm_S[set] == S && dp->tm_T[set] == T && dp->tm_frames[set] == frames_cap) return UGD_OK;
    dp->tm_S[set] = S; dp->tm_T[set] = T; dp->tm_frames[set] = frames_cap; dp->tm_ok[set] = false;
    EncodeTiledFn enc = encode_tiled_fn();
    if (!enc) return UGD_OK;                           // no driver entry point: the kernels stage with LDGSTS
    const Plan& p = dp->host;
    const Cascade& c = *d->casc;
    const ScanGeom g = make_scan_geom(c.win_w, c.win_h, c.has_tilted);
    std::vector<CUtensorMap> maps(p.levels.size() * 4);
    memset(maps.data(), 0, maps.size() * sizeof(CUtensorMap));
    for (size_t l = 0; l < p.levels.size(); ++l) {
        const LevelInfo& L = p.levels[l];
        const int si = L.step - 1;
        uint32_t box[3], es[3];
        if (si && !d->deint2) continue;                // natural rows on a grid-step-2 level: no box (plain staging)
        if (!scan_tma_box(g, si, box, es)) return UGD_OK;
        for (int pl = 0; pl < (T ? 2 : 1); ++pl) {
            // grid step 2: rows are stored [even columns][odd columns] (deint2): one map per half-row
            for (int hf = 0; hf < (si ? 2 : 1); ++hf) {
                const int32_t* base = (pl ? T : S) + L.int_off + (si ? hf * (L.P / 2) : 0);
                const cuuint64_t dims[3] = {(cuuint64_t)(si ? L.P / 2 : L.P), (cuuint64_t)(L.sh + 1), (cuuint64_t)frames_cap};
                const cuuint64_t strides[2] = {(cuuint64_t)L.P * 4, (cuuint64_t)p.int_ints * 4};
                CUtensorMap m;
                if (enc(&m, CU_TENSOR_MAP_DATA_TYPE_INT32, 3, const_cast<int32_t*>(base), dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                        CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
                    return UGD_OK;                     // a shape the encoder refuses: plain staging
                maps[(l * 2 + pl) * 2 + hf] = m;
            }
        }
    }
    static_assert(sizeof(CUtensorMap) == 128, "tensor map size");
    CK(dp->tmaps[set].reserve(std::max<size_t>(maps.size() * sizeof(CUtensorMap), 128)));
    CK(cudaMemcpyAsync(dp->tmaps[set].p, maps.data(), maps.size() * sizeof(CUtensorMap), cudaMemcpyHostToDevice, d->stream));
    CK(cudaStreamSynchronize(d->stream));
    dp->tm_ok[set] = true;
    return UGD_OK;
}

int get_legacy_plan(ugd_detector* d, int w, int h, const ugd_detector::LegacyDev** out) {
    auto key = std::make_pair(w, h);
    auto it = d->lg_plans.find(key);
    if (it != d->lg_plans.end()) { *out = it->second.get(); return UGD_OK; }
    std::unique_ptr<ugd_detector::LegacyDev> lp(new ugd_detector::LegacyDev);
    const Cascade& c = *d->casc;
    build_legacy_plan(w, h, c.win_w, c.win_h, d->prm.scale_factor, d->prm.min_w, d->prm.min_h, d->prm.max_w, d->prm.max_h, lp->host);
    int rc;
    if ((rc = upload(d, lp->levels, lp->host.levels))) return rc;
    if ((rc = upload(d, lp->tab, lp->host.tab))) return rc;
    if ((rc = upload(d, lp->c0, lp->host.c0))) return rc;
    if (!d->lg_nodes.p) {                              // the hidden cascade of 2.4, once per detector
        LegacyCascadeHost lc;
        build_legacy_cascade(c, lc);
        if ((rc = upload(d, d->lg_nodes, lc.nodes))) return rc;
        if ((rc = upload(d, d->lg_thr24, lc.thr24))) return rc;
        if ((rc = upload(d, d->lg_child, lc.child))) return rc;
        if ((rc = upload(d, d->lg_parent, c.stage_parent))) return rc;
        if ((rc = upload(d, d->lg_next, c.stage_next))) return rc;
    }
    CK(cudaStreamSynchronize(d->stream));
    *out = lp.get();
    d->lg_plans[key] = std::move(lp);
    return UGD_OK;
}

// The scan of the OpenCV-2.4 compatibility mode in place of resize .. k_scan (everything on `st`).
int run_legacy_scan(ugd_detector* d, const DevBufs& b, int nframes, cudaStream_t st) {
    const ugd_detector::LegacyDev* lp = nullptr;
    int rc;
    if ((rc = get_legacy_plan(d, b.w, b.h, &lp))) return rc;
    const LegacyPlan& p = lp->host;
    const size_t F = (size_t)nframes;
    CK(d->lg_img.reserve(F * p.img_bytes + 256));
    CK(d->lg_S.reserve(F * p.int_entries * 4 + 256));
    CK(d->lg_SQ.reserve(F * p.int_entries * 8 + 256));
    if (d->want_map) CK(d->stage_map.reserve(std::max<size_t>(p.map_entries, 1) * 2 + 256));
    LegacyArgs a{};
    a.levels = lp->levels.as<LegacyLevel>(); a.tab = lp->tab.as<uint32_t>(); a.c0 = lp->c0.as<int16_t>();
    a.src = b.src_gray; a.src_stride = b.src_stride; a.src_pitch = b.src_pitch; a.w = b.w; a.h = b.h;
    a.img = d->lg_img.as<uint8_t>(); a.img_stride = p.img_bytes;
    a.S = d->lg_S.as<int32_t>(); a.SQ = d->lg_SQ.as<double>(); a.int_stride = p.int_entries;
    a.win_w = b.win_w; a.win_h = b.win_h; a.nstages = b.nstages;
    a.nodes = d->lg_nodes.as<LegacyNode>(); a.stage_first_tree = b.stage_first_tree; a.stage_ntrees = b.stage_ntrees;
    a.tree_first_node = b.tree_first_node;
    a.thr24 = d->lg_thr24.as<float>(); a.parent = d->lg_parent.as<int32_t>(); a.next = d->lg_next.as<int32_t>(); a.child = d->lg_child.as<int32_t>();
    a.raw = b.raw; a.rawsum = b.rawsum; a.rawcap = b.rawcap; a.rawcount = b.rawcount;
    a.stage_map = d->want_map ? d->stage_map.as<int16_t>() : nullptr; a.counters = b.counters;
    if (a.stage_map) CK(cudaMemsetAsync(a.stage_map, 0x80, std::max<size_t>(p.map_entries, 1) * 2, st));
    launch_legacy(a, p, nframes, st);
    d->last_lg = lp;
    return UGD_OK;
}

// One group of <= max_batch frames already resident on the device (frames f0.. of the call).  With `defer` the
// function only enqueues (kernels + result copies to pinned memory); the caller synchronises and calls finish_group.
// `set` >= 0 (deferred calls only): the group uses buffer set `set` and its image stage goes to img_stream, so that it
// overlaps the scan stage of the group before it (which works on the other set).
int run_group(ugd_detector* d, DevPlan* dp, const uint8_t* dev_frames, size_t frame_stride, bool is_bgr, int nframes,
              ugd_rect* out, int cap, int* counts, int f0, bool defer, int set = -1) {
    const bool ovl = set >= 0 && defer;
    const bool second = ovl && set == 1;
    const Plan& p = dp->host;
    const Cascade& c = *d->casc;
    int rc;
    DevBufs b{};
    b.w = p.w; b.h = p.h; b.nlevels = (int)p.levels.size(); b.nframes = nframes;
    b.win_w = c.win_w; b.win_h = c.win_h; b.nstages = c.nstages(); b.dense_stages = d->dense_stages;
    b.stumps_only = c.max_tree_nodes == 1;
    b.fast_scan = scan_mode(c); b.word_off = scan_word_offsets(c) ? 1 : 0;
    b.seq_sum = c.sum_exact ? 0 : 1;
    b.qbase = d->qbase.as<double>();
    { static const int kmin = getenv("UGD_SCAN_KMIN") ? atoi(getenv("UGD_SCAN_KMIN")) : 2; b.scan_kmin = std::max(kmin, 1); }   // 2: measured optimum (1: -2 %, 3: -3 %, 4: -8 %)
    b.levels = dp->levels.as<LevelInfo>(); b.xtab = dp->xtab.as<uint32_t>(); b.ytab = dp->ytab.as<uint32_t>();
    b.nodeC = d->nodeC.as<uint4>();
    b.nodeM = d->nodeM.as<uint32_t>();
    for (int si = 0; si < 2; ++si) { b.nodeTA[si] = d->nodeTA[si].as<uint4>(); b.nodeTB[si] = d->nodeTB[si].as<uint4>(); }
    b.stage_first_tree = d->st_first_tree.as<int32_t>(); b.stage_ntrees = d->st_ntrees.as<int32_t>();
    b.stage_thr = d->st_thr.as<float>(); b.tree_first_node = d->tree_first_node.as<int32_t>();
    b.stage_first_node = d->st_first_node.as<int32_t>();
    b.px_stride = round_up((size_t)p.w * p.h, 256);
    b.gray_raw = d->gray_raw.as<uint8_t>(); b.gray_eq = d->gray_eq.as<uint8_t>(); b.mask = (second ? d->mask2 : d->mask).as<uint8_t>();
    b.hist = d->hist.as<uint32_t>();
    memcpy(b.skin_lo, d->prm.skin_lower, 3); memcpy(b.skin_hi, d->prm.skin_upper, 3);
    b.use_skin = is_bgr ? 1 : 0;
    if (is_bgr) {
        b.bgr = dev_frames; b.bgr_stride = frame_stride;
        b.area2_src = d->area2_src; b.area2_row_stride = d->area2_row_stride; b.area2_frame_stride = d->area2_frame_stride;
        b.src_gray = b.gray_eq; b.src_stride = b.px_stride; b.src_pitch = p.w;
    } else {
        b.src_gray = dev_frames; b.src_stride = frame_stride; b.src_pitch = p.w;
    }
    b.pyr = d->pyr.as<uint8_t>(); b.pyr_stride = p.pyr_bytes;
    b.S = (second ? d->S2 : d->S).as<int32_t>(); b.SQ = (second ? d->SQ2 : d->SQ).as<int32_t>(); b.int_stride = p.int_ints;
    b.T = d->casc->has_tilted ? (second ? d->T2 : d->T).as<int32_t>() : nullptr;
    b.deint2 = d->deint2 ? 1 : 0;
    if (d->tma) {
        if ((rc = ensure_tmaps(d, dp, second ? 1 : 0, b.S, b.T, d->max_batch))) return rc;
        if (!dp->tm_ok[second ? 1 : 0]) { d->err = "tensor maps could not be encoded"; return UGD_ERR_CUDA; }
        b.tmaps = dp->tmaps[second ? 1 : 0].p;
    }
    b.carryS = d->carryS.as<uint32_t>(); b.carryQ = d->carryQ.as<uint32_t>(); b.carry_stride = p.carry_ints;
    b.pos_stride = p.pos_entries;
    b.tile_min = d->tile_min;
    b.raw = (second ? d->raw2 : d->raw).as<ugd_rect>(); b.rawsum = (second ? d->rawsum2 : d->rawsum).as<double>(); b.rawcap = d->rawcap;
    b.rawcount = (second ? d->rawcount2 : d->rawcount).as<int>();
    b.stage_map = d->want_map ? d->stage_map.as<int16_t>() : nullptr;
    b.grp_scratch = (second ? d->grp2 : d->grp).as<int>(); b.grp_stride = 11 * (size_t)d->rawcap;
    b.out = (second ? d->out2 : d->out).as<ugd_rect>(); b.outcap = cap; b.outcount = (second ? d->outcount2 : d->outcount).as<int>();
    const bool legacy = d->prm.legacy_24 != 0;
    b.min_neighbors = legacy && d->prm.min_neighbors != 0 ? std::max(d->prm.min_neighbors, 1) : d->prm.min_neighbors;
    b.no_clip = legacy ? 1 : 0; b.gray14 = legacy ? 1 : 0;
    b.counters = (second ? d->counters2 : d->counters).as<unsigned long long>();
    cudaStream_t st = second ? d->scan2 : d->stream;  // scan stage (and everything, without the group pipeline); set 1 has its own stream
    cudaStream_t si = ovl ? d->img_stream : st;       // image stage

    if (ovl && d->set_used[set]) CK(cudaStreamWaitEvent(si, d->ev_scan_done[set], 0));   // the set's last reader (group g-2) is done
    mark(d, 0);
    if (is_bgr) {
        CK(cudaMemsetAsync(b.hist, 0, (size_t)nframes * 256 * 4, si));
        launch_preprocess(b, si);
        if (d->area2_consumed) CK(cudaEventRecord(d->area2_consumed, si));
    }
    mark(d, 1);
    if (is_bgr) launch_equalize(b, si);
    mark(d, 2);
    if (!legacy) launch_resize_colsum(b, dp->wk_resize.as<WorkItem>(), (int)p.wk_resize.size(), si);    // pyramid + strip column sums in one kernel
    mark(d, 3);
    mark(d, 4);
    if (!legacy) launch_carry(b, dp->wk_carry.as<WorkItem>(), (int)p.wk_carry.size(), si);
    mark(d, 5);
    if (!legacy) launch_integral_bands(b, dp->wk_bands.as<WorkItem>(), (int)p.wk_bands.size(), si);
    mark(d, 6);
    if (b.T && !legacy) launch_tilted(b, p.max_level_w, si);
    mark(d, 7);
    if (ovl) {
        CK(cudaEventRecord(d->ev_img_done[set], si));
        CK(cudaStreamWaitEvent(st, d->ev_img_done[set], 0));
    }
    CK(cudaMemsetAsync(b.rawcount, 0, (size_t)nframes * 4, st));
    CK(cudaMemsetAsync(b.counters, 0, (size_t)nframes * kCtrPerFrame * 8, st));
    if (b.stage_map && !legacy) CK(cudaMemsetAsync(b.stage_map, 0x80, p.pos_entries * 2, st));
    if (legacy) {
        if ((rc = run_legacy_scan(d, b, nframes, st))) return rc;
    } else {
        // Too few blocks for their length -- a block that walks its rows from end to end makes a one-frame call last as long as
        // its longest block, and a group of small frames ends in a long tail -- so calls below 6000 row-walking blocks run one
        // block per (row band, TILE): the x-skip state at the left edge of a tile comes from a short backward pass over stage 0
        // (k_scan, chunked mode).  Measured, chunked against row-walking: one 1280x720 frame 0.44 / 0.81 ms per call, 32 such
        // frames 3.92 / 4.10 ms, 64 frames 640x360 2.10 / 2.35 ms; 64 frames 1280x720 (9216 blocks) 7.08 / 7.06, 1080p and 4K
        // groups 3 to 6 % slower chunked (the backward pass and the extra tile staging are not free).
        static const int chunk_below = getenv("UGD_CHUNK_BELOW") ? atoi(getenv("UGD_CHUNK_BELOW")) : 6000;
        // (two-node-tree cascades -- frontalface_alt2 -- are faster chunked at every size measured: 720p groups 6.68 k -> 6.89 k frames/s,
        //  1080p 2.94 k -> 3.07 k: their tiles take longer, and the tail of a launch of row-walking blocks with them)
        static const bool chunk_env = getenv("UGD_CHUNK_BELOW") != nullptr;
        const bool chunked = !dp->host_wk_scan_chunk.empty() &&
                             ((!chunk_env && b.fast_scan == 2) || (long long)nframes * (long long)p.wk_scan.size() < chunk_below);
        const std::vector<WorkItem>& wk = chunked ? dp->host_wk_scan_chunk : p.wk_scan;
        int n2 = 0;                                   // scan blocks of the grid-step-2 levels (they lead the table)
        while (n2 < (int)wk.size() && p.levels[wk[n2].level].step == 2) ++n2;
        b.scan_chunked = chunked ? 1 : 0;
        if ((rc = scan_with_constants(d, b, (chunked ? dp->wk_scan_chunk : dp->wk_scan).as<WorkItem>(), (int)wk.size(), n2, st, second ? 1 : 0))) return rc;
    }
    mark(d, 8);
    launch_group(b, st);
    mark(d, 9);
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(d->h_out + (size_t)f0 * cap, b.out, (size_t)nframes * cap * sizeof(ugd_rect), cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(d->h_counts + f0, b.outcount, (size_t)nframes * 4, cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(d->h_ctr + (size_t)f0 * kCtrPerFrame, b.counters, (size_t)nframes * kCtrPerFrame * 8, cudaMemcpyDeviceToHost, st));
    if (d->shard.call) {                              // this rank's message for the gather: rect slots, then counts
        uint8_t* msg = d->shard.send.as<uint8_t>();
        CK(cudaMemcpyAsync(msg + (size_t)f0 * cap * sizeof(ugd_rect), b.out, (size_t)nframes * cap * sizeof(ugd_rect), cudaMemcpyDeviceToDevice, st));
        CK(cudaMemcpyAsync(msg + (size_t)d->shard.slot_frames * cap * sizeof(ugd_rect) + (size_t)f0 * 4, b.outcount, (size_t)nframes * 4, cudaMemcpyDeviceToDevice, st));
    }
    d->last = b; d->last_plan = dp; d->last_was_bgr = is_bgr;
    if (ovl) { CK(cudaEventRecord(d->ev_scan_done[set], st)); d->set_used[set] = true; }
    if (defer) return UGD_OK;
    CK(cudaStreamSynchronize(st));
    bool overflow = false;
    finish_group(d, p, f0, nframes, out, cap, counts, &overflow);
    if (d->profiling) {
        const int launches[UGD_K_COUNT] = {is_bgr ? 1 : 0, is_bgr ? 1 : 0, 1, 0, 1, 1, b.T ? 1 : 0, 1, 1, 0};   // (k_colsum is part of k_resize since r02)   // k_scan: its two launches (grid step 2 / step 1 levels) are timed as one
        for (int k = 0; k < UGD_K_AREA; ++k) {
            float ms = 0;
            cudaEventElapsedTime(&ms, d->ev[k], d->ev[k + 1]);
            d->ktime[k] += ms; d->klaunch[k] += launches[k];
        }
    }
    d->last = b; d->last_plan = dp; d->last_was_bgr = is_bgr;
    if (overflow) { d->err = "per-frame candidate list overflowed (raise ugd_params.max_raw_per_frame)"; return UGD_ERR_CAPACITY; }
    return UGD_OK;
}

int truncation_status(ugd_detector* d) {
    if (!d->truncated) return UGD_OK;
    d->err = "a frame has more rectangles than cap: counts hold the true numbers, call again with a larger cap";
    return UGD_ERR_TRUNCATED;
}

int get_area_plan(ugd_detector* d, int sw, int sh, int dw, int dh, DevAreaPlan** out) {
    const std::vector<int> key{sw, sh, dw, dh};
    auto it = d->area_plans.find(key);
    if (it != d->area_plans.end()) { *out = it->second.get(); return UGD_OK; }
    std::unique_ptr<DevAreaPlan> ap(new DevAreaPlan);
    if (!build_area_plan(sw, sh, dw, dh, ap->host)) { d->err = "INTER_AREA path covers decimation only (scan size <= frame size)"; return UGD_ERR_UNSUPPORTED; }
    int rc;
    if ((rc = upload(d, ap->xfirst, ap->host.xfirst))) return rc;
    if ((rc = upload(d, ap->yfirst, ap->host.yfirst))) return rc;
    if ((rc = upload(d, ap->xtaps, ap->host.xtaps))) return rc;
    if ((rc = upload(d, ap->ytaps, ap->host.ytaps))) return rc;
    CK(cudaStreamSynchronize(d->stream));
    *out = ap.get();
    d->area_plans[key] = std::move(ap);
    return UGD_OK;
}

AreaArgs area_args(const DevAreaPlan& ap, const uint8_t* src, size_t row_stride, size_t frame_stride, uint8_t* dst, size_t dst_frame_stride) {
    AreaArgs a{};
    a.src = src; a.src_row_stride = row_stride; a.src_frame_stride = frame_stride;
    a.dst = dst; a.dst_frame_stride = dst_frame_stride;
    a.dw = ap.host.dw; a.dh = ap.host.dh; a.ix = ap.host.ix; a.iy = ap.host.iy; a.fast = ap.host.fast ? 1 : 0;
    a.xfirst = ap.xfirst.as<int32_t>(); a.yfirst = ap.yfirst.as<int32_t>();
    a.xtaps = ap.xtaps.as<AreaTap>(); a.ytaps = ap.ytaps.as<AreaTap>();
    return a;
}

int wait_pending(ugd_detector* d, ugd_rect* out, int* counts);

// async_submit: enqueue everything (copies, kernels, result copies to pinned memory) and return without waiting;
// ugd_wait_batch finishes the call.  out/counts are not touched until then.
int detect_body(ugd_detector* d, const void* frames, bool is_device, bool is_bgr, int nframes, int w, int h,
                size_t host_row_stride, ugd_rect* out, int cap, int* counts, bool async_submit);

// A failure in the middle of the group loop leaves enqueued work behind: drain it before the error goes up, so the
// caller may free or reuse its buffers.
void drain(ugd_detector* d) {
    if (d->copy_stream) cudaStreamSynchronize(d->copy_stream);
    if (d->img_stream) cudaStreamSynchronize(d->img_stream);
    if (d->scan2) cudaStreamSynchronize(d->scan2);
    if (d->stream) cudaStreamSynchronize(d->stream);
    d->pending.active = false;
}

// True when `p` is ordinary (pageable) host memory: not pinned, not registered, not managed.
bool is_pageable_host(const void* p) {
    cudaPointerAttributes a{};
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return true; }
    return a.type == cudaMemoryTypeUnregistered;
}

// memcpy with a few threads (a single thread moves ~10 GB/s; the H2D link takes 55)
void parallel_memcpy(uint8_t* dst, const uint8_t* src, size_t bytes) {
    static const int nthreads = [] { const char* e = getenv("UGD_COPY_THREADS"); int n = e ? atoi(e) : 4; return std::max(1, std::min(n, 16)); }();
    if (nthreads == 1 || bytes < (4u << 20)) { memcpy(dst, src, bytes); return; }
    const size_t chunk = (bytes / nthreads + 4095) & ~(size_t)4095;
    std::vector<std::thread> th;
    for (int t = 1; t < nthreads; ++t) {
        const size_t o = (size_t)t * chunk;
        if (o >= bytes) break;
        th.emplace_back([=] { memcpy(dst + o, src + o, std::min(chunk, bytes - o)); });
    }
    memcpy(dst, src, std::min(chunk, bytes));
    for (auto& t : th) t.join();
}

// One group of tight host frames -> device staging rows of `dpitch` bytes, on the copy stream.  Pageable frames go through
// pinned buffer `bi` of the detector (free again once ev_copied[bi], the copy of two groups ago, has completed).
int copy_group_h2d(ugd_detector* d, uint8_t* dst, size_t dpitch, const uint8_t* src, size_t tight, int nf, int bi, int g, bool pageable) {
    if (pageable) {
        const size_t need = (size_t)d->max_batch * tight;
        if (d->pin_bytes < need) {
            CK(cudaStreamSynchronize(d->copy_stream));
            for (int i = 0; i < 2; ++i) { if (d->pin[i]) cudaFreeHost(d->pin[i]); d->pin[i] = nullptr; }
            d->pin_bytes = 0;
            for (int i = 0; i < 2; ++i) CK(cudaHostAlloc((void**)&d->pin[i], need, cudaHostAllocDefault));
            d->pin_bytes = need;
        } else if (g >= 2) CK(cudaEventSynchronize(d->ev_copied[bi]));
        parallel_memcpy(d->pin[bi], src, (size_t)nf * tight);
        src = d->pin[bi];
    }
    CK(cudaMemcpy2DAsync(dst, dpitch, src, tight, tight, nf, cudaMemcpyHostToDevice, d->copy_stream));
    return UGD_OK;
}

int detect_common(ugd_detector* d, const void* frames, bool is_device, bool is_bgr, int nframes, int w, int h,
                  size_t host_row_stride, ugd_rect* out, int cap, int* counts, bool async_submit = false) {
    const int rc = detect_body(d, frames, is_device, is_bgr, nframes, w, h, host_row_stride, out, cap, counts, async_submit);
    if (rc != UGD_OK && rc != UGD_ERR_TRUNCATED && rc != UGD_ERR_ARG && d) drain(d);
    return rc;
}

int detect_body(ugd_detector* d, const void* frames, bool is_device, bool is_bgr, int nframes, int w, int h,
                size_t host_row_stride, ugd_rect* out, int cap, int* counts, bool async_submit) {
    if (!d || !frames || nframes < 0 || cap <= 0 || w <= 0 || h <= 0) return UGD_ERR_ARG;
    if (!async_submit && (!out || !counts)) return UGD_ERR_ARG;
    if (d->pending.active) { d->err = "a submitted batch is still pending: call ugd_wait_batch first"; return UGD_ERR_ARG; }
    if (w > d->max_w || h > d->max_h) { d->err = "frame larger than the detector was created for"; return UGD_ERR_CAPACITY; }
    CK(cudaSetDevice(d->device));
    DevPlan* dp = nullptr;
    int rc;
    if ((rc = get_plan(d, w, h, &dp))) return rc;
    d->ctr = ugd_counters{};
    d->truncated = false;
    memset(d->ktime, 0, sizeof d->ktime); memset(d->klaunch, 0, sizeof d->klaunch);
    const size_t ch = is_bgr ? 3 : 1;
    const size_t tight = (size_t)w * h * ch;
    const size_t st_stride = round_up(tight, 256);
    // Host frames in more than one group: copies of group g+1 (copy stream) overlap the kernels of group g.
    const bool tight_host = !is_device && host_row_stride == (size_t)w * ch;
    const bool pipelined = tight_host && !d->profiling && nframes > d->max_batch;
    // More than one group: the image stage of group g+1 overlaps the scan stage of group g (two buffer sets, two streams).
    static const bool no_overlap = getenv("UGD_NO_OVERLAP") != nullptr;
    const bool overlap = !no_overlap && !d->profiling && !d->want_map && !d->prm.legacy_24 && nframes > d->max_batch;
    const bool defer = pipelined || async_submit || overlap;
    static const bool no_pin_ring = getenv("UGD_NO_PIN_RING") != nullptr;
    const bool pageable = pipelined && !no_pin_ring && is_pageable_host(frames);
    if (pipelined) CK(d->stage2.reserve((size_t)d->max_batch * st_stride + 256));
    cudaStream_t in_stream = overlap ? d->img_stream : d->stream;      // the stream that consumes the input frames
    if (overlap) {
        d->set_used[0] = d->set_used[1] = false;
        CK(cudaEventRecord(d->ev_call, d->stream));                   // the image stream starts after whatever the caller queued before
        CK(cudaStreamWaitEvent(d->img_stream, d->ev_call, 0));
        CK(cudaStreamWaitEvent(d->scan2, d->ev_call, 0));
    }
    int g = 0;
    for (int f0 = 0, nf = 0; f0 < nframes; f0 += nf, ++g) {
        nf = std::min(d->max_batch, nframes - f0);
        // host frames: a short first group (a quarter) lets the kernels start after a quarter of the first copy
        if (pipelined && g == 0) nf = std::max(1, std::min(nf, d->max_batch / 4));
        if ((rc = ensure_buffers(d, dp->host, std::min(d->max_batch, nframes), is_bgr, cap, nframes, overlap))) return rc;
        const uint8_t* dev; size_t stride;
        uint8_t* staging = is_bgr ? d->bgr.as<uint8_t>() : d->gray_in.as<uint8_t>();
        if (pipelined && (g & 1)) staging = d->stage2.as<uint8_t>();
        if (is_device) {
            const uint8_t* src = static_cast<const uint8_t*>(frames) + (size_t)f0 * tight;
            if (is_bgr && (((uintptr_t)src & 3) || (tight & 3))) {   // K1 reads 32-bit words: realign
                CK(cudaMemcpy2DAsync(staging, st_stride, src, tight, tight, nf, cudaMemcpyDeviceToDevice, in_stream));
                dev = staging; stride = st_stride;
            } else { dev = src; stride = tight; }
        } else {
            const uint8_t* src = static_cast<const uint8_t*>(frames);
            if (pipelined) {
                const int bi = g & 1;
                if (g >= 2) CK(cudaStreamWaitEvent(d->copy_stream, d->ev_consumed[bi], 0));   // staging[bi] free again
                if ((rc = copy_group_h2d(d, staging, st_stride, src + (size_t)f0 * tight, tight, nf, bi, g, pageable))) return rc;
                CK(cudaEventRecord(d->ev_copied[bi], d->copy_stream));
                CK(cudaStreamWaitEvent(in_stream, d->ev_copied[bi], 0));
            } else if (tight_host) {
                CK(cudaMemcpy2DAsync(staging, st_stride, src + (size_t)f0 * tight, tight, tight, nf, cudaMemcpyHostToDevice, d->stream));
            } else {                                                   // single strided frame (cv::Mat with step)
                CK(cudaMemcpy2DAsync(staging, (size_t)w * ch, src, host_row_stride, (size_t)w * ch, h, cudaMemcpyHostToDevice, d->stream));
            }
            dev = staging; stride = st_stride;
        }
        if ((rc = run_group(d, dp, dev, stride, is_bgr, nf, defer ? nullptr : out + (size_t)f0 * cap, cap, defer ? nullptr : counts + f0, f0, defer,
                            overlap ? (g & 1) : -1))) return rc;
        if (pipelined) CK(cudaEventRecord(d->ev_consumed[g & 1], in_stream));      // the image stage was the last reader of the staging buffer
    }
    if (overlap && g > 1) {                            // the groups of set 1 ran on their own stream: the call ends when both have
        CK(cudaEventRecord(d->ev_set1, d->scan2));
        CK(cudaStreamWaitEvent(d->stream, d->ev_set1, 0));
    }
    if (defer) {
        d->pending.active = true; d->pending.nframes = nframes; d->pending.cap = cap; d->pending.plan = dp;
        if (!async_submit) return wait_pending(d, out, counts);
        return UGD_OK;
    }
    return truncation_status(d);
}

int wait_pending(ugd_detector* d, ugd_rect* out, int* counts) {
    if (!d->pending.active) { d->err = "no submitted batch to wait for"; return UGD_ERR_ARG; }
    d->pending.active = false;
    CK(cudaStreamSynchronize(d->stream));
    const int nframes = d->pending.nframes, cap = d->pending.cap;
    bool overflow = false;
    for (int f0 = 0; f0 < nframes; f0 += d->max_batch)
        finish_group(d, d->pending.plan->host, f0, std::min(d->max_batch, nframes - f0), out + (size_t)f0 * cap, cap, counts + f0, &overflow);
    if (overflow) { d->err = "per-frame candidate list overflowed (raise ugd_params.max_raw_per_frame)"; return UGD_ERR_CAPACITY; }
    return truncation_status(d);
}

// ---- NCCL, bound at run time --------------------------------------------------------------------
struct NcclApi {
    void* h = nullptr;
    decltype(&ncclGetUniqueId) GetUniqueId = nullptr;
    decltype(&ncclCommInitRank) CommInitRank = nullptr;
    decltype(&ncclCommDestroy) CommDestroy = nullptr;
    decltype(&ncclSend) Send = nullptr;
    decltype(&ncclRecv) Recv = nullptr;
    decltype(&ncclGroupStart) GroupStart = nullptr;
    decltype(&ncclGroupEnd) GroupEnd = nullptr;
    decltype(&ncclGetErrorString) GetErrorString = nullptr;
    decltype(&ncclGetVersion) GetVersion = nullptr;
    std::string err;
};

NcclApi* nccl_api() {
    static NcclApi api;
    static std::once_flag once;
    std::call_once(once, [] {
        // a process that already carries an NCCL (torch's bundled one) gets that same library back by soname
        const char* names[] = {getenv("UGD_NCCL_LIB"), "libnccl.so.2", "libnccl.so"};
        for (const char* n : names) {
            if (!n) continue;
            api.h = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
            if (api.h) break;
        }
        if (!api.h) { api.err = std::string("cannot load libnccl.so.2: ") + (dlerror() ? dlerror() : "?"); return; }
#define UGD_SYM(f) api.f = reinterpret_cast<decltype(api.f)>(dlsym(api.h, "nccl" #f)); if (!api.f) { api.err = "libnccl lacks nccl" #f; api.h = nullptr; return; }
        UGD_SYM(GetUniqueId) UGD_SYM(CommInitRank) UGD_SYM(CommDestroy) UGD_SYM(Send) UGD_SYM(Recv) UGD_SYM(GroupStart)
        UGD_SYM(GroupEnd) UGD_SYM(GetErrorString) UGD_SYM(GetVersion)
#undef UGD_SYM
    });
    return api.h ? &api : nullptr;
}

int fail_nccl(ugd_detector* d, ncclResult_t r, const char* what) {
    NcclApi* n = nccl_api();
    d->err = std::string(what) + ": " + (n ? n->GetErrorString(r) : "NCCL not loaded");
    return UGD_ERR_COMM;
}
#define NK(call)                                                        \
    do { ncclResult_t r_ = (call); if (r_ != ncclSuccess) return fail_nccl(d, r_, #call); } while (0)

void comm_release(ugd_detector* d) {
    if (d->shard.comm && d->shard.own) { NcclApi* n = nccl_api(); if (n) n->CommDestroy(d->shard.comm); }
    d->shard.comm = nullptr; d->shard.own = false; d->shard.rank = 0; d->shard.world = 1;
}

}  // namespace

// ---- C ABI ---------------------------------------------------------------------------------------
extern "C" {

int ugd_cascade_load_xml(const char* path, ugd_cascade** out) {
    if (!path || !out) return UGD_ERR_ARG;
    *out = nullptr;
    std::unique_ptr<ugd_cascade> c(new ugd_cascade);
    std::string err;
    int rc = load_cascade_xml(path, c->c, err);
    if (rc != UGD_OK) { fprintf(stderr, "ugdet: %s: %s\n", path, err.c_str()); return rc; }
    static std::atomic<uint64_t> next_uid{1};
    c->uid = next_uid.fetch_add(1);
    *out = c.release();
    return UGD_OK;
}

int ugd_cascade_get_info(const ugd_cascade* c, ugd_cascade_info* i) {
    if (!c || !i) return UGD_ERR_ARG;
    i->win_w = c->c.win_w; i->win_h = c->c.win_h; i->nstages = c->c.nstages(); i->ntrees = c->c.ntrees();
    i->nnodes = c->c.nnodes(); i->max_tree_nodes = c->c.max_tree_nodes; i->old_format = c->c.old_format;
    i->sum_exact = c->c.sum_exact; i->has_tilted = c->c.has_tilted;
    return UGD_OK;
}

int ugd_cascade_dump(const ugd_cascade* cc, int32_t* sft, int32_t* snt, float* sthr, int32_t* tfn, int32_t* tnn,
                     int32_t* nrects, float* nw, float* nthr, int32_t* nl, int32_t* nr, float* nlv, float* nrv, int32_t* ntl) {
    if (!cc) return UGD_ERR_ARG;
    const Cascade& c = cc->c;
    for (int s = 0; s < c.nstages(); ++s) {
        if (sft) sft[s] = c.stage_first_tree[s];
        if (snt) snt[s] = c.stage_ntrees[s];
        if (sthr) sthr[s] = c.stage_thr[s];
    }
    for (int t = 0; t < c.ntrees(); ++t) { if (tfn) tfn[t] = c.tree_first_node[t]; if (tnn) tnn[t] = c.tree_nnodes[t]; }
    for (int n = 0; n < c.nnodes(); ++n) {
        const HostNode& h = c.nodes[n];
        if (nrects) memcpy(nrects + 12 * n, h.rect, 48);
        if (nw) memcpy(nw + 3 * n, h.weight, 12);
        if (nthr) nthr[n] = h.thr;
        if (nl) nl[n] = h.left;
        if (nr) nr[n] = h.right;
        if (nlv) nlv[n] = h.left_val;
        if (nrv) nrv[n] = h.right_val;
        if (ntl) ntl[n] = h.tilted;
    }
    return UGD_OK;
}

void ugd_cascade_free(ugd_cascade* c) { delete c; }

int ugd_create(const ugd_cascade* c, const ugd_params* p, int cuda_device, int max_w, int max_h, int max_batch,
               ugd_detector** out) {
    if (!c || !out || max_w <= 0 || max_h <= 0 || max_batch <= 0) return UGD_ERR_ARG;
    *out = nullptr;
    if (max_w > 4095 + c->c.win_w || max_h > 4095 + c->c.win_h) return UGD_ERR_CAPACITY;   // window ids and tile offsets are sized for frames up to 4K
    std::unique_ptr<ugd_detector> up(new ugd_detector);
    ugd_detector* d = up.get();
    d->casc = &c->c; d->casc_uid = c->uid;
    if (p) d->prm = *p; else ugd_params_default(&d->prm);
    if (!(d->prm.scale_factor > 1.0)) return UGD_ERR_ARG;
    d->device = cuda_device; d->max_w = max_w; d->max_h = max_h; d->max_batch = max_batch;
    d->rawcap = d->prm.max_raw_per_frame > 0 ? (int)round_up((size_t)d->prm.max_raw_per_frame, 4) : 8192;
    // in-tile stages: default 8 (measured optimum for frontalface_alt at 720p), bounded by the __constant__ node budget
    int D = d->prm.dense_stages > 0 ? d->prm.dense_stages : 9;     // r02: 9 measured best on alt / alt_tree, equal to 8..10 on alt2 / default
    if (const char* e = getenv("UGD_TILE_MIN")) d->tile_min = atoi(e);
    D = std::min(D, std::min(d->casc->nstages(), kMaxDenseStages));
    while (D > 1 && d->casc->stage_first_node[D] > kMaxDenseNodes) --D;
    if (d->casc->stage_first_node[D] > kMaxDenseNodes) return UGD_ERR_UNSUPPORTED;
    d->dense_stages = D; d->dense_nodes = d->casc->stage_first_node[D];
    cudaError_t e = cudaSetDevice(cuda_device);
    if (e != cudaSuccess) { fprintf(stderr, "ugdet: cudaSetDevice(%d): %s\n", cuda_device, cudaGetErrorString(e)); return UGD_ERR_CUDA; }
    if ((e = cudaStreamCreateWithFlags(&d->own_stream, cudaStreamNonBlocking)) != cudaSuccess) {
        fprintf(stderr, "ugdet: cudaStreamCreate: %s\n", cudaGetErrorString(e)); return UGD_ERR_CUDA;
    }
    d->stream = d->own_stream;
    cudaStreamCreateWithFlags(&d->copy_stream, cudaStreamNonBlocking);
    {
        // UGD_IMG_PRIO=low|normal|high (default low): low lets the image stage of the next group fill the tail of the running scan launches
        int lo = 0, hi = 0;
        cudaDeviceGetStreamPriorityRange(&lo, &hi);
        const char* pr = getenv("UGD_IMG_PRIO");
        const int prio = pr && !strcmp(pr, "high") ? hi : pr && !strcmp(pr, "normal") ? 0 : lo;      // measured (alt, 720p, 4 groups): low +2.1 %, normal +1.8 %, high +0.2 %
        cudaStreamCreateWithPriority(&d->img_stream, cudaStreamNonBlocking, prio);
        for (int i = 0; i < 2; ++i) { cudaEventCreateWithFlags(&d->ev_img_done[i], cudaEventDisableTiming); cudaEventCreateWithFlags(&d->ev_scan_done[i], cudaEventDisableTiming); }
        cudaEventCreateWithFlags(&d->ev_call, cudaEventDisableTiming);
    }
    for (int i = 0; i < 2; ++i) { cudaEventCreateWithFlags(&d->ev_copied[i], cudaEventDisableTiming); cudaEventCreateWithFlags(&d->ev_consumed[i], cudaEventDisableTiming); }
    cudaDeviceGetAttribute(&d->sm_count, cudaDevAttrMultiProcessorCount, cuda_device);
    for (auto& ev : d->ev) cudaEventCreate(&ev);
    for (auto& ev : d->ev_area) cudaEventCreate(&ev);
    cudaStreamCreateWithFlags(&d->scan2, cudaStreamNonBlocking);
    cudaEventCreateWithFlags(&d->ev_set1, cudaEventDisableTiming);
    cudaEventCreateWithFlags(&d->ev_bank_upload, cudaEventDisableTiming);
    for (int i = 0; i < 2; ++i) {
        cudaEventCreateWithFlags(&d->ev_const_use[i], cudaEventDisableTiming);
        cudaStreamCreateWithFlags(&d->scan_side[i], cudaStreamNonBlocking);
        cudaEventCreateWithFlags(&d->ev_scan_fork[i], cudaEventDisableTiming);
        cudaEventCreateWithFlags(&d->ev_scan_join[i], cudaEventDisableTiming);
    }
    if ((e = init_kernel_attributes(cuda_device, d->casc->win_w, d->casc->win_h, d->casc->has_tilted)) != cudaSuccess) {
        fprintf(stderr, "ugdet: kernel attributes: %s\n", cudaGetErrorString(e)); ugd_destroy(up.release()); return UGD_ERR_CUDA;
    }
    {   // both tile layouts through TMA <=> the encoder exists, both boxes fit, one plane per tile (the tilted plane keeps the natural layout)
        const ScanGeom g = make_scan_geom(d->casc->win_w, d->casc->win_h, d->casc->has_tilted);
        uint32_t box[3], es[3];
        d->tma = !d->prm.legacy_24 && encode_tiled_fn() && scan_tma_box(g, 0, box, es);
        d->deint2 = d->tma && !d->casc->has_tilted && scan_tma_box(g, 1, box, es);
    }
    if (build_scan_constants(*d->casc, d->dense_stages, d->const_img) != cudaSuccess) {
        fprintf(stderr, "ugdet: a dense-stage node has a corner count the scan evaluators do not cover\n"); ugd_destroy(up.release()); return UGD_ERR_UNSUPPORTED;
    }
    int rc = upload_cascade(d);
    if (rc != UGD_OK) { fprintf(stderr, "ugdet: %s\n", d->err.c_str()); ugd_destroy(up.release()); return rc; }
    *out = up.release();
    return UGD_OK;
}

void ugd_destroy(ugd_detector* d) {
    if (!d) return;
    cudaSetDevice(d->device);
    if (d->copy_stream) cudaStreamSynchronize(d->copy_stream);
    for (int i = 0; i < 2; ++i) if (d->pin[i]) cudaFreeHost(d->pin[i]);
    if (d->stream) cudaStreamSynchronize(d->stream);
    if (d->img_stream) cudaStreamSynchronize(d->img_stream);
    if (d->scan2) cudaStreamSynchronize(d->scan2);
    {   // the streams are idle: nothing of this detector reads the constant bank any more
        std::lock_guard<std::mutex> lk(g_const_mu);
        auto& us = g_bank[d->device & 63].users;
        us.erase(std::remove(us.begin(), us.end(), d), us.end());
    }
    for (int i = 0; i < 2; ++i) {
        if (d->ev_const_use[i]) cudaEventDestroy(d->ev_const_use[i]);
        if (d->scan_side[i]) { cudaStreamSynchronize(d->scan_side[i]); cudaStreamDestroy(d->scan_side[i]); }
        if (d->ev_scan_fork[i]) cudaEventDestroy(d->ev_scan_fork[i]);
        if (d->ev_scan_join[i]) cudaEventDestroy(d->ev_scan_join[i]);
    }
    if (d->scan2) { cudaStreamSynchronize(d->scan2); cudaStreamDestroy(d->scan2); }
    if (d->ev_set1) cudaEventDestroy(d->ev_set1);
    if (d->ev_bank_upload) cudaEventDestroy(d->ev_bank_upload);
    d->raw2.release(); d->rawsum2.release(); d->rawcount2.release(); d->grp2.release(); d->out2.release(); d->outcount2.release(); d->counters2.release();
    comm_release(d);
    d->shard.send.release(); d->shard.recv.release(); d->shard.packed.release();
    if (d->shard.h_packed) cudaFreeHost(d->shard.h_packed);
    for (auto& kv : d->plans) kv.second->release();
    for (auto& kv : d->area_plans) kv.second->release();
    d->src_full.release(); d->area_dst.release();
    for (auto& kv : d->lg_plans) { kv.second->levels.release(); kv.second->tab.release(); kv.second->c0.release(); }
    DevPtr* all[] = {&d->lg_nodes, &d->lg_thr24, &d->lg_parent, &d->lg_next, &d->lg_child, &d->lg_img, &d->lg_S, &d->lg_SQ, &d->qbase, &d->nodeC, &d->nodeM, &d->nodeTA[0], &d->nodeTA[1], &d->nodeTB[0], &d->nodeTB[1], &d->st_first_tree, &d->st_ntrees, &d->st_thr, &d->tree_first_node,
                     &d->st_first_node, &d->bgr, &d->gray_raw, &d->gray_eq, &d->mask, &d->hist, &d->pyr, &d->S, &d->SQ, &d->T,
                     &d->carryS, &d->carryQ, &d->raw, &d->rawsum, &d->rawcount, &d->grp,
                     &d->out, &d->outcount, &d->counters, &d->stage_map, &d->gray_in, &d->S2, &d->SQ2, &d->T2, &d->mask2};
    for (DevPtr* p : all) p->release();
    if (d->h_out) cudaFreeHost(d->h_out);
    if (d->h_counts) cudaFreeHost(d->h_counts);
    if (d->h_ctr) cudaFreeHost(d->h_ctr);
    for (auto& ev : d->ev) if (ev) cudaEventDestroy(ev);
    for (auto& ev : d->ev_area) if (ev) cudaEventDestroy(ev);
    for (int i = 0; i < 2; ++i) { if (d->ev_copied[i]) cudaEventDestroy(d->ev_copied[i]); if (d->ev_consumed[i]) cudaEventDestroy(d->ev_consumed[i]); }
    if (d->copy_stream) cudaStreamDestroy(d->copy_stream);
    if (d->img_stream) cudaStreamDestroy(d->img_stream);
    for (int i = 0; i < 2; ++i) { if (d->ev_img_done[i]) cudaEventDestroy(d->ev_img_done[i]); if (d->ev_scan_done[i]) cudaEventDestroy(d->ev_scan_done[i]); }
    if (d->ev_call) cudaEventDestroy(d->ev_call);
    d->stage2.release(); d->crop_items.release(); d->crop_out.release();
    if (d->own_stream) cudaStreamDestroy(d->own_stream);
    delete d;
}

int ugd_set_stream(ugd_detector* d, void* s) {
    if (!d) return UGD_ERR_ARG;
    if (d->stream) cudaStreamSynchronize(d->stream);
    d->stream = s ? static_cast<cudaStream_t>(s) : d->own_stream;
    return UGD_OK;
}

int ugd_detect_bgr(ugd_detector* d, const uint8_t* bgr, int w, int h, size_t stride, ugd_rect* out, int cap, int* n) {
    if (!n) return UGD_ERR_ARG;
    if (stride == 0) stride = (size_t)w * 3;
    if (stride < (size_t)w * 3) return UGD_ERR_ARG;
    return detect_common(d, bgr, false, true, 1, w, h, stride, out, cap, n);
}

int ugd_detect_gray(ugd_detector* d, const uint8_t* gray, int w, int h, size_t stride, ugd_rect* out, int cap, int* n) {
    if (!n) return UGD_ERR_ARG;
    if (stride == 0) stride = (size_t)w;
    if (stride < (size_t)w) return UGD_ERR_ARG;
    return detect_common(d, gray, false, false, 1, w, h, stride, out, cap, n);
}

int ugd_detect_batch(ugd_detector* d, const void* frames, int is_device, int is_bgr, int nframes, int w, int h,
                     ugd_rect* out, int cap, int* counts) {
    return detect_common(d, frames, is_device != 0, is_bgr != 0, nframes, w, h, (size_t)w * (is_bgr ? 3 : 1), out, cap, counts);
}

int ugd_resize_area(ugd_detector* d, const uint8_t* src, int sw, int sh, int channels, size_t stride, uint8_t* dst, int dw, int dh) {
    if (!d || !src || !dst || (channels != 1 && channels != 3) || sw <= 0 || sh <= 0) return UGD_ERR_ARG;
    if (stride == 0) stride = (size_t)sw * channels;
    if (stride < (size_t)sw * channels) return UGD_ERR_ARG;
    CK(cudaSetDevice(d->device));
    DevAreaPlan* ap = nullptr;
    int rc;
    if ((rc = get_area_plan(d, sw, sh, dw, dh, &ap))) return rc;
    const size_t sbytes = (size_t)sw * channels * sh, dbytes = (size_t)dw * channels * dh;
    CK(d->src_full.reserve(sbytes + 256));
    CK(d->area_dst.reserve(dbytes + 256));
    CK(cudaMemcpy2DAsync(d->src_full.p, (size_t)sw * channels, src, stride, (size_t)sw * channels, sh, cudaMemcpyHostToDevice, d->stream));
    launch_area(area_args(*ap, d->src_full.as<uint8_t>(), (size_t)sw * channels, sbytes, d->area_dst.as<uint8_t>(), dbytes), channels, 1, d->stream);
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(dst, d->area_dst.p, dbytes, cudaMemcpyDeviceToHost, d->stream));
    CK(cudaStreamSynchronize(d->stream));
    return UGD_OK;
}

int ugd_crop_faces(ugd_detector* d, const void* dev_frames, int nframes, int w, int h, int channels, const ugd_rect* rects,
                   const int* counts, int cap, uint8_t* out, size_t out_cap, size_t* offsets, size_t* total) {
    if (!d || !dev_frames || !rects || !counts || !offsets || !total || nframes < 0 || cap <= 0 || w <= 0 || h <= 0 ||
        (channels != 1 && channels != 3))
        return UGD_ERR_ARG;
    std::vector<CropItem> items;
    size_t bytes = 0;
    for (int f = 0; f < nframes; ++f) {
        const int n = std::min(counts[f], cap);
        for (int i = 0; i < n; ++i) {
            const ugd_rect& r = rects[(size_t)f * cap + i];
            if (r.x < 0 || r.y < 0 || r.width <= 0 || r.height <= 0 || r.x + r.width > w || r.y + r.height > h) {
                d->err = "ugd_crop_faces: a rectangle reaches outside the frame";     // (detect calls return clipped rects)
                return UGD_ERR_ARG;
            }
            offsets[(size_t)f * cap + i] = bytes;
            items.push_back({f, r.x, r.y, r.width, r.height, 0, (uint64_t)bytes});
            bytes += (size_t)r.width * r.height * channels;
        }
    }
    *total = bytes;
    if (bytes == 0) return UGD_OK;
    if (!out || out_cap < bytes) { d->err = "ugd_crop_faces: output buffer smaller than *total"; return UGD_ERR_CAPACITY; }
    CK(cudaSetDevice(d->device));
    CK(d->crop_items.reserve(items.size() * sizeof(CropItem)));
    CK(d->crop_out.reserve(bytes));
    CK(cudaMemcpyAsync(d->crop_items.p, items.data(), items.size() * sizeof(CropItem), cudaMemcpyHostToDevice, d->stream));
    launch_crop(static_cast<const uint8_t*>(dev_frames), (size_t)w * h * channels, (size_t)w * channels, channels,
                d->crop_items.as<CropItem>(), (int)items.size(), d->crop_out.as<uint8_t>(), d->stream);
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(out, d->crop_out.p, bytes, cudaMemcpyDeviceToHost, d->stream));
    CK(cudaStreamSynchronize(d->stream));          // `items` is a local: the copies above must be done before it goes
    return UGD_OK;
}

int ugd_detect_batch_scaled(ugd_detector* d, const void* frames, int is_device, int nframes, int src_w, int src_h, int scan_w,
                            int scan_h, ugd_rect* out, int cap, int* counts) {
    if (!d || !frames || !out || !counts || nframes < 0 || cap <= 0 || src_w <= 0 || src_h <= 0) return UGD_ERR_ARG;
    if (scan_w <= 0 || scan_h <= 0) { scan_w = src_w; scan_h = src_h; }           // scan.frame_size absent: structure.cc:64-65
    if (scan_w == src_w && scan_h == src_h) return ugd_detect_batch(d, frames, is_device, 1, nframes, src_w, src_h, out, cap, counts);
    if (d->pending.active) { d->err = "a submitted batch is still pending: call ugd_wait_batch first"; return UGD_ERR_ARG; }
    if (scan_w > d->max_w || scan_h > d->max_h) { d->err = "scan size larger than the detector was created for"; return UGD_ERR_CAPACITY; }
    CK(cudaSetDevice(d->device));
    DevAreaPlan* ap = nullptr;
    DevPlan* dp = nullptr;
    int rc;
    if ((rc = get_area_plan(d, src_w, src_h, scan_w, scan_h, &ap))) return rc;
    if ((rc = get_plan(d, scan_w, scan_h, &dp))) return rc;
    d->ctr = ugd_counters{};
    d->truncated = false;
    memset(d->ktime, 0, sizeof d->ktime); memset(d->klaunch, 0, sizeof d->klaunch);
    const size_t stight = (size_t)src_w * src_h * 3, dstride = round_up((size_t)scan_w * scan_h * 3, 256);
    // host frames in more than one group: the H2D copy of group g+1 overlaps the kernels of group g (two input buffers)
    const bool pipelined = !is_device && !d->profiling && nframes > d->max_batch;
    const size_t in_stride = round_up((size_t)d->max_batch * stight, 256);
    static const bool no_pin_ring = getenv("UGD_NO_PIN_RING") != nullptr;
    const bool pageable = pipelined && !no_pin_ring && is_pageable_host(frames);
    if (!is_device) CK(d->src_full.reserve((pipelined ? 2 : 1) * in_stride + 256));
    int g = 0;
    for (int f0 = 0; f0 < nframes; f0 += d->max_batch, ++g) {
        const int nf = std::min(d->max_batch, nframes - f0);
        if ((rc = ensure_buffers(d, dp->host, nf, true, cap, nframes))) return rc;
        const uint8_t* src = static_cast<const uint8_t*>(frames) + (size_t)f0 * stight;
        if (!is_device) {
            uint8_t* in = d->src_full.as<uint8_t>() + (pipelined && (g & 1) ? in_stride : 0);
            if (pipelined) {
                const int bi = g & 1;
                if (g >= 2) CK(cudaStreamWaitEvent(d->copy_stream, d->ev_consumed[bi], 0));
                if ((rc = copy_group_h2d(d, in, stight, src, stight, nf, bi, g, pageable))) return rc;
                CK(cudaEventRecord(d->ev_copied[bi], d->copy_stream));
                CK(cudaStreamWaitEvent(d->stream, d->ev_copied[bi], 0));
            } else {
                CK(cudaMemcpyAsync(in, src, (size_t)nf * stight, cudaMemcpyHostToDevice, d->stream));
            }
            src = in;
        }
        // A 2x2 decimation (the shipped cfg: 1280x720 -> 640x360) is done inside k_preprocess: no k_area launch, the scan-size
        // BGR frame is never written
        static const bool no_fuse = getenv("UGD_NO_AREA_FUSE") != nullptr;
        const bool fuse = !no_fuse && ap->host.fast && ap->host.ix == 2 && ap->host.iy == 2 &&
                          preprocess_can_fuse_area2(src, (size_t)src_w * 3, stight, scan_w);
        if (fuse) {
            d->area2_src = src; d->area2_row_stride = (size_t)src_w * 3; d->area2_frame_stride = stight;
            d->area2_consumed = pipelined ? d->ev_consumed[g & 1] : nullptr;
        } else {
            if (d->profiling) cudaEventRecord(d->ev_area[0], d->stream);
            launch_area(area_args(*ap, src, (size_t)src_w * 3, stight, d->bgr.as<uint8_t>(), dstride), 3, nf, d->stream);
            if (d->profiling) cudaEventRecord(d->ev_area[1], d->stream);
            if (pipelined) CK(cudaEventRecord(d->ev_consumed[g & 1], d->stream));  // the input buffer is free once k_area has run
        }
        rc = run_group(d, dp, d->bgr.as<uint8_t>(), dstride, true, nf, out + (size_t)f0 * cap, cap, counts + f0, f0, pipelined);
        d->area2_src = nullptr; d->area2_consumed = nullptr;
        if (rc) return rc;
        if (fuse) continue;
        if (d->profiling) {                            // run_group synchronised the stream
            float ms = 0;
            cudaEventElapsedTime(&ms, d->ev_area[0], d->ev_area[1]);
            d->ktime[UGD_K_AREA] += ms; d->klaunch[UGD_K_AREA] += 1;
        }
    }
    if (pipelined) {
        CK(cudaStreamSynchronize(d->stream));
        bool overflow = false;
        for (int f0 = 0; f0 < nframes; f0 += d->max_batch)
            finish_group(d, dp->host, f0, std::min(d->max_batch, nframes - f0), out + (size_t)f0 * cap, cap, counts + f0, &overflow);
        if (overflow) { d->err = "per-frame candidate list overflowed (raise ugd_params.max_raw_per_frame)"; return UGD_ERR_CAPACITY; }
    }
    return truncation_status(d);
}

int ugd_submit_batch(ugd_detector* d, const void* frames, int is_device, int is_bgr, int nframes, int w, int h, int cap) {
    return detect_common(d, frames, is_device != 0, is_bgr != 0, nframes, w, h, (size_t)w * (is_bgr ? 3 : 1), nullptr, cap, nullptr, true);
}

int ugd_wait_batch(ugd_detector* d, ugd_rect* out, int* counts) {
    if (!d || !out || !counts) return UGD_ERR_ARG;
    CK(cudaSetDevice(d->device));
    return wait_pending(d, out, counts);
}

int ugd_get_counters(const ugd_detector* d, ugd_counters* c) {
    if (!d || !c) return UGD_ERR_ARG;
    *c = d->ctr;
    return UGD_OK;
}

int ugd_set_profiling(ugd_detector* d, int on) { if (!d) return UGD_ERR_ARG; d->profiling = on != 0; return UGD_OK; }

int ugd_get_kernel_times(const ugd_detector* d, float* t, int64_t* l) {
    if (!d) return UGD_ERR_ARG;
    for (int k = 0; k < UGD_K_COUNT; ++k) { if (t) t[k] = d->ktime[k]; if (l) l[k] = d->klaunch[k]; }
    return UGD_OK;
}

const char* ugd_kernel_name(int k) {
    static const char* names[UGD_K_COUNT] = {"k_preprocess", "k_equalize", "k_resize", "k_colsum", "k_carry",
                                             "k_integral", "k_tilted", "k_scan", "k_group", "k_area"};
    return k >= 0 && k < UGD_K_COUNT ? names[k] : "?";
}

int ugd_plan_info(ugd_detector* d, int w, int h, int* nlevels, float* scales, int32_t* lw, int32_t* lh, int32_t* step, int cap) {
    if (!d || w <= 0 || h <= 0) return UGD_ERR_ARG;
    Plan p;
    const Cascade& c = *d->casc;
    build_plan(w, h, c.win_w, c.win_h, d->prm.scale_factor, d->prm.min_w, d->prm.min_h, d->prm.max_w, d->prm.max_h, p);
    if (nlevels) *nlevels = (int)p.levels.size();
    for (int i = 0; i < (int)p.levels.size() && i < cap; ++i) {
        if (scales) scales[i] = p.levels[i].scale;
        if (lw) lw[i] = p.levels[i].sw;
        if (lh) lh[i] = p.levels[i].sh;
        if (step) step[i] = p.levels[i].step;
    }
    return UGD_OK;
}

int ugd_plan_query(const ugd_cascade* c, const ugd_params* prm, int w, int h, int* nlevels, float* scales, int32_t* lw,
                   int32_t* lh, int32_t* step, int cap, int64_t* grid_windows, int64_t* level_pixels) {
    if (!c || w <= 0 || h <= 0) return UGD_ERR_ARG;
    ugd_params dp;
    if (!prm) { ugd_params_default(&dp); prm = &dp; }
    Plan p;
    build_plan(w, h, c->c.win_w, c->c.win_h, prm->scale_factor, prm->min_w, prm->min_h, prm->max_w, prm->max_h, p);
    if (nlevels) *nlevels = (int)p.levels.size();
    if (grid_windows) *grid_windows = p.grid_windows;
    if (level_pixels) *level_pixels = p.level_pixels;
    for (int i = 0; i < (int)p.levels.size() && i < cap; ++i) {
        if (scales) scales[i] = p.levels[i].scale;
        if (lw) lw[i] = p.levels[i].sw;
        if (lh) lh[i] = p.levels[i].sh;
        if (step) step[i] = p.levels[i].step;
    }
    return UGD_OK;
}

int ugd_get_raw(ugd_detector* d, ugd_rect* out, double* sums, int cap, int* n) {
    if (!d || !n || !d->last_plan) return UGD_ERR_ARG;
    CK(cudaSetDevice(d->device));
    int cnt = 0;
    CK(cudaMemcpy(&cnt, d->last.rawcount, 4, cudaMemcpyDeviceToHost));
    *n = cnt;
    int m = std::min(std::min(cnt, cap), d->last.rawcap);
    if (out && m > 0) CK(cudaMemcpy(out, d->last.raw, (size_t)m * sizeof(ugd_rect), cudaMemcpyDeviceToHost));
    if (sums && m > 0) CK(cudaMemcpy(sums, d->last.rawsum, (size_t)m * 8, cudaMemcpyDeviceToHost));
    return UGD_OK;
}

int ugd_set_stage_map(ugd_detector* d, int on) { if (!d) return UGD_ERR_ARG; d->want_map = on != 0; return UGD_OK; }

int ugd_get_stage_map(ugd_detector* d, int16_t* map, size_t cap, size_t* n) {
    if (!d || !n || !d->last_plan || !d->last.stage_map) return UGD_ERR_ARG;
    CK(cudaSetDevice(d->device));
    const Plan& p = d->last_plan->host;
    if (d->prm.legacy_24 && d->last_lg) {              // the legacy scan writes the map in scan order already
        *n = d->last_lg->host.map_entries;
        if (map) CK(cudaMemcpy(map, d->stage_map.p, std::min(cap, *n) * 2, cudaMemcpyDeviceToHost));
        return UGD_OK;
    }
    *n = p.map_entries;
    if (!map) return UGD_OK;
    std::vector<int16_t> padded(p.pos_entries);
    CK(cudaMemcpy(padded.data(), d->last.stage_map, p.pos_entries * 2, cudaMemcpyDeviceToHost));
    for (const LevelInfo& L : p.levels)
        for (int r = 0; r < L.nyg; ++r)
            for (int x = 0; x < L.nxg; ++x) {
                size_t o = (size_t)L.map_off + (size_t)r * L.nxg + x;
                if (o < cap) map[o] = padded[(size_t)L.pos_off + (size_t)r * L.nxg_pad + x];
            }
    return UGD_OK;
}

int ugd_get_preprocessed(ugd_detector* d, uint8_t* gray_eq, uint8_t* mask) {
    if (!d || !d->last_plan || !d->last_was_bgr) return UGD_ERR_ARG;
    CK(cudaSetDevice(d->device));
    size_t n = (size_t)d->last.w * d->last.h;
    if (gray_eq) CK(cudaMemcpy(gray_eq, d->last.gray_eq, n, cudaMemcpyDeviceToHost));
    if (mask) CK(cudaMemcpy(mask, d->last.mask, n, cudaMemcpyDeviceToHost));
    return UGD_OK;
}

int ugd_get_level(ugd_detector* d, int level, uint8_t* img, int32_t* S, int32_t* SQ) {
    if (!d || !d->last_plan) return UGD_ERR_ARG;
    const Plan& p = d->last_plan->host;
    if (level < 0 || level >= (int)p.levels.size()) return UGD_ERR_ARG;
    CK(cudaSetDevice(d->device));
    const LevelInfo& L = p.levels[level];
    if (img) CK(cudaMemcpy2D(img, L.sw, d->last.pyr + L.img_off, L.img_pitch, L.sw, L.sh, cudaMemcpyDeviceToHost));
    size_t wbytes = (size_t)(L.sw + 1) * 4;
    if (d->last.deint2 && L.step == 2) {               // rows stored [even columns][odd columns]: put them back in order
        std::vector<int32_t> row((size_t)L.P * (L.sh + 1));
        for (int pl = 0; pl < 2; ++pl) {
            int32_t* dst = pl ? SQ : S;
            if (!dst) continue;
            CK(cudaMemcpy(row.data(), (pl ? d->last.SQ : d->last.S) + L.int_off, row.size() * 4, cudaMemcpyDeviceToHost));
            for (int y = 0; y <= L.sh; ++y)
                for (int x = 0; x <= L.sw; ++x) dst[(size_t)y * (L.sw + 1) + x] = row[(size_t)y * L.P + (x & 1) * (L.P / 2) + (x >> 1)];
        }
        return UGD_OK;
    }
    if (S) CK(cudaMemcpy2D(S, wbytes, d->last.S + L.int_off, (size_t)L.P * 4, wbytes, L.sh + 1, cudaMemcpyDeviceToHost));
    if (SQ) CK(cudaMemcpy2D(SQ, wbytes, d->last.SQ + L.int_off, (size_t)L.P * 4, wbytes, L.sh + 1, cudaMemcpyDeviceToHost));
    return UGD_OK;
}

int ugd_get_level_tilted(ugd_detector* d, int level, int32_t* T) {
    if (!d || !d->last_plan || !T) return UGD_ERR_ARG;
    const Plan& p = d->last_plan->host;
    if (level < 0 || level >= (int)p.levels.size()) return UGD_ERR_ARG;
    if (!d->last.T) return UGD_ERR_UNSUPPORTED;          // the cascade has no tilted features: the plane is never built
    CK(cudaSetDevice(d->device));
    const LevelInfo& L = p.levels[level];
    const size_t wbytes = (size_t)(L.sw + 1) * 4;
    CK(cudaMemcpy2D(T, wbytes, d->last.T + L.int_off, (size_t)L.P * 4, wbytes, L.sh + 1, cudaMemcpyDeviceToHost));
    return UGD_OK;
}

// ---- multi-GPU: frames sharded over ranks, lists gathered to the root over NCCL ------------------------------
int ugd_comm_unique_id(uint8_t* id) {
    if (!id) return UGD_ERR_ARG;
    NcclApi* n = nccl_api();
    if (!n) return UGD_ERR_COMM;
    ncclUniqueId u;
    static_assert(sizeof u == UGD_COMM_ID_BYTES, "ncclUniqueId size");
    if (n->GetUniqueId(&u) != ncclSuccess) return UGD_ERR_COMM;
    memcpy(id, &u, sizeof u);
    return UGD_OK;
}

int ugd_comm_init(ugd_detector* d, const uint8_t* id, int rank, int world) {
    if (!d || !id || world < 1 || rank < 0 || rank >= world) return UGD_ERR_ARG;
    NcclApi* n = nccl_api();
    if (!n) { d->err = "NCCL is not available: libnccl.so.2 could not be loaded"; return UGD_ERR_COMM; }
    CK(cudaSetDevice(d->device));
    comm_release(d);
    ncclUniqueId u;
    memcpy(&u, id, sizeof u);
    NK(n->CommInitRank(&d->shard.comm, world, u, rank));
    d->shard.own = true; d->shard.rank = rank; d->shard.world = world;
    return UGD_OK;
}

int ugd_comm_attach(ugd_detector* d, void* nccl_comm, int rank, int world) {
    if (!d || world < 1 || rank < 0 || rank >= world || (!nccl_comm && world > 1)) return UGD_ERR_ARG;
    if (world > 1 && !nccl_api()) { d->err = "NCCL is not available: libnccl.so.2 could not be loaded"; return UGD_ERR_COMM; }
    comm_release(d);
    d->shard.comm = static_cast<ncclComm_t>(nccl_comm); d->shard.own = false; d->shard.rank = rank; d->shard.world = world;
    return UGD_OK;
}

int ugd_comm_info(const ugd_detector* d, int* rank, int* world, int* nccl_version) {
    if (!d) return UGD_ERR_ARG;
    if (rank) *rank = d->shard.rank;
    if (world) *world = d->shard.world;
    if (nccl_version) { *nccl_version = 0; if (NcclApi* n = nccl_api()) n->GetVersion(nccl_version); }
    return UGD_OK;
}

void ugd_shard_range(int nframes, int rank, int world, int* lo, int* hi) {
    if (lo) *lo = (int)((int64_t)rank * nframes / world);
    if (hi) *hi = (int)((int64_t)(rank + 1) * nframes / world);
}

int ugd_detect_batch_sharded(ugd_detector* d, const void* frames, int is_device, int is_bgr, int nframes, int w, int h,
                             int slot_frames, int root, ugd_rect* out, int cap, int* counts) {
    if (!d || !out || !counts || nframes < 0 || slot_frames < nframes || cap <= 0 || slot_frames <= 0) return UGD_ERR_ARG;
    const int world = d->shard.world, rank = d->shard.rank;
    if (root < 0 || root >= world) return UGD_ERR_ARG;
    if (world > 1 && !d->shard.comm) { d->err = "no communicator: call ugd_comm_init or ugd_comm_attach first"; return UGD_ERR_ARG; }
    if (d->pending.active) { d->err = "a submitted batch is still pending: call ugd_wait_batch first"; return UGD_ERR_ARG; }
    CK(cudaSetDevice(d->device));
    NcclApi* nc = world > 1 ? nccl_api() : nullptr;
    const bool is_root = rank == root;
    const size_t rect_bytes = (size_t)slot_frames * cap * sizeof(ugd_rect);
    const size_t msg = rect_bytes + (size_t)slot_frames * 4;                     // one rank's message
    const size_t nslots = (size_t)world * slot_frames;
    const size_t hdr = round_up(4 + nslots * 4, 16);                             // packed: total, counts, then the rects
    ugd_rect* lout = is_root ? out + (size_t)rank * slot_frames * cap : out;     // where this rank's own lists go
    int* lcounts = is_root ? counts + (size_t)rank * slot_frames : counts;
    int rc = UGD_OK;
    if (world > 1) {
        CK(d->shard.send.reserve(msg + 256));
        CK(cudaMemsetAsync(d->shard.send.as<uint8_t>() + rect_bytes, 0xff, (size_t)slot_frames * 4, d->stream));   // absent frames: count -1
        if (is_root) {
            CK(d->shard.recv.reserve((size_t)world * msg + 256));
            CK(d->shard.packed.reserve(hdr + nslots * cap * sizeof(ugd_rect) + 256));
            const size_t need = hdr + nslots * cap * sizeof(ugd_rect);           // pinned: header, then every rect that can arrive
            if (need > d->shard.h_packed_cap || !d->shard.h_packed) {
                if (d->shard.h_packed) cudaFreeHost(d->shard.h_packed);
                d->shard.h_packed = nullptr;
                CK(cudaMallocHost((void**)&d->shard.h_packed, need));
                d->shard.h_packed_cap = need;
            }
        }
        d->shard.call = true; d->shard.slot_frames = slot_frames;
    }
    // everything of this rank's frames is enqueued (copies, kernels, result copies, the message for the gather) ...
    if (nframes > 0)
        rc = detect_body(d, frames, is_device != 0, is_bgr != 0, nframes, w, h, (size_t)w * (is_bgr ? 3 : 1), nullptr, cap, nullptr, true);
    d->shard.call = false;
    if (rc != UGD_OK) { drain(d); return rc; }
    // ... then the exchange step on the same stream: every peer sends its slots to the root, which packs what
    // arrived into one dense list before it crosses PCIe
    if (world > 1) {
        NK(nc->GroupStart());
        if (is_root) {
            for (int r = 0; r < world; ++r)
                if (r != root) NK(nc->Recv(d->shard.recv.as<uint8_t>() + (size_t)r * msg, msg, ncclUint8, r, d->shard.comm, d->stream));
        } else {
            NK(nc->Send(d->shard.send.as<uint8_t>(), msg, ncclUint8, root, d->shard.comm, d->stream));
        }
        NK(nc->GroupEnd());
        if (is_root) {
            launch_pack_lists(d->shard.recv.as<uint8_t>(), msg, world, slot_frames, cap, root, d->shard.packed.as<int>(), hdr, d->stream);
            CK(cudaGetLastError());
            CK(cudaMemcpyAsync(d->shard.h_packed, d->shard.packed.p, hdr, cudaMemcpyDeviceToHost, d->stream));
        }
    }
    if (nframes > 0) rc = wait_pending(d, lout, lcounts);                        // synchronises the stream
    else CK(cudaStreamSynchronize(d->stream));
    if (rc != UGD_OK && rc != UGD_ERR_TRUNCATED) return rc;
    if (is_root) {
        for (int f = nframes; f < slot_frames; ++f) lcounts[f] = -1;
        if (world > 1) {
            const int total = d->shard.h_packed[0];
            const int* pc = d->shard.h_packed + 1;
            ugd_rect* rects = reinterpret_cast<ugd_rect*>(reinterpret_cast<uint8_t*>(d->shard.h_packed) + hdr);
            if (total > 0) {                                                     // the rects that exist, not world x frames x cap slots
                CK(cudaMemcpyAsync(rects, d->shard.packed.as<uint8_t>() + hdr, (size_t)total * sizeof(ugd_rect), cudaMemcpyDeviceToHost, d->stream));
                CK(cudaStreamSynchronize(d->stream));
            }
            size_t at = 0;
            for (int r = 0; r < world; ++r) {
                if (r == root) continue;
                for (int f = 0; f < slot_frames; ++f) {
                    const size_t e = (size_t)r * slot_frames + f;
                    const int n = pc[e];
                    counts[e] = n;
                    const int take = std::max(std::min(n, cap), 0);
                    if (take) memcpy(out + e * cap, rects + at, (size_t)take * sizeof(ugd_rect));
                    at += (size_t)take;
                    if (n > cap) rc = UGD_ERR_TRUNCATED;
                }
            }
            if (rc == UGD_ERR_TRUNCATED) d->err = "a frame has more rectangles than cap: counts hold the true numbers, call again with a larger cap";
        }
    }
    return rc;
}

const char* ugd_strerror(int code) {
    switch (code) {
        case UGD_OK: return "ok";
        case UGD_ERR_ARG: return "bad argument";
        case UGD_ERR_IO: return "cannot read file";
        case UGD_ERR_PARSE: return "malformed cascade or config file";
        case UGD_ERR_UNSUPPORTED: return "unsupported cascade feature";
        case UGD_ERR_CUDA: return "CUDA error";
        case UGD_ERR_CAPACITY: return "capacity exceeded";
        case UGD_ERR_NOMEM: return "out of memory";
        case UGD_ERR_TRUNCATED: return "more rectangles than cap (true counts returned)";
        case UGD_ERR_COMM: return "NCCL error";
        default: return "unknown error";
    }
}

const char* ugd_last_error(const ugd_detector* d) { return d ? d->err.c_str() : ""; }
const char* ugd_version(void) { return "ugdet-b200 0.1 (sm_100a)"; }

}  // extern "C"
