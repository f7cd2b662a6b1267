// Host-side cascade model (immutable after load) and the corner-list form the device layouts are built from.
// Replaces the cv::CascadeClassifier object the reference loads at src/file_io.cc:11.
#pragma once
#include <cstdint>
#include <string>
#include <vector>

namespace ugd {

struct HostNode {
    int32_t rect[3][4];     // x, y, w, h ; unused rect = zeros
    float   weight[3];      // 0 = rect absent
    float   thr;
    int32_t left, right;    // > 0: child node index inside the tree, 0: leaf
    float   left_val, right_val;
    int32_t tilted;         // 1: the rects are 45-degree tilted (read the tilted integral)
};

// Corner-list form of a node: feature = sum_j coef[j] * S[corner j], duplicates merged
// (6, 8 or 9 distinct corners for the shipped cascades instead of 8 / 12).  Exact in int32.
struct CornerNode {
    int nc;                 // corners used, padded with coef 0 up to 6, 8, 9 or 12
    int8_t cx[12], cy[12];  // corner position inside the window
    int8_t coef[12];
    int8_t plane;           // 0: upright integral S, 1: tilted integral T
};

struct Cascade {
    int win_w = 0, win_h = 0;
    bool old_format = false;
    bool sum_exact = false;
    bool has_tilted = false;                // some node reads the tilted integral
    bool delta_exact = false;               // base + sum of (left - right) is exact too (see cascade_xml.cc)
    int max_tree_nodes = 0;
    std::vector<int32_t> stage_first_tree, stage_ntrees;
    std::vector<float>   stage_thr;              // f32(xml) - 1e-5f
    std::vector<float>   stage_thr_raw;          // f32(xml): the OpenCV-2.4 legacy mode applies its own bias (1e-4)
    std::vector<int32_t> stage_parent, stage_next;   // the stage tree of the old format (-1 = none); a linear chain otherwise
    std::vector<int32_t> tree_first_node, tree_nnodes;
    std::vector<HostNode> nodes;
    // derived
    std::vector<CornerNode> corners;             // one per node, same order
    std::vector<int32_t> stage_first_node;       // [nstages+1]
    int nstages() const { return (int)stage_thr.size(); }
    int ntrees() const { return (int)tree_first_node.size(); }
    int nnodes() const { return (int)nodes.size(); }
};

// Returns UGD_OK or an error code; err receives a one-line reason.
int load_cascade_xml(const std::string& path, Cascade& out, std::string& err);

}  // namespace ugd

struct ugd_cascade { ugd::Cascade c; uint64_t uid = 0; };   // uid: process-unique, never reused (keys the constant bank)
