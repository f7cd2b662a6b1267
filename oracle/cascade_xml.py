"""TEST INFRASTRUCTURE (oracle side) -- not part of the product path.

Independent Python (ElementTree) reader for OpenCV Haar cascade XML files, used to
(a) feed the C restatement in ``haar_oracle.c`` and (b) cross-check the product's own
C++ XML loader (``ugproj-detector_b200/csrc/cascade_xml.cc``) value by value.

Handles both formats the reference can meet:
  * old format ``type_id="opencv-haar-classifier"`` -- what the reference ships in
    ``data/haarcascade_frontalface_*.xml`` (e.g. ``data/haarcascade_frontalface_alt.xml:46``)
    and loads at ``src/file_io.cc:11``;
  * new format ``type_id="opencv-cascade-classifier"`` -- what cv2 4.13 converts the old
    files to internally and ships under ``cv2/data/``.

The flat layout returned here is the one documented in ``oracle/haar_oracle.h``.
"""
from __future__ import annotations

import xml.etree.ElementTree as ET
from dataclasses import dataclass

import numpy as np

THRESHOLD_EPS = np.float32(1e-5)   # stage_thr = f32(xml) - 1e-5f


@dataclass
class FlatCascade:
    win_w: int
    win_h: int
    stage_first_tree: np.ndarray   # int32 [nstages]
    stage_ntrees: np.ndarray       # int32 [nstages]
    stage_thr: np.ndarray          # float32 [nstages]  (already f32(xml) - 1e-5f)
    tree_first_node: np.ndarray    # int32 [ntrees]
    tree_nnodes: np.ndarray        # int32 [ntrees]
    node_rects: np.ndarray         # int32 [nnodes, 3, 4]   x,y,w,h  (unused rect = zeros)
    node_weights: np.ndarray       # float32 [nnodes, 3]   (0 = rect absent)
    node_thr: np.ndarray           # float32 [nnodes]
    node_left: np.ndarray          # int32 [nnodes]  >0: child node index inside the tree, 0: leaf
    node_right: np.ndarray         # int32 [nnodes]
    node_left_val: np.ndarray      # float32 [nnodes]  leaf value when node_left == 0
    node_right_val: np.ndarray     # float32 [nnodes]
    node_tilted: np.ndarray = None  # int32 [nnodes]  1: the node's rects are 45-degree tilted (read the tilted integral)
    tilted: bool = False
    # the stage tree of the old format (<parent>/<next>, -1 = none; a linear chain elsewhere) and the stage thresholds as
    # written in the file -- only the OpenCV-2.4 legacy restatement (oracle/haar_legacy.c) reads them
    stage_thr_raw: np.ndarray = None   # float32 [nstages]  f32(xml)
    stage_parent: np.ndarray = None    # int32 [nstages]
    stage_next: np.ndarray = None      # int32 [nstages]

    @property
    def nstages(self):
        return len(self.stage_thr)

    @property
    def ntrees(self):
        return len(self.tree_first_node)

    @property
    def nnodes(self):
        return len(self.node_thr)


def _f32(txt: str) -> np.float32:
    return np.float32(float(txt))


def _finish(win, stages, trees, nodes, tilted, tree_info=None):
    n = len(nodes)
    rects = np.zeros((n, 3, 4), np.int32)
    weights = np.zeros((n, 3), np.float32)
    thr = np.zeros(n, np.float32)
    left = np.zeros(n, np.int32)
    right = np.zeros(n, np.int32)
    lval = np.zeros(n, np.float32)
    rval = np.zeros(n, np.float32)
    tl = np.zeros(n, np.int32)
    for i, nd in enumerate(nodes):
        tl[i] = int(bool(nd.get("tilted", False)))
        for k, (x, y, w, h, wt) in enumerate(nd["rects"]):
            rects[i, k] = (x, y, w, h)
            weights[i, k] = wt
        thr[i] = nd["thr"]
        left[i], right[i] = nd["left"], nd["right"]
        lval[i], rval[i] = nd["lval"], nd["rval"]
    return FlatCascade(
        win_w=win[0], win_h=win[1],
        stage_first_tree=np.array([s[0] for s in stages], np.int32),
        stage_ntrees=np.array([s[1] for s in stages], np.int32),
        stage_thr=np.array([s[2] for s in stages], np.float32),
        tree_first_node=np.array([t[0] for t in trees], np.int32),
        tree_nnodes=np.array([t[1] for t in trees], np.int32),
        node_rects=rects, node_weights=weights, node_thr=thr,
        node_left=left, node_right=right, node_left_val=lval, node_right_val=rval,
        node_tilted=tl, tilted=bool(tl.any()),
        stage_thr_raw=np.array([s[3] for s in stages], np.float32),
        stage_parent=np.array([s[4] for s in stages], np.int32),
        stage_next=np.array([s[5] for s in stages], np.int32))


def _parse_rect(txt):
    p = txt.split()
    return int(p[0]), int(p[1]), int(p[2]), int(p[3]), np.float32(float(p[4]))


def _load_old(root_el):
    win = tuple(int(v) for v in root_el.find("size").text.split())
    stages, trees, nodes = [], [], []
    tilted = False
    for st in root_el.find("stages"):
        first_tree = len(trees)
        for tr in st.find("trees"):
            first_node = len(nodes)
            tnodes = list(tr)
            for nd in tnodes:
                feat = nd.find("feature")
                rects = [_parse_rect(r.text) for r in feat.find("rects")]
                t = feat.find("tilted")
                d = {"rects": rects, "thr": _f32(nd.find("threshold").text),
                     "tilted": t is not None and int(t.text) != 0,
                     "left": 0, "right": 0, "lval": np.float32(0), "rval": np.float32(0)}
                if nd.find("left_val") is not None:
                    d["lval"] = _f32(nd.find("left_val").text)
                else:
                    d["left"] = int(nd.find("left_node").text)
                if nd.find("right_val") is not None:
                    d["rval"] = _f32(nd.find("right_val").text)
                else:
                    d["right"] = int(nd.find("right_node").text)
                nodes.append(d)
            trees.append((first_node, len(tnodes)))
        raw = _f32(st.find("stage_threshold").text)
        thr = np.float32(raw - THRESHOLD_EPS)
        par = int(st.find("parent").text) if st.find("parent") is not None else len(stages) - 1
        nxt = int(st.find("next").text) if st.find("next") is not None else -1
        stages.append((first_tree, len(trees) - first_tree, thr, raw, par, nxt))
    return _finish(win, stages, trees, nodes, tilted)


def _load_new(root_el):
    if root_el.find("featureType").text.strip() != "HAAR":
        raise ValueError("only HAAR cascades are supported")
    win = (int(root_el.find("width").text), int(root_el.find("height").text))
    feats = []
    tilted = False
    for f in root_el.find("features"):
        rects = [_parse_rect(r.text) for r in f.find("rects")]
        t = f.find("tilted")
        feats.append((rects, t is not None and int(t.text) != 0))
    stages, trees, nodes = [], [], []
    for st in root_el.find("stages"):
        first_tree = len(trees)
        for wc in st.find("weakClassifiers"):
            inodes = wc.find("internalNodes").text.split()
            leaves = [_f32(v) for v in wc.find("leafValues").text.split()]
            nn = len(inodes) // 4
            first_node = len(nodes)
            for k in range(nn):
                l, r, fi = int(inodes[4 * k]), int(inodes[4 * k + 1]), int(inodes[4 * k + 2])
                d = {"rects": feats[fi][0], "tilted": feats[fi][1], "thr": _f32(inodes[4 * k + 3]),
                     "left": max(l, 0), "right": max(r, 0),
                     "lval": leaves[-l] if l <= 0 else np.float32(0),
                     "rval": leaves[-r] if r <= 0 else np.float32(0)}
                nodes.append(d)
            trees.append((first_node, nn))
        raw = _f32(st.find("stageThreshold").text)
        thr = np.float32(raw - THRESHOLD_EPS)
        stages.append((first_tree, len(trees) - first_tree, thr, raw, len(stages) - 1, -1))      # the new format has no stage tree
    return _finish(win, stages, trees, nodes, tilted)


def load_cascade(path: str) -> FlatCascade:
    root = ET.parse(path).getroot()
    for el in root:
        tid = el.get("type_id")
        if tid == "opencv-haar-classifier":
            return _load_old(el)
        if tid == "opencv-cascade-classifier" or (tid is None and el.find("stageType") is not None):
            return _load_new(el)          # some shipped files (frontalcatface*) carry no type_id
    raise ValueError(f"{path}: no Haar cascade node found")
