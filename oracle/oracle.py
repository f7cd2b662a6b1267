"""TEST INFRASTRUCTURE -- ctypes front end of the C oracle (``haar_oracle.c``).

Only tests/, ``__graft_entry__.smoke()`` and bench.py's cpu_baseline leg may import this.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

from .cascade_xml import FlatCascade, load_cascade  # noqa: F401

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

LOWVAR, SKIP = -1, -2


class _Rect(C.Structure):
    _fields_ = [("x", C.c_int32), ("y", C.c_int32), ("w", C.c_int32), ("h", C.c_int32)]


class _Cascade(C.Structure):
    _fields_ = [("win_w", C.c_int32), ("win_h", C.c_int32), ("nstages", C.c_int32), ("ntrees", C.c_int32),
                ("nnodes", C.c_int32)] + [(n, C.c_void_p) for n in (
                    "stage_first_tree", "stage_ntrees", "stage_thr", "tree_first_node", "tree_nnodes",
                    "node_rects", "node_weights", "node_thr", "node_left", "node_right", "node_left_val",
                    "node_right_val", "node_tilted")] + [("has_tilted", C.c_int32)]


class Counters(C.Structure):
    _fields_ = [(n, C.c_int64) for n in ("levels", "level_pixels", "grid_windows", "windows_evaluated",
                                         "lowvar", "weak_evals", "raw", "grouped")]

    def as_dict(self):
        return {n: int(getattr(self, n)) for n, _ in self._fields_}


def build(force: bool = False) -> str:
    """Compile ``libhaar_oracle.so`` with gcc (recipe: oracle/Makefile)."""
    so = os.path.join(_HERE, "libhaar_oracle.so")
    srcs = [os.path.join(_HERE, f) for f in ("haar_oracle.c", "haar_legacy.c", "haar_oracle.h")]
    if force or not os.path.exists(so) or os.path.getmtime(so) < max(os.path.getmtime(f) for f in srcs):
        subprocess.check_call(["make", "-C", _HERE, "-s", "-B", "CC=gcc"])
    return so


def lib():
    global _LIB
    if _LIB is None:
        _LIB = C.CDLL(build())
        _LIB.orc_run_at.restype = C.c_int
    return _LIB


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


class OracleCascade:
    """Keeps the numpy arrays alive next to the C view of them."""

    def __init__(self, flat: FlatCascade):
        self.flat = flat
        f = flat
        self._keep = [np.ascontiguousarray(a) for a in (
            f.stage_first_tree, f.stage_ntrees, f.stage_thr, f.tree_first_node, f.tree_nnodes,
            f.node_rects.reshape(-1), f.node_weights.reshape(-1), f.node_thr, f.node_left, f.node_right,
            f.node_left_val, f.node_right_val, f.node_tilted)]
        self.c = _Cascade(f.win_w, f.win_h, f.nstages, f.ntrees, f.nnodes, *[_p(a) for a in self._keep],
                          int(f.tilted))

    @classmethod
    def load(cls, path):
        return cls(load_cascade(path))


def _rects_to_np(buf, n):
    out = np.zeros((n, 4), np.int32)
    for i in range(n):
        out[i] = (buf[i].x, buf[i].y, buf[i].w, buf[i].h)
    return out


def bgr2gray(bgr):
    bgr = np.ascontiguousarray(bgr, np.uint8)
    out = np.empty(bgr.shape[:2], np.uint8)
    lib().orc_bgr2gray(_p(bgr), C.c_int(out.size), _p(out))
    return out


def bgr2ycrcb(bgr):
    bgr = np.ascontiguousarray(bgr, np.uint8)
    out = np.empty_like(bgr)
    lib().orc_bgr2ycrcb(_p(bgr), C.c_int(bgr.shape[0] * bgr.shape[1]), _p(out))
    return out


def inrange3(img, lo, hi):
    img = np.ascontiguousarray(img, np.uint8)
    out = np.empty(img.shape[:2], np.uint8)
    lo = (C.c_uint8 * 3)(*lo)
    hi = (C.c_uint8 * 3)(*hi)
    lib().orc_inrange3(_p(img), C.c_int(out.size), lo, hi, _p(out))
    return out


def equalize_hist(gray):
    gray = np.ascontiguousarray(gray, np.uint8)
    out = np.empty_like(gray)
    lib().orc_equalize_hist(_p(gray), C.c_int(gray.size), _p(out))
    return out


def resize_area(img, dw, dh):
    """cv2.resize(img, (dw, dh), interpolation=INTER_AREA) for decimation (face_tracker.cc:112-114)."""
    img = np.ascontiguousarray(img, np.uint8)
    h, w = img.shape[:2]
    cn = 1 if img.ndim == 2 else img.shape[2]
    out = np.zeros((dh, dw) if img.ndim == 2 else (dh, dw, cn), np.uint8)
    rc = lib().orc_resize_area(_p(img), w, h, cn, C.c_size_t(w * cn), _p(out), dw, dh)
    if rc:
        raise ValueError("INTER_AREA restatement covers decimation only")
    return out


def resize_exact(gray, dw, dh):
    gray = np.ascontiguousarray(gray, np.uint8)
    h, w = gray.shape
    out = np.empty((dh, dw), np.uint8)
    lib().orc_resize_exact(_p(gray), w, h, C.c_size_t(w), _p(out), dw, dh)
    return out


def integral(gray):
    gray = np.ascontiguousarray(gray, np.uint8)
    h, w = gray.shape
    S = np.empty((h + 1, w + 1), np.int32)
    SQ = np.empty((h + 1, w + 1), np.int32)
    lib().orc_integral(_p(gray), w, h, _p(S), _p(SQ))
    return S, SQ


def integral_tilted(gray):
    """Third output of cv2.integral3 (CV_32S): the 45-degree rotated integral tilted cascades read."""
    gray = np.ascontiguousarray(gray, np.uint8)
    h, w = gray.shape
    T = np.empty((h + 1, w + 1), np.int32)
    lib().orc_integral_tilted(_p(gray), w, h, _p(T))
    return T


def scale_list(W, H, ow, oh, scale_factor=1.1, min_size=(30, 30), max_size=(0, 0)):
    buf = np.zeros(256, np.float32)
    n = lib().orc_scale_list(W, H, ow, oh, C.c_double(scale_factor), min_size[0], min_size[1], max_size[0],
                             max_size[1], _p(buf), 256)
    return buf[:n].copy()


def level_size(W, H, s):
    a, b = C.c_int(), C.c_int()
    lib().orc_level_size(W, H, C.c_float(s), C.byref(a), C.byref(b))
    return a.value, b.value


def detect_raw(casc: OracleCascade, gray, scale_factor=1.1, min_size=(30, 30), max_size=(0, 0),
               want_map=False, cap=1 << 16):
    """All accepted windows (unclipped) + final-stage sums + counters (+ per-position stage map)."""
    gray = np.ascontiguousarray(gray, np.uint8)
    H, W = gray.shape
    rects = (_Rect * cap)()
    sums = np.zeros(cap, np.float64)
    ctr = Counters()
    smap = None
    map_cap = 0
    if want_map:
        map_cap = int(W) * int(H) * 8
        smap = np.full(map_cap, -32768, np.int16)
    n = lib().orc_detect_raw(C.byref(casc.c), _p(gray), W, H, C.c_size_t(W), C.c_double(scale_factor),
                             min_size[0], min_size[1], max_size[0], max_size[1], rects, _p(sums), cap,
                             C.byref(ctr), _p(smap) if want_map else None, C.c_size_t(map_cap))
    assert n <= cap
    out = _rects_to_np(rects, n)
    if want_map:
        return out, sums[:n].copy(), ctr.as_dict(), smap[:ctr.grid_windows].copy()
    return out, sums[:n].copy(), ctr.as_dict()


def group_rectangles(rects, min_neighbors=2, eps=0.2):
    rects = np.asarray(rects, np.int32).reshape(-1, 4)
    n = len(rects)
    inb = (_Rect * max(n, 1))(*[_Rect(*map(int, r)) for r in rects])
    outb = (_Rect * max(n, 1))()
    cnt = np.zeros(max(n, 1), np.int32)
    m = lib().orc_group_rectangles(inb, n, min_neighbors, C.c_double(eps), outb, _p(cnt))
    return _rects_to_np(outb, m), cnt[:m].copy()


def detect_multiscale(casc: OracleCascade, gray, scale_factor=1.1, min_neighbors=2, min_size=(30, 30),
                      max_size=(0, 0), cap=4096):
    gray = np.ascontiguousarray(gray, np.uint8)
    H, W = gray.shape
    rects = (_Rect * cap)()
    ctr = Counters()
    n = lib().orc_detect_multiscale(C.byref(casc.c), _p(gray), W, H, C.c_size_t(W), C.c_double(scale_factor),
                                    min_neighbors, min_size[0], min_size[1], max_size[0], max_size[1], rects,
                                    cap, C.byref(ctr))
    return _rects_to_np(rects, min(n, cap)), ctr.as_dict()


def detect_faces(casc: OracleCascade, bgr, scale_factor=1.1, min_neighbors=2, min_size=(30, 30),
                 max_size=(0, 0), skin_lower=(0, 133, 77), skin_upper=(255, 173, 127), cap=4096):
    """== FaceDetector::detectFaces (detector.cc:10-43) on one BGR frame."""
    bgr = np.ascontiguousarray(bgr, np.uint8)
    H, W = bgr.shape[:2]
    rects = (_Rect * cap)()
    ctr = Counters()
    lo = (C.c_uint8 * 3)(*skin_lower)
    hi = (C.c_uint8 * 3)(*skin_upper)
    n = lib().orc_detect_faces(C.byref(casc.c), _p(bgr), W, H, C.c_double(scale_factor), min_neighbors,
                               min_size[0], min_size[1], max_size[0], max_size[1], lo, hi, rects, cap,
                               C.byref(ctr))
    return _rects_to_np(rects, n), ctr.as_dict()
