"""TEST INFRASTRUCTURE -- ctypes front end of the OpenCV-2.4 legacy restatement (``haar_legacy.c``).

PARITY UNPINNED: there is no OpenCV 2.4 in this image; the GPU engine's legacy mode is checked against this
restatement only.  Only tests/ may import this module."""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import oracle as orc


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def _extras(casc: orc.OracleCascade):
    f = casc.flat
    keep = [np.ascontiguousarray(f.stage_thr_raw, np.float32), np.ascontiguousarray(f.stage_parent, np.int32),
            np.ascontiguousarray(f.stage_next, np.int32)]
    return keep


def resize_linear(gray, dw, dh):
    """cvResize(..., CV_INTER_LINEAR) of 2.4 for 8UC1 (11-bit coefficients)."""
    gray = np.ascontiguousarray(gray, np.uint8)
    h, w = gray.shape
    out = np.empty((dh, dw), np.uint8)
    orc.lib().orc24_resize_linear(_p(gray), w, h, C.c_size_t(w), _p(out), dw, dh)
    return out


def integral(gray):
    gray = np.ascontiguousarray(gray, np.uint8)
    h, w = gray.shape
    S = np.empty((h + 1, w + 1), np.int32)
    SQ = np.empty((h + 1, w + 1), np.float64)
    orc.lib().orc24_integral(_p(gray), w, h, _p(S), _p(SQ))
    return S, SQ


def bgr2gray(bgr):
    bgr = np.ascontiguousarray(bgr, np.uint8)
    out = np.empty(bgr.shape[:2], np.uint8)
    orc.lib().orc24_bgr2gray(_p(bgr), C.c_int(out.size), _p(out))
    return out


def scale_list(W, H, ow, oh, scale_factor=1.1, min_size=(30, 30), max_size=(0, 0)):
    buf = np.zeros(256, np.float64)
    n = orc.lib().orc24_scale_list(W, H, ow, oh, C.c_double(scale_factor), min_size[0], min_size[1], max_size[0], max_size[1],
                                   _p(buf), 256)
    return buf[:n].copy()


def detect_raw(casc: orc.OracleCascade, gray, scale_factor=1.1, min_size=(30, 30), max_size=(0, 0), want_map=False, cap=1 << 17):
    gray = np.ascontiguousarray(gray, np.uint8)
    H, W = gray.shape
    thr, par, nxt = _extras(casc)
    rects = (orc._Rect * cap)()
    ctr = orc.Counters()
    map_cap = int(W) * int(H) * 8 if want_map else 0
    smap = np.full(max(map_cap, 1), -32768, np.int16)
    n = orc.lib().orc24_detect_raw(C.byref(casc.c), _p(thr), _p(par), _p(nxt), _p(gray), W, H, C.c_size_t(W), C.c_double(scale_factor),
                                   min_size[0], min_size[1], max_size[0], max_size[1], rects, cap, C.byref(ctr),
                                   _p(smap) if want_map else None, C.c_size_t(map_cap))
    assert n <= cap
    out = orc._rects_to_np(rects, n)
    if want_map:
        return out, ctr.as_dict(), smap[:ctr.grid_windows].copy()
    return out, ctr.as_dict()


def detect_multiscale(casc: orc.OracleCascade, gray, scale_factor=1.1, min_neighbors=2, min_size=(30, 30), max_size=(0, 0), cap=1 << 16):
    gray = np.ascontiguousarray(gray, np.uint8)
    H, W = gray.shape
    thr, par, nxt = _extras(casc)
    rects = (orc._Rect * cap)()
    ctr = orc.Counters()
    n = orc.lib().orc24_detect_multiscale(C.byref(casc.c), _p(thr), _p(par), _p(nxt), _p(gray), W, H, C.c_size_t(W),
                                          C.c_double(scale_factor), min_neighbors, min_size[0], min_size[1], max_size[0],
                                          max_size[1], rects, cap, C.byref(ctr))
    return orc._rects_to_np(rects, min(n, cap)), ctr.as_dict()


def detect_faces(casc: orc.OracleCascade, bgr, scale_factor=1.1, min_neighbors=2, min_size=(30, 30), skin_lower=(0, 133, 77),
                 skin_upper=(255, 173, 127)):
    """detector.cc:10-43 with the 2.4 pieces: 14-bit BGR2GRAY, equalizeHist, AND mask, legacy detectMultiScale, skin filter
    (rects are not clipped in 2.4, so the mask mean is taken over the part of the rect inside the frame)."""
    bgr = np.ascontiguousarray(bgr, np.uint8)
    H, W = bgr.shape[:2]
    mask = orc.inrange3(orc.bgr2ycrcb(bgr), skin_lower, skin_upper)
    gray = orc.equalize_hist(bgr2gray(bgr)) & mask
    rects, ctr = detect_multiscale(casc, gray, scale_factor, min_neighbors, min_size)
    keep = []
    for x, y, w, h in rects.tolist():
        x0, y0, x1, y1 = max(x, 0), max(y, 0), min(x + w, W), min(y + h, H)
        if x1 <= x0 or y1 <= y0:
            continue
        m = float(mask[y0:y1, x0:x1].astype(np.float64).mean())
        if m / 256 < 0.8:
            continue
        keep.append((x, y, w, h))
    return np.asarray(keep, np.int32).reshape(-1, 4), ctr
