"""TEST INFRASTRUCTURE -- the reference's own CPU path, call for call, through cv2 4.13.

``detect_faces_cv2`` is a 1:1 transliteration of ``FaceDetector::detectFaces``
(/root/reference/src/tracker/detector.cc:10-43): the same ten OpenCV calls in the same order.
OpenCV is the reference's un-vendored dependency (CMakeLists.txt:13); the only OpenCV in this
image is opencv-python-headless 4.13.0, so this module is both what pins the C oracle and the
"reference" arm of bench.py (`--impl reference`).  It never touches the GPU engine.
"""
from __future__ import annotations

import numpy as np

import cv2


def cascade_dir() -> str:
    """Where Haar cascade XMLs can be found on this machine: the reference's data/ when the
    reference tree is mounted, else the copies OpenCV ships (same numbers, new XML format)."""
    import os
    ref = "/root/reference/data"
    return ref if os.path.isdir(ref) else cv2.data.haarcascades


def cascade_path(name: str) -> str:
    """"alt" -> haarcascade_frontalface_alt.xml (the reference's four); any other cv2 Haar cascade by its
    short name ("smile", "upperbody", ...) or an explicit file path."""
    import os
    if os.path.isfile(name):
        return name
    p = os.path.join(cascade_dir(), f"haarcascade_frontalface_{name}.xml")
    if os.path.isfile(p):
        return p
    return os.path.join(cv2.data.haarcascades, f"haarcascade_{name}.xml")


def load(path: str) -> "cv2.CascadeClassifier":
    c = cv2.CascadeClassifier()
    if not c.load(path):                              # file_io.cc:11
        raise IOError(f"cannot load cascade {path}")
    return c


def detect_multiscale_cv2(cascade, gray, scale=1.1, min_neighbors=2, min_size=(30, 30), max_size=None):
    kw = {} if not max_size or max_size == (0, 0) else {"maxSize": tuple(max_size)}
    r = cascade.detectMultiScale(gray, scaleFactor=scale, minNeighbors=min_neighbors,
                                 flags=cv2.CASCADE_SCALE_IMAGE, minSize=tuple(min_size), **kw)
    return np.asarray(r, np.int32).reshape(-1, 4)


def detect_faces_cv2(cascade, frame_bgr, scale=1.1, skin_lower=(0, 133, 77), skin_upper=(255, 173, 127),
                     min_neighbors=2, min_size=(30, 30)):
    ycrcb = cv2.cvtColor(frame_bgr, cv2.COLOR_BGR2YCrCb)                       # detector.cc:17
    mask = cv2.inRange(ycrcb, np.array(skin_lower, np.float64), np.array(skin_upper, np.float64))  # :18
    gray = cv2.cvtColor(frame_bgr, cv2.COLOR_BGR2GRAY)                         # :20
    gray = cv2.equalizeHist(gray)                                              # :21
    gray = cv2.bitwise_and(gray, mask)                                         # :22
    faces = detect_multiscale_cv2(cascade, gray, scale, min_neighbors, min_size)  # :25-31
    out = []
    for (x, y, w, h) in faces:                                                 # :33-42
        m = np.linalg.norm(cv2.mean(mask[y:y + h, x:x + w]))
        if m / 256 < 0.8:
            continue
        out.append((x, y, w, h))
    return np.asarray(out, np.int32).reshape(-1, 4)
