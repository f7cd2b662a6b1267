"""Generates the golden fixtures in this directory from the REFERENCE path itself: cv2 4.13
(opencv-python-headless 4.13.0.92) driven exactly as /root/reference/src/tracker/detector.cc:10-43
drives OpenCV, on seeded synthetic frames.  Run from the repo root:

    python tests/golden/make_golden.py

Output: golden_detections.json (small, committed).  Frames are NOT stored -- they are regenerated
from their seeds by ugproj-detector_b200/synth.py, and each entry carries a checksum of the frame so
a drift of the generator is caught before any comparison is made.
"""
import hashlib
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402

ge.load_package()
import importlib  # noqa: E402

synth = importlib.import_module(ge.PKG_NAME + ".synth")
from oracle import cv2_ref  # noqa: E402
import cv2  # noqa: E402

CASES = [  # (cascade, w, h, nfaces, seed, scale, min_size, kind)
    ("alt", 640, 360, 6, 0, 1.1, (30, 30), "gray"),
    ("alt", 640, 360, 6, 1, 1.1, (30, 30), "gray"),
    ("alt2", 640, 360, 6, 0, 1.1, (30, 30), "gray"),
    ("default", 640, 360, 6, 1, 1.1, (30, 30), "gray"),
    ("alt_tree", 640, 360, 6, 0, 1.1, (30, 30), "gray"),
    ("alt", 1280, 720, 8, 1, 1.1, (30, 30), "gray"),          # BASELINE configs[0]
    ("alt2", 1280, 720, 8, 1, 1.1, (30, 30), "gray"),         # configs[1]
    ("alt_tree", 1280, 720, 8, 1, 1.1, (30, 30), "gray"),     # configs[3]
    ("alt", 480, 270, 5, 2, 1.05, (24, 24), "gray"),          # configs[4] parameters on a small frame
    ("alt2", 640, 360, 6, 0, 1.1, (30, 30), "bgr"),           # the cfg as shipped: full detectFaces
    ("alt2", 640, 360, 6, 3, 1.1, (30, 30), "bgr"),
]


def main():
    out = {"generator": "tests/golden/make_golden.py", "cv2": cv2.__version__, "cases": []}
    for name, w, h, nf, seed, sf, ms, kind in CASES:
        cc = cv2_ref.load(cv2_ref.cascade_path(name))
        if kind == "gray":
            f = synth.gray_frame(w, h, nf, seed)
            raw = cc.detectMultiScale(f, scaleFactor=sf, minNeighbors=0, flags=cv2.CASCADE_SCALE_IMAGE, minSize=ms)
            rects = cv2_ref.detect_multiscale_cv2(cc, f, sf, 2, ms)
        else:
            f = synth.bgr_frame(w, h, nf, seed)
            raw = np.zeros((0, 4), np.int32)
            rects = cv2_ref.detect_faces_cv2(cc, f, sf, min_size=ms)
        raw = np.asarray(raw, np.int32).reshape(-1, 4)
        out["cases"].append({
            "cascade": name, "w": w, "h": h, "nfaces": nf, "seed": seed, "scale": sf, "min_size": list(ms),
            "kind": kind, "frame_sha1": hashlib.sha1(f.tobytes()).hexdigest(),
            "raw_count": int(len(raw)),
            "raw_sha1": hashlib.sha1(np.array(sorted(map(tuple, raw.tolist())), np.int32).tobytes()).hexdigest(),
            "rects": sorted(map(list, rects.tolist())),
        })
        print(name, w, h, kind, "raw", len(raw), "rects", len(rects))
    with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden_detections.json"), "w") as fh:
        json.dump(out, fh, indent=1)


if __name__ == "__main__":
    main()
