"""Seeded synthetic frames for tests and bench (SURVEY.md A.7 generator, restated).

The reference ships no sample video (its cfg points at a URL), so every measurement in this repo
runs on these frames: band-limited noise background + procedural face-like patches + pixel noise.
cv2 is used only to synthesise inputs (blur / bicubic up-sampling), never on the detection path.
"""
from __future__ import annotations

import numpy as np


def _face(sz, r):
    import cv2
    y, x = np.mgrid[0:sz, 0:sz].astype(np.float32) / sz
    img = np.full((sz, sz), float(r.integers(110, 170)), np.float32)
    img[((x - .5) / .42) ** 2 + ((y - .5) / .52) ** 2 < 1] = float(r.integers(170, 210))

    def blob(cx, cy, rx, ry, v):
        img[((x - cx) / rx) ** 2 + ((y - cy) / ry) ** 2 < 1] = v

    def j():
        return float(r.normal(0, .01))

    blob(.32 + j(), .40 + j(), .11, .055, 60); blob(.68 + j(), .40 + j(), .11, .055, 60)
    blob(.32, .30, .13, .03, 90); blob(.68, .30, .13, .03, 90)
    blob(.5, .55, .05, .12, 215)
    blob(.5, .78 + j(), .17, .045, 90)
    blob(.25, .6, .1, .1, 210); blob(.75, .6, .1, .1, 210)
    return cv2.GaussianBlur(img, (0, 0), sz / 40)


def gray_frame(w, h, nfaces, seed, noise=3):
    """uint8 [h,w] gray frame with `nfaces` procedural faces."""
    import cv2
    r = np.random.default_rng(seed)
    img = cv2.resize(r.normal(110, 25, (max(h // 8, 1), max(w // 8, 1))).astype(np.float32), (w, h),
                     interpolation=cv2.INTER_CUBIC)
    for _ in range(nfaces):
        hi = max(min(h, w) // 3, 31)
        s = int(r.integers(30, hi))
        px = int(r.integers(0, w - s)); py = int(r.integers(0, h - s))
        img[py:py + s, px:px + s] = _face(s, r)
    img += r.normal(0, noise, img.shape)
    return np.clip(img, 0, 255).astype(np.uint8)


def bgr_frame(w, h, nfaces, seed, noise=3):
    """uint8 [h,w,3] BGR frame whose face pixels mostly fall inside the default skin box."""
    g = gray_frame(w, h, nfaces, seed, noise).astype(np.float32)
    r = np.random.default_rng(seed + 100003)
    out = np.stack([g * 0.75, g * 0.85, g], axis=-1) + r.normal(0, 1.5, (h, w, 3))
    return np.clip(out, 0, 255).astype(np.uint8)


def default_nfaces(w, h):
    return 8 if h <= 720 else (12 if h <= 1080 else 24)


def gray_batch(w, h, n, seed0=0, nfaces=None, noise=3):
    nf = default_nfaces(w, h) if nfaces is None else nfaces
    return np.stack([gray_frame(w, h, nf, seed0 + i, noise) for i in range(n)])
