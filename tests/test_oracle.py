"""CPU tests (no GPU): the C oracle is pinned against the reference's own arithmetic -- live cv2 4.13
and the committed golden fixtures (tests/golden/, generated from cv2 by make_golden.py)."""
import hashlib
import json
import os

import numpy as np
import pytest

from conftest import CASCADES, all_cv2_haar_cascades, pass_everything_cascade, rectset, truncate_cascade

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "golden_detections.json")


def _clip(raw, w, h):
    r = raw.copy()
    x1 = np.minimum(r[:, 0] + r[:, 2], w); y1 = np.minimum(r[:, 1] + r[:, 3], h)
    r[:, 2] = x1 - r[:, 0]; r[:, 3] = y1 - r[:, 1]
    return r


def test_building_blocks_match_cv2(orc, synth):
    import cv2
    rng = np.random.default_rng(0)
    bgr = rng.integers(0, 256, (211, 317, 3), dtype=np.uint8)
    assert np.array_equal(orc.bgr2gray(bgr), cv2.cvtColor(bgr, cv2.COLOR_BGR2GRAY))
    ycc = cv2.cvtColor(bgr, cv2.COLOR_BGR2YCrCb)
    assert np.array_equal(orc.bgr2ycrcb(bgr), ycc)
    assert np.array_equal(orc.inrange3(ycc, (0, 133, 77), (255, 173, 127)), cv2.inRange(ycc, (0, 133, 77), (255, 173, 127)))
    g = synth.gray_frame(640, 360, 6, 1)
    assert np.array_equal(orc.equalize_hist(g), cv2.equalizeHist(g))
    flat = np.full((50, 70), 9, np.uint8)
    assert np.array_equal(orc.equalize_hist(flat), cv2.equalizeHist(flat))        # constant image: unchanged
    for dw, dh in [(397, 224), (100, 57), (639, 359), (640, 360), (38, 21)]:
        assert np.array_equal(orc.resize_exact(g, dw, dh), cv2.resize(g, (dw, dh), interpolation=cv2.INTER_LINEAR_EXACT))
    S, SQ = orc.integral(g)
    s2, q2 = cv2.integral2(g, sdepth=cv2.CV_32S, sqdepth=cv2.CV_32S)
    assert np.array_equal(S, s2) and np.array_equal(SQ, q2)


def test_inter_area_matches_cv2(orc):
    """8(f) N1: the INTER_AREA restatement (integral and rational factors, 1 and 3 channels) vs cv2.resize."""
    import cv2
    rng = np.random.default_rng(3)
    for shape, (dw, dh) in [((720, 1280, 3), (640, 360)), ((360, 640, 3), (213, 120)), ((90, 160, 3), (107, 60)),
                            ((90, 160), (100, 50)), ((97, 161, 3), (64, 36)), ((200, 300, 3), (300, 200)), ((200, 300, 3), (100, 50))]:
        src = rng.integers(0, 256, shape, dtype=np.uint8)
        assert np.array_equal(orc.resize_area(src, dw, dh), cv2.resize(src, (dw, dh), interpolation=cv2.INTER_AREA)), (shape, dw, dh)


def test_integral_wraps_like_cv2(orc):
    import cv2
    g = np.full((720, 1280), 255, np.uint8)
    S, SQ = orc.integral(g)
    s2, q2 = cv2.integral2(g, sdepth=cv2.CV_32S, sqdepth=cv2.CV_32S)
    assert (q2 < 0).any()
    assert np.array_equal(S, s2) and np.array_equal(SQ, q2)


def test_tilted_integral_matches_cv2(orc):
    """8(f) N4: the tilted integral restatement vs the third output of cv2.integral3 (CV_32S)."""
    import cv2
    rng = np.random.default_rng(1)
    for shape in [(1, 1), (1, 5), (5, 1), (7, 9), (270, 480), (447, 795)]:
        g = rng.integers(0, 256, shape, dtype=np.uint8)
        assert np.array_equal(orc.integral_tilted(g), cv2.integral3(g, sdepth=cv2.CV_32S, sqdepth=cv2.CV_32S)[2]), shape
    g = np.full((720, 1280), 255, np.uint8)
    assert np.array_equal(orc.integral_tilted(g), cv2.integral3(g, sdepth=cv2.CV_32S, sqdepth=cv2.CV_32S)[2])


@pytest.mark.parametrize("path", all_cv2_haar_cascades(), ids=lambda p: os.path.basename(p)[12:-4])
def test_every_cv2_haar_cascade_matches_cv2(orc, cv2ref, synth, tmp_path, path):
    """8(f) N4: any Haar cascade cv2 can load -- tilted features, non-square windows, files without a
    type_id -- gives the same raw windows as cv2; on a 4-stage truncation (thousands of accepted windows)
    also the same final-stage sums, bitwise."""
    import cv2
    g = synth.gray_frame(320, 180, 3, 7)
    for p, cap in ((path, 1 << 16), (truncate_cascade(path, 4, tmp_path / "t4.xml"), 1 << 18)):
        oc = orc.OracleCascade.load(p)
        cc = cv2.CascadeClassifier(p)
        raw, sums, _ = orc.detect_raw(oc, g, cap=cap)
        r3, _lv, wt = cc.detectMultiScale3(g, scaleFactor=1.1, minNeighbors=0, minSize=(30, 30), outputRejectLevels=True)
        a = sorted(zip(map(tuple, _clip(raw, 320, 180).tolist()), sums.tolist()))
        b = sorted(zip(map(tuple, np.asarray(r3).reshape(-1, 4).tolist()), np.asarray(wt).reshape(-1).tolist()))
        assert a == b, p
    assert len(a) > 100                          # the truncated cascade really accepted many windows


def test_window_grid_matches_cv2_row_stripes(orc, tmp_path):
    """The rows cv2 visits: every level is cut into ns = ceil(first level's working width / 32) stripes of
    ceil(floor(ny / step) / ns) * step rows, so some levels lose their last grid row (320x180 does, at the
    39-pixel window).  A pass-everything cascade on noise enumerates the grid; compare it with cv2's."""
    import cv2
    rng = np.random.default_rng(11)
    sizes = [(320, 180), (201, 202), (640, 360)] + [(int(rng.integers(60, 700)), int(rng.integers(60, 400))) for _ in range(9)]
    lost = 0
    for win in ((20, 20), (24, 24)):
        p = pass_everything_cascade(tmp_path / f"all{win[0]}.xml", win)
        oc = orc.OracleCascade.load(p)
        cc = cv2.CascadeClassifier(p)
        for (w, h) in sizes:
            sf, ms = float(rng.choice([1.05, 1.1, 1.25])), [(0, 0), (30, 30), (50, 50)][int(rng.integers(0, 3))]
            g = rng.integers(0, 256, (h, w), dtype=np.uint8)
            raw, _, ctr = orc.detect_raw(oc, g, sf, ms, cap=1 << 20)
            want = cc.detectMultiScale(g, scaleFactor=sf, minNeighbors=0, minSize=ms)
            assert rectset(_clip(raw, w, h)) == rectset(np.asarray(want).reshape(-1, 4)), (win, w, h, sf, ms)
            full = sum(len(range(0, lh + 1 - win[1], 1 if s >= 2 else 2)) * len(range(0, lw + 1 - win[0], 1 if s >= 2 else 2))
                       for s in orc.scale_list(w, h, win[0], win[1], sf, ms)
                       for lw, lh in [orc.level_size(w, h, float(s))])
            lost += full - ctr["grid_windows"]
    assert lost > 0                              # the rule really removed rows somewhere in this sample


def test_group_rectangles_matches_cv2(orc):
    import cv2
    rng = np.random.default_rng(3)
    for trial in range(12):
        centers = rng.integers(40, 500, (5, 2))
        rects = []
        for _ in range(int(rng.integers(0, 60))):
            c = centers[rng.integers(0, 5)]
            s = int(rng.integers(30, 90))
            rects.append([int(c[0] + rng.integers(-6, 7)), int(c[1] + rng.integers(-6, 7)), s, s])
        got, _ = orc.group_rectangles(rects, 2, 0.2)
        want, _w = cv2.groupRectangles([list(r) for r in rects], 2, 0.2)
        assert rectset(got) == rectset(np.asarray(want).reshape(-1, 4))
    assert len(orc.group_rectangles([], 2)[0]) == 0


@pytest.mark.parametrize("name", CASCADES)
def test_raw_windows_sums_and_groups_match_cv2(orc, cv2ref, synth, name):
    import cv2
    oc = orc.OracleCascade.load(cv2ref.cascade_path(name))
    cc = cv2ref.load(cv2ref.cascade_path(name))
    g = synth.gray_frame(480, 270, 5, 7)
    raw, sums, ctr = orc.detect_raw(oc, g)
    r0 = cv2ref.detect_multiscale_cv2(cc, g, 1.1, 0)
    assert rectset(_clip(raw, 480, 270)) == rectset(r0)
    assert ctr["windows_evaluated"] <= ctr["grid_windows"]
    # final-stage sums, bitwise, through detectMultiScale3(outputRejectLevels)
    r3, lv, wt = cc.detectMultiScale3(g, scaleFactor=1.1, minNeighbors=0, minSize=(30, 30), outputRejectLevels=True)
    a = sorted(zip(map(tuple, _clip(raw, 480, 270).tolist()), sums.tolist()))
    b = sorted(zip(map(tuple, np.asarray(r3).reshape(-1, 4).tolist()), np.asarray(wt).reshape(-1).tolist()))
    assert a == b
    got, _ = orc.detect_multiscale(oc, g)
    assert rectset(got) == rectset(cv2ref.detect_multiscale_cv2(cc, g))


def test_detect_faces_matches_cv2(orc, cv2ref, synth):
    oc = orc.OracleCascade.load(cv2ref.cascade_path("alt2"))
    cc = cv2ref.load(cv2ref.cascade_path("alt2"))
    for seed in (0, 3):
        bgr = synth.bgr_frame(640, 360, 6, seed)
        got, _ = orc.detect_faces(oc, bgr)
        assert rectset(got) == rectset(cv2ref.detect_faces_cv2(cc, bgr))


def test_oracle_against_golden_fixtures(orc, cv2ref, synth):
    with open(GOLDEN) as fh:
        gold = json.load(fh)
    assert len(gold["cases"]) >= 10
    for c in gold["cases"]:
        oc = orc.OracleCascade.load(cv2ref.cascade_path(c["cascade"]))
        ms = tuple(c["min_size"])
        if c["kind"] == "gray":
            f = synth.gray_frame(c["w"], c["h"], c["nfaces"], c["seed"])
            assert hashlib.sha1(f.tobytes()).hexdigest() == c["frame_sha1"], "synthetic frame generator drifted"
            raw, _, _ = orc.detect_raw(oc, f, c["scale"], ms)
            raw = _clip(raw, c["w"], c["h"])
            assert len(raw) == c["raw_count"]
            assert hashlib.sha1(np.array(rectset(raw), np.int32).tobytes()).hexdigest() == c["raw_sha1"]
            got, _ = orc.detect_multiscale(oc, f, c["scale"], 2, ms)
        else:
            f = synth.bgr_frame(c["w"], c["h"], c["nfaces"], c["seed"])
            assert hashlib.sha1(f.tobytes()).hexdigest() == c["frame_sha1"]
            got, _ = orc.detect_faces(oc, f, c["scale"], 2, ms)
        assert rectset(got) == rectset(c["rects"]), c


def test_inexact_stage_sums_match_cv2_bitwise(orc, cv2ref, synth, tmp_path):
    """Pins the ORDER of the oracle's stage sums: on a cascade whose sums are not exactly representable
    (conftest.inexact_cascade) the final-stage sums equal cv2's detectMultiScale3 weights bit for bit."""
    import cv2
    from conftest import inexact_cascade
    p = inexact_cascade(cv2ref.cascade_path("alt"), tmp_path / "inexact.xml")
    oc = orc.OracleCascade.load(p)
    cc = cv2.CascadeClassifier(p)
    g = synth.gray_frame(640, 360, 6, 3)
    raw, sums, _ = orc.detect_raw(oc, g)
    r, _, wt = cc.detectMultiScale3(g, scaleFactor=1.1, minNeighbors=0, minSize=(30, 30), outputRejectLevels=True)
    assert len(raw) == len(r) > 0
    assert sorted(sums.tolist()) == sorted(np.asarray(wt).reshape(-1).tolist())


# ---- the OpenCV-2.4 legacy restatement (oracle/haar_legacy.c).  Its evaluator is UNPINNED (no OpenCV 2.4
# exists here); the pieces cv2 4.13 still computes the same way are pinned below. ----

def test_legacy_resize_and_integrals_match_cv2(synth):
    """cvResize's 11-bit INTER_LINEAR survives unchanged as cv2.INTER_LINEAR for 8-bit images, and the int32 / double
    integral pair is cv2.integral2(CV_32S, CV_64F): both legacy building blocks are bit-equal to cv2 4.13."""
    import cv2
    from oracle import legacy
    g = synth.gray_frame(640, 360, 6, 0)
    for dw, dh in ((582, 327), (291, 164), (100, 57), (397, 224), (640, 360), (33, 19)):
        assert np.array_equal(legacy.resize_linear(g, dw, dh), cv2.resize(g, (dw, dh), interpolation=cv2.INTER_LINEAR)), (dw, dh)
    small = legacy.resize_linear(g, 291, 164)
    S, SQ = legacy.integral(small)
    s2, q2 = cv2.integral2(small, sdepth=cv2.CV_32S, sqdepth=cv2.CV_64F)
    assert np.array_equal(S, s2) and np.array_equal(SQ, q2)
    bgr = synth.bgr_frame(320, 180, 2, 1)
    y = cv2.cvtColor(bgr, cv2.COLOR_BGR2YCrCb)[:, :, 0]                   # the 14-bit coefficients are what 4.x keeps for Y
    assert np.array_equal(legacy.bgr2gray(bgr), y)


def test_legacy_stage_tree_is_walked_as_a_tree(orc, cv2ref, synth, tmp_path):
    """The <parent>/<next> tree: on a crafted three-stage tree every window is accepted through the `next` branch, while
    the 4.x evaluation in file order rejects everything; on the real alt_tree file the two branches after stage 4 are
    both walked (a window failing in the first moves on to the second)."""
    from conftest import stage_tree_cascade
    from oracle import legacy
    p = stage_tree_cascade(tmp_path / "tree.xml")
    oc = orc.OracleCascade.load(p)
    assert oc.flat.stage_parent.tolist() == [-1, 0, 0] and oc.flat.stage_next.tolist() == [-1, 2, -1]
    g = synth.gray_frame(160, 120, 1, 0)
    raw, ctr, smap = legacy.detect_raw(oc, g, want_map=True)
    assert len(raw) == ctr["grid_windows"] > 0 and (smap == 3).all()       # every window accepted (through stage 2)
    raw4, _, ctr4 = orc.detect_raw(oc, g)
    assert len(raw4) == 0                                                   # file order: stage 1 kills everything
    ot = orc.OracleCascade.load(cv2ref.cascade_path("alt_tree"))
    if ot.flat.stage_next.max() >= 0:                                       # the old-format file (the new format has no tree)
        g2 = synth.gray_frame(640, 360, 6, 0)
        _, _, m2 = legacy.detect_raw(ot, g2, want_map=True)
        last = set(np.unique(m2).tolist())
        # a window that fails in branch A (5, 7, 9, ... 39) moves on to branch B (6, 8, ... 46): no window can END in A
        assert not (last & set(range(5, 40, 2))) and (last & {6, 8, 10, 12})
        assert ot.flat.nstages in last                                      # and some are accepted


def test_legacy_scale_list(orc):
    from oracle import legacy
    f = legacy.scale_list(1280, 720, 20, 20)
    assert len(f) > 25 and f[0] >= 1.5 and np.allclose(f[1:] / f[:-1], 1.1)
    assert int(np.rint(20 * f[0])) >= 30 > int(np.rint(20 * f[0] / 1.1))
