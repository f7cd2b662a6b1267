"""SURVEY 8(f) N3 on the GPU: ugd_params.legacy_24 = 1 (OpenCV-2.4 evaluator, stage tree walked as a tree) against its
CPU restatement oracle/haar_legacy.c -- per-window outcome, candidate rects, grouped rects.  PARITY UNPINNED against the
real OpenCV 2.4 (none exists here): this is self-consistency, as SURVEY 8(f) says it can only be."""
import numpy as np
import pytest

from conftest import rectset, stage_tree_cascade

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name", ["alt", "alt_tree", "alt2", "default"])
def test_legacy_mode_matches_its_restatement(ug, orc, cv2ref, synth, name):
    from oracle import legacy
    path = cv2ref.cascade_path(name)
    casc = ug.Cascade(path)
    oc = orc.OracleCascade.load(path)
    det = ug.Detector(casc, ug.default_params(legacy_24=1, max_raw_per_frame=1 << 15), max_size=(640, 360), max_batch=2)
    det.set_stage_map(True)
    for seed in (0, 1):
        g = synth.gray_frame(640, 360, 6, seed)
        got = det.detect_gray(g)
        raw, _ = det.raw()
        smap = det.stage_map()
        oraw, octr, omap = legacy.detect_raw(oc, g, want_map=True)
        assert len(smap) == len(omap) == octr["grid_windows"]
        bad = np.nonzero(smap != omap)[0]
        assert len(bad) == 0, f"{name}: {len(bad)} window outcomes differ, first at {bad[:5]}: {smap[bad[:5]]} vs {omap[bad[:5]]}"
        assert rectset(raw) == rectset(oraw) and len(oraw) > 0
        want, _ = legacy.detect_multiscale(oc, g)
        assert rectset(got) == rectset(want) and len(want) > 0
        ctr = det.counters()
        assert ctr["windows_evaluated"] == octr["windows_evaluated"] and ctr["weak_evals"] == octr["weak_evals"]
        assert ctr["grid_windows"] == octr["grid_windows"] and ctr["levels"] == octr["levels"]
    # a batch (two groups) equals the single-frame calls
    frames = synth.gray_batch(640, 360, 3, seed0=5, nfaces=5)
    b = det.detect_batch(frames)
    for i in range(3):
        assert rectset(b[i]) == rectset(det.detect_gray(frames[i]))
    det.close()


def test_legacy_stage_tree_on_the_gpu(ug, orc, synth, tmp_path):
    """The crafted tree of the CPU test: every window accepted through the `next` branch in legacy mode, none in 4.x mode."""
    from oracle import legacy
    p = stage_tree_cascade(tmp_path / "tree.xml")
    casc = ug.Cascade(p)
    oc = orc.OracleCascade.load(p)
    g = synth.gray_frame(160, 120, 1, 0)
    oraw, octr = legacy.detect_raw(oc, g)
    d24 = ug.Detector(casc, ug.default_params(legacy_24=1, min_neighbors=0, max_raw_per_frame=1 << 16), max_size=(160, 120), max_batch=1)
    got = d24.detect_gray(g, cap=1 << 16)
    assert len(got) == octr["grid_windows"] > 0 and rectset(got) == rectset(oraw)      # minNeighbors 0: the raw list, unclipped
    d4 = ug.Detector(casc, ug.default_params(min_neighbors=0), max_size=(160, 120), max_batch=1)
    assert len(d4.detect_gray(g)) == 0
    d24.close(); d4.close()


def test_legacy_detect_faces_bgr(ug, orc, cv2ref, synth):
    """detector.cc:10-43 in legacy mode: 14-bit BGR2GRAY, equalizeHist, mask, legacy detectMultiScale, skin filter."""
    from oracle import legacy
    path = cv2ref.cascade_path("alt2")
    casc = ug.Cascade(path)
    oc = orc.OracleCascade.load(path)
    det = ug.Detector(casc, ug.default_params(legacy_24=1), max_size=(640, 360), max_batch=1)
    nonempty = 0
    for seed in range(3):
        bgr = synth.bgr_frame(640, 360, 6, seed)
        got = det.detect_bgr(bgr)
        want, _ = legacy.detect_faces(oc, bgr)
        assert rectset(got) == rectset(want)
        nonempty += len(want) > 0
    assert nonempty > 0
    det.close()
