"""Parity at the BASELINE sizes, per window and per level (VERDICT r1 task 1c): the 128-column grid-step-1 tiles, the
two-word survivor masks and the deep pyramids only show up at 720p / 1080p / 4K.  Every grid position's fate
(x-skipped, low variance, rejected at stage k, accepted) against the oracle's stage map; every level's image and
integral planes against cv2; plus the robustness cases of ADVICE r1 (two cascades on one device, inexact stage sums,
truncated output lists)."""
import numpy as np
import pytest

from conftest import inexact_cascade, long_failure_runs_cascade, rectset

pytestmark = pytest.mark.gpu


def _stage_map_case(ug, orc, cv2ref, synth, name, w, h, nfaces, seed, sf=1.1, ms=(30, 30), mx=(0, 0), check_cv2=True):
    casc = ug.Cascade(cv2ref.cascade_path(name))
    oc = orc.OracleCascade.load(cv2ref.cascade_path(name))
    det = ug.Detector(casc, ug.default_params(scale_factor=sf, min_size=ms, max_size=mx, max_raw_per_frame=1 << 16),
                      max_size=(w, h), max_batch=1)
    det.set_stage_map(True)
    g = synth.gray_frame(w, h, nfaces, seed)
    got = det.detect_gray(g)
    raw, sums = det.raw()
    smap = det.stage_map()
    ctr = det.counters()
    oraw, osums, octr, omap = orc.detect_raw(oc, g, sf, ms, mx, want_map=True)
    assert len(smap) == len(omap) == octr["grid_windows"]
    bad = np.nonzero(smap != omap)[0]
    assert len(bad) == 0, f"{name} {w}x{h}: {len(bad)} window decisions differ, first at {bad[:5]}: {smap[bad[:5]]} vs {omap[bad[:5]]}"
    for k in ("levels", "grid_windows", "windows_evaluated", "lowvar"):
        assert ctr[k] == octr[k], (name, k, ctr[k], octr[k])
    a = sorted(zip(map(tuple, raw.tolist()), sums.tolist()))
    b = sorted(zip(map(tuple, oraw.tolist()), osums.tolist()))
    assert a == b and len(a) > 0
    if check_cv2:
        cc = cv2ref.load(cv2ref.cascade_path(name))
        assert rectset(got) == rectset(cv2ref.detect_multiscale_cv2(cc, g, sf, 2, ms, mx if mx != (0, 0) else None))
    det.close()
    return ctr


@pytest.mark.parametrize("name", ["alt", "alt2", "alt_tree"])
def test_stage_map_720p(ug, orc, cv2ref, synth, name):
    """BASELINE configs 1, 2, 4: one 1280x720 frame, 1.19 M grid positions each."""
    ctr = _stage_map_case(ug, orc, cv2ref, synth, name, 1280, 720, 8, 2000)
    assert ctr["levels"] == 33 and ctr["grid_windows"] == 1185515


def test_stage_map_1080p_default(ug, orc, cv2ref, synth):
    """BASELINE config 3: frontalface_default (24x24 window) at 1920x1080, 3.26 M grid positions."""
    ctr = _stage_map_case(ug, orc, cv2ref, synth, "default", 1920, 1080, 12, 2001)
    assert ctr["levels"] == 37


@pytest.mark.parametrize("ms,mx", [((24, 24), (30, 30)), ((48, 48), (120, 120))], ids=["step2-levels", "step1-levels"])
def test_stage_map_4k_level_subsets(ug, orc, cv2ref, synth, ms, mx):
    """BASELINE config 5 (3840x2160, scale 1.05, min 24) on two level subsets picked with min/max size: the five
    largest grid-step-2 levels (5.7 M positions) and nineteen grid-step-1 levels (12.4 M positions)."""
    _stage_map_case(ug, orc, cv2ref, synth, "alt", 3840, 2160, 24, 5, sf=1.05, ms=ms, mx=mx, check_cv2=False)


@pytest.mark.parametrize("w,h,sf,ms,name", [(1280, 720, 1.1, (30, 30), "alt"), (1920, 1080, 1.1, (30, 30), "default"),
                                            (3840, 2160, 1.05, (24, 24), "alt")], ids=["720p", "1080p", "4k-sf1.05"])
def test_every_level_image_and_integrals(ug, cv2ref, synth, w, h, sf, ms, name):
    """a7 / a8 at full size: EVERY pyramid level's image, S and SQ against cv2.resize(INTER_LINEAR_EXACT) and
    cv2.integral2 (int32, wrapping) -- 33, 37 and 92 levels."""
    import cv2
    casc = ug.Cascade(cv2ref.cascade_path(name))
    det = ug.Detector(casc, ug.default_params(scale_factor=sf, min_size=ms), max_size=(w, h), max_batch=1)
    g = synth.gray_frame(w, h, 8, 77)
    g[: h // 3] = np.maximum(g[: h // 3], 240)              # a bright band makes SQ wrap on the large levels
    det.detect_gray(g)
    sc, lw, lh, _ = det.plan_info(w, h)
    assert len(sc) == {"720p": 33, "1080p": 37}.get(f"{h}p", 92)
    wrapped = False
    for i in range(len(sc)):
        img, S, SQ = det.level(i, int(lw[i]), int(lh[i]))
        ref = cv2.resize(g, (int(lw[i]), int(lh[i])), interpolation=cv2.INTER_LINEAR_EXACT)
        assert np.array_equal(img, ref), f"level {i} image"
        s2, q2 = cv2.integral2(ref, sdepth=cv2.CV_32S, sqdepth=cv2.CV_32S)
        assert np.array_equal(S, s2), f"level {i} S"
        assert np.array_equal(SQ, q2), f"level {i} SQ"
        wrapped |= bool((q2 < 0).any())
    assert wrapped
    det.close()


def test_two_cascades_on_one_device_do_not_corrupt_each_other(ug, cv2ref, synth):
    """the dense-stage nodes live in one __constant__ image per device.  Detector A (alt) has a
    submitted batch in flight while detector B (default, another cascade) runs; both must still match their own
    stand-alone results, whichever order the streams run in.  Repeated so that the image really ping-pongs."""
    ca, cb = ug.Cascade(cv2ref.cascade_path("alt")), ug.Cascade(cv2ref.cascade_path("default"))
    A = ug.Detector(ca, ug.default_params(), max_size=(640, 360), max_batch=4)
    B = ug.Detector(cb, ug.default_params(), max_size=(640, 360), max_batch=4)
    C = ug.Detector(ca, ug.default_params(), max_size=(640, 360), max_batch=4)       # same cascade as A: shares the image
    fa = synth.gray_batch(640, 360, 24, seed0=1200, nfaces=5)
    fb = synth.gray_batch(640, 360, 6, seed0=1300, nfaces=5)
    want_a = A.detect_batch(fa)
    want_b = [B.detect_gray(f) for f in fb]
    assert sum(map(len, want_a)) > 0 and sum(map(len, want_b)) > 0
    cc = cv2ref.load(cv2ref.cascade_path("default"))
    assert rectset(want_b[0]) == rectset(cv2ref.detect_multiscale_cv2(cc, fb[0]))
    for rep in range(3):
        A.submit_batch(fa)
        got_b = [B.detect_gray(f) for f in fb]                                     # takes the constant bank over mid-flight
        got_c = C.detect_batch(fa[:5])
        got_a = A.wait_batch()
        for x, y in zip(want_a, got_a):
            assert np.array_equal(x, y), rep
        for x, y in zip(want_b, got_b):
            assert np.array_equal(x, y), rep
        for x, y in zip(want_a[:5], got_c):
            assert np.array_equal(x, y), rep
    for d in (A, B, C):
        d.close()


def test_inexact_stage_sums_follow_file_order(ug, orc, cv2ref, synth, tmp_path):
    """a cascade whose stage sums do not fit fp64 exactly (sum_exact == 0) must be summed in file order.
    Stage maps and final-stage sums (bitwise) against the oracle with most stages in the per-lane late evaluator
    (dense_stages 2) and with the default split; final rectangles against cv2 on the same file."""
    import cv2
    p = inexact_cascade(cv2ref.cascade_path("alt"), tmp_path / "inexact.xml")
    casc = ug.Cascade(p)
    assert casc.info.sum_exact == 0
    oc = orc.OracleCascade.load(p)
    g = synth.gray_frame(640, 360, 6, 3)
    oraw, osums, octr, omap = orc.detect_raw(oc, g, want_map=True)
    assert len(oraw) > 0
    for dense in (2, 0):
        det = ug.Detector(casc, ug.default_params(dense_stages=dense), max_size=(640, 360), max_batch=1)
        det.set_stage_map(True)
        got = det.detect_gray(g)
        raw, sums = det.raw()
        assert np.array_equal(det.stage_map(), omap), dense
        assert sorted(zip(map(tuple, raw.tolist()), sums.tolist())) == sorted(zip(map(tuple, oraw.tolist()), osums.tolist())), dense
        cc = cv2.CascadeClassifier(p)
        assert rectset(got) == rectset(cc.detectMultiScale(g, scaleFactor=1.1, minNeighbors=2, minSize=(30, 30)))
        det.close()


@pytest.mark.parametrize("name", ["alt", "alt2"])
def test_queue_takes_every_survivor_of_stage_0(ug, orc, cv2ref, synth, name):
    """dense_stages = 1 sends every window that passes stage 0 (three quarters of the grid: thousands per tile) to the
    block-local queue: the survivor list of stump cascades then overflows the dead mask buffers it lives in and is drained
    in passes, tree cascades pick every window out of the masks; both on the 128 x 32 and the 64 x 16 tiles of a 1280x720
    frame.  Per-window decisions and final-stage sums against the oracle."""
    casc = ug.Cascade(cv2ref.cascade_path(name))
    oc = orc.OracleCascade.load(cv2ref.cascade_path(name))
    g = synth.gray_frame(1280, 720, 8, 77)
    det = ug.Detector(casc, ug.default_params(dense_stages=1, max_raw_per_frame=1 << 16), max_size=(1280, 720), max_batch=1)
    det.set_stage_map(True)
    got = det.detect_gray(g)
    raw, sums = det.raw()
    ctr = det.counters()
    oraw, osums, octr, omap = orc.detect_raw(oc, g, want_map=True)
    assert np.array_equal(det.stage_map(), omap)
    assert sorted(zip(map(tuple, raw.tolist()), sums.tolist())) == sorted(zip(map(tuple, oraw.tolist()), osums.tolist())) and len(raw) > 0
    assert ctr["queued"] == int((omap >= 1).sum())                  # every stage-0 survivor went through the queue
    assert rectset(got) == rectset(orc.detect_multiscale(oc, g)[0])
    det.close()


@pytest.mark.parametrize("name", ["alt", "alt2"])
def test_stage_map_720p_in_a_batch(ug, orc, cv2ref, synth, name):
    """Small calls (below 6000 row-walking scan blocks: every one-frame call) run one scan block per tile, with the x-skip state
    at the tile's left edge from a backward pass over stage 0; large groups walk whole rows.  The same frame in a group of 44
    (6336 blocks) takes the row-walking path: per-window decisions (frame 0) against the oracle, every list equal to the
    one-frame call's (the chunked path)."""
    casc = ug.Cascade(cv2ref.cascade_path(name))
    oc = orc.OracleCascade.load(cv2ref.cascade_path(name))
    g = synth.gray_frame(1280, 720, 8, 2000)
    det = ug.Detector(casc, ug.default_params(max_raw_per_frame=1 << 16), max_size=(1280, 720), max_batch=44)
    single = det.detect_gray(g)
    det.set_stage_map(True)
    lists = det.detect_batch(np.stack([g] * 44))
    smap = det.stage_map()
    raw, sums = det.raw()
    oraw, osums, octr, omap = orc.detect_raw(oc, g, want_map=True)
    assert np.array_equal(smap, omap)
    assert sorted(zip(map(tuple, raw.tolist()), sums.tolist())) == sorted(zip(map(tuple, oraw.tolist()), osums.tolist()))
    for f in range(44):
        assert rectset(lists[f]) == rectset(single)
    assert len(single) > 0
    det.close()


@pytest.mark.parametrize("thr", [0.002, 1e6], ids=["runs-of-dozens", "every-window-fails"])
def test_x_skip_across_tile_boundaries(ug, orc, synth, tmp_path, thr):
    """The x-skip rule carries the parity of a run of stage-0 failures along the row -- in a register when a block walks the row,
    through a backward pass over stage 0 when a block owns one tile (small calls).  A crafted cascade whose stage 0 fails in runs
    of dozens of windows (or at every window: the backward pass then walks to the start of the row) against the oracle's stage
    map, for a one-frame call (one block per tile) and for frame 0 of a group of 44 (row-walking blocks), at 1280x720."""
    p = long_failure_runs_cascade(tmp_path / "runs.xml", thr=thr)
    casc = ug.Cascade(p)
    oc = orc.OracleCascade.load(p)
    g = synth.gray_frame(1280, 720, 8, 31)
    oraw, osums, octr, omap = orc.detect_raw(oc, g, want_map=True, cap=1 << 21)
    assert (omap == -2).mean() > 0.25                                                   # a third of the grid is x-skipped
    det = ug.Detector(casc, ug.default_params(max_raw_per_frame=64), max_size=(1280, 720), max_batch=44)
    det.set_stage_map(True)
    for batch in (1, 44):
        try:
            if batch == 1:
                det.detect_gray(g)
            else:
                det.detect_batch(np.stack([g] * batch))
            assert len(oraw) <= 64
        except ug.UgdError as e:                                                        # hundreds of thousands of accepted windows:
            assert e.code == ug.ERR_CAPACITY and len(oraw) > 64                         # the candidate list overflows, the map is complete
        assert np.array_equal(det.stage_map(), omap), batch
        ctr = det.counters()
        assert ctr["windows_evaluated"] == octr["windows_evaluated"] * batch
    det.close()


def test_truncated_lists_are_reported_not_cut(ug, cv2ref, synth):
    """more rectangles than `cap` is UGD_ERR_TRUNCATED with the true counts, and the wrappers fetch again."""
    import ctypes as C
    casc = ug.Cascade(cv2ref.cascade_path("alt"))
    det = ug.Detector(casc, ug.default_params(min_neighbors=0), max_size=(640, 360), max_batch=2)
    g = synth.gray_frame(640, 360, 6, 21)
    full = det.detect_gray(g, cap=8192)
    assert len(full) > 16
    out = (ug.Rect * 16)()
    n = C.c_int()
    rc = ug.lib().ugd_detect_gray(det._h, g.ctypes.data_as(C.c_void_p), 640, 360, C.c_size_t(640), out, 16, C.byref(n))
    assert rc == ug.ERR_TRUNCATED and n.value == len(full)
    assert len(det.detect_gray(g, cap=16)) == len(full)                            # wrapper retries with the true count
    both = det.detect_batch(np.stack([g, g]), cap=16)
    assert len(both[0]) == len(both[1]) == len(full)
    det.close()
