"""GPU parity tests proper: every kernel of the path against the CPU oracle (and live cv2 4.13),
called through the C ABI (ctypes -> libugdet.so).  Bit-exact: integers, bytes, decisions, rects."""
import os

import numpy as np
import pytest

from conftest import CASCADES, all_cv2_haar_cascades, pass_everything_cascade, rectset, truncate_cascade

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def dets(ug, cv2ref):
    out = {}
    for name in CASCADES:
        c = ug.Cascade(cv2ref.cascade_path(name))
        out[name] = (c, ug.Detector(c, ug.default_params(), max_size=(1280, 720), max_batch=4))
    yield out
    for _, d in out.values():
        d.close()


def test_pyramid_and_integrals_bit_exact(dets, synth, orc):
    import cv2
    _, det = dets["alt"]
    g = synth.gray_frame(640, 360, 6, 5)
    det.detect_gray(g)
    sc, lw, lh, st = det.plan_info(640, 360)
    want_sc = orc.scale_list(640, 360, 20, 20)
    assert np.array_equal(sc, want_sc)
    for i in range(len(sc)):
        assert (lw[i], lh[i]) == orc.level_size(640, 360, float(sc[i]))
        img, S, SQ = det.level(i, int(lw[i]), int(lh[i]))
        ref = cv2.resize(g, (int(lw[i]), int(lh[i])), interpolation=cv2.INTER_LINEAR_EXACT)
        assert np.array_equal(img, ref), f"level {i} image"
        assert np.array_equal(img, orc.resize_exact(g, int(lw[i]), int(lh[i])))
        s2, q2 = cv2.integral2(ref, sdepth=cv2.CV_32S, sqdepth=cv2.CV_32S)
        assert np.array_equal(S, s2), f"level {i} S"
        assert np.array_equal(SQ, q2), f"level {i} SQ"


def test_integral_wraps_like_cv2_at_720p(dets, synth):
    import cv2
    _, det = dets["alt"]
    g = np.full((720, 1280), 255, np.uint8)
    g[::7, ::5] = 3
    det.detect_gray(g)
    sc, lw, lh, _ = det.plan_info(1280, 720)
    img, S, SQ = det.level(0, int(lw[0]), int(lh[0]))
    s2, q2 = cv2.integral2(img, sdepth=cv2.CV_32S, sqdepth=cv2.CV_32S)
    assert (q2 < 0).any()                      # the squared integral really wraps here
    assert np.array_equal(S, s2) and np.array_equal(SQ, q2)


def test_preprocess_bit_exact(dets, synth, orc):
    import cv2
    _, det = dets["alt2"]
    rng = np.random.default_rng(1)
    for bgr in (synth.bgr_frame(640, 360, 6, 2), rng.integers(0, 256, (123, 157, 3), dtype=np.uint8)):
        h, w = bgr.shape[:2]
        det.detect_bgr(bgr)
        geq, mask = det.preprocessed(w, h)
        ycc = cv2.cvtColor(bgr, cv2.COLOR_BGR2YCrCb)
        m = cv2.inRange(ycc, (0, 133, 77), (255, 173, 127))
        want = cv2.bitwise_and(cv2.equalizeHist(cv2.cvtColor(bgr, cv2.COLOR_BGR2GRAY)), m)
        assert np.array_equal(mask, m)
        assert np.array_equal(geq, want)
        assert np.array_equal(mask, orc.inrange3(orc.bgr2ycrcb(bgr), (0, 133, 77), (255, 173, 127)))


@pytest.mark.parametrize("name", CASCADES)
def test_window_decisions_raw_and_grouped(dets, synth, orc, cv2ref, name):
    casc, det = dets[name]
    oc = orc.OracleCascade.load(cv2ref.cascade_path(name))
    det.set_stage_map(True)
    for seed in (0, 1):
        g = synth.gray_frame(640, 360, 6, seed)
        got = det.detect_gray(g)
        raw, sums = det.raw()
        smap = det.stage_map()
        ctr = det.counters()
        oraw, osums, octr, omap = orc.detect_raw(oc, g, want_map=True)
        # per-window accept/reject decisions, every grid position, bit-exact
        assert len(smap) == len(omap)
        bad = np.nonzero(smap != omap)[0]
        assert len(bad) == 0, f"{len(bad)} window decisions differ, first at {bad[:5]}: {smap[bad[:5]]} vs {omap[bad[:5]]}"
        assert ctr["windows_evaluated"] == octr["windows_evaluated"]
        assert ctr["grid_windows"] == octr["grid_windows"]
        assert ctr["lowvar"] == octr["lowvar"]
        # accepted windows + final-stage sums (bitwise)
        a = sorted(zip(map(tuple, raw.tolist()), sums.tolist()))
        b = sorted(zip(map(tuple, oraw.tolist()), osums.tolist()))
        assert a == b
        # grouped rectangles vs oracle and live cv2
        want, _ = orc.detect_multiscale(oc, g)
        assert rectset(got) == rectset(want)
        cc = cv2ref.load(cv2ref.cascade_path(name))
        assert rectset(got) == rectset(cv2ref.detect_multiscale_cv2(cc, g))
    det.set_stage_map(False)


def test_detect_faces_bgr_end_to_end(dets, synth, orc, cv2ref):
    casc, det = dets["alt2"]
    oc = orc.OracleCascade.load(cv2ref.cascade_path("alt2"))
    cc = cv2ref.load(cv2ref.cascade_path("alt2"))
    nonempty = 0
    for seed in range(4):
        bgr = synth.bgr_frame(640, 360, 6, seed)
        got = det.detect_bgr(bgr)
        want, _ = orc.detect_faces(oc, bgr)
        assert rectset(got) == rectset(want)
        assert rectset(got) == rectset(cv2ref.detect_faces_cv2(cc, bgr))
        nonempty += len(got) > 0
    assert nonempty > 0


def test_batch_equals_single_and_720p(dets, synth, orc, cv2ref):
    casc, det = dets["alt"]
    cc = cv2ref.load(cv2ref.cascade_path("alt"))
    frames = synth.gray_batch(1280, 720, 6, seed0=10)       # 6 frames > max_batch 4: two groups
    got = det.detect_batch(frames)
    for f in range(len(frames)):
        assert rectset(got[f]) == rectset(det.detect_gray(frames[f]))
        assert rectset(got[f]) == rectset(cv2ref.detect_multiscale_cv2(cc, frames[f]))


def test_pipelined_host_batches(dets, synth):
    """Host batches larger than max_batch run as a copy/compute pipeline over two staging buffers:
    9 frames = groups of 4, 4, 1 -- results equal the one-frame calls, for the gray and the BGR entry."""
    _, det = dets["alt2"]
    gray = synth.gray_batch(640, 360, 9, seed0=300, nfaces=5)
    got = det.detect_batch(gray)
    for f in range(9):
        assert rectset(got[f]) == rectset(det.detect_gray(gray[f]))
    bgr = np.stack([synth.bgr_frame(640, 360, 5, 400 + i) for i in range(9)])
    got = det.detect_batch(bgr)
    assert det.counters()["frames"] == 9
    for f in range(9):
        assert rectset(got[f]) == rectset(det.detect_bgr(bgr[f]))


def test_group_pipeline_device_batches(ug, synth, cv2ref):
    """Multi-group calls run as a two-stream pipeline (image stage of group g+1 next to the scan stage of group g, S / SQ /
    mask double-buffered).  Device-resident batches of 11 frames with max_batch 2 = six groups, gray and BGR (the BGR
    pointer deliberately misaligned so the realign copy joins the image stream), twice in a row on the same detector:
    every frame must equal its single-frame call, and the counters must add up."""
    import torch
    casc = ug.Cascade(cv2ref.cascade_path("alt2"))
    det = ug.Detector(casc, ug.default_params(), max_size=(640, 360), max_batch=2)
    gray = synth.gray_batch(640, 360, 11, seed0=900, nfaces=5)
    bgr = np.stack([synth.bgr_frame(640, 360, 5, 950 + i) for i in range(11)])
    want_g = [det.detect_gray(f) for f in gray]
    want_b = [det.detect_bgr(f) for f in bgr]
    dg = torch.from_numpy(gray).cuda()
    pad = torch.zeros(bgr.size + 16, dtype=torch.uint8, device="cuda")
    pad[1:1 + bgr.size] = torch.from_numpy(bgr).reshape(-1).cuda()          # data_ptr + 1: not 4-byte aligned
    torch.cuda.synchronize()
    for _ in range(2):
        got = det.detect_batch_device(dg.data_ptr(), False, 11, 640, 360)
        assert det.counters()["frames"] == 11
        for f in range(11):
            assert rectset(got[f]) == rectset(want_g[f])
        got = det.detect_batch_device(pad.data_ptr() + 1, True, 11, 640, 360)
        for f in range(11):
            assert rectset(got[f]) == rectset(want_b[f])
    assert sum(len(x) for x in want_g) > 0 and sum(len(x) for x in want_b) > 0
    det.close()


def test_submit_wait_prefetch(dets, synth, ug):
    """8(f) N2: submit returns before the batch is done, the host does other work, wait delivers the same lists as the
    blocking call; a second submit (or a blocking call) while one is pending is refused."""
    _, det = dets["alt"]
    a = synth.gray_batch(640, 360, 7, seed0=600, nfaces=5)
    b = synth.gray_batch(640, 360, 3, seed0=700, nfaces=5)
    want_a, want_b = det.detect_batch(a), det.detect_batch(b)
    det.submit_batch(a)
    with pytest.raises(ug.UgdError):
        det.detect_gray(b[0])
    host_work = sum(int(x.sum()) for x in b)                                   # the tracker's share, overlapped
    got_a = det.wait_batch()
    det.submit_batch(b)
    got_b = det.wait_batch()
    assert host_work > 0
    for w_, g_ in zip(want_a + want_b, got_a + got_b):
        assert np.array_equal(w_, g_)
    with pytest.raises(ug.UgdError):
        lib_rc = det._ck(ug.lib().ugd_wait_batch(det._h, None, None))


def test_inter_area_decimation_bit_exact(dets, synth, orc):
    """cv::resize(orig, scan.frame_size, INTER_AREA) (face_tracker.cc:112-114) on the device,
    bit-exact with cv2 and the oracle: the shipped cfg's 2x case, other integral factors, rational factors, gray."""
    import cv2
    _, det = dets["alt2"]
    rng = np.random.default_rng(7)
    cases = [((720, 1280, 3), (640, 360)), ((1080, 1920, 3), (640, 360)), ((720, 1280, 3), (427, 240)),
             ((720, 1280, 3), (854, 480)), ((97, 161, 3), (64, 36)), ((360, 640), (320, 180)), ((200, 300, 3), (150, 200))]
    for shape, (dw, dh) in cases:
        src = rng.integers(0, 256, shape, dtype=np.uint8)
        got = det.resize_area(src, dw, dh)
        assert np.array_equal(got, cv2.resize(src, (dw, dh), interpolation=cv2.INTER_AREA)), (shape, dw, dh)
        assert np.array_equal(got, orc.resize_area(src, dw, dh))
    with pytest.raises(Exception):
        det.resize_area(np.zeros((100, 100, 3), np.uint8), 200, 200)           # enlarging is not the reference's use


def test_scaled_detection_matches_reference_pipeline(dets, synth, orc, cv2ref):
    """Full-size BGR frames -> INTER_AREA to the cfg's 640x360 scan size -> detectFaces, as the tracker does
    (face_tracker.cc:112-114 then :295): equal to the oracle run on cv2-resized frames."""
    import cv2
    _, det = dets["alt2"]
    oc = orc.OracleCascade.load(cv2ref.cascade_path("alt2"))
    frames = np.stack([synth.bgr_frame(1280, 720, 8, 500 + i) for i in range(5)])      # 5 frames > max_batch 4
    got = det.detect_batch_scaled(frames, (640, 360))
    for f in range(5):
        small = cv2.resize(frames[f], (640, 360), interpolation=cv2.INTER_AREA)
        want, _ = orc.detect_faces(oc, small)
        assert rectset(got[f]) == rectset(want)
        assert rectset(got[f]) == rectset(det.detect_bgr(small))


def test_scaled_fused_decimation_images(ug, synth, orc, cv2ref):
    """The 2x2 decimation runs inside k_preprocess (no scan-size BGR frame is written): the equalised, masked gray image
    and the skin mask it leaves are those of the reference pipeline on the cv2-resized frame (detector.cc:13-22 after
    face_tracker.cc:112-114), byte for byte; sizes the fused form does not take (3x3 factor, scan width not a multiple of
    4, rational factor) go through k_area and give the same kind of agreement."""
    import cv2
    casc = ug.Cascade(cv2ref.cascade_path("alt"))
    oc = orc.OracleCascade.load(cv2ref.cascade_path("alt"))
    for (W, H), (sw, sh) in (((1280, 720), (640, 360)), ((1284, 720), (642, 360)), ((960, 540), (320, 180)), ((1280, 720), (854, 480))):
        det = ug.Detector(casc, ug.default_params(), max_size=(sw, sh), max_batch=2)
        frames = np.stack([synth.bgr_frame(W, H, 6, 900 + i) for i in range(3)])         # 3 frames > max_batch 2
        got = det.detect_batch_scaled(frames, (sw, sh))
        last_group_first = 2                                                             # frame 0 of the last launch group
        small = cv2.resize(frames[last_group_first], (sw, sh), interpolation=cv2.INTER_AREA)
        ycc = cv2.cvtColor(small, cv2.COLOR_BGR2YCrCb)
        mask = cv2.inRange(ycc, (0, 133, 77), (255, 173, 127))
        gray = cv2.equalizeHist(cv2.cvtColor(small, cv2.COLOR_BGR2GRAY)) & mask
        g, m = det.preprocessed(sw, sh)
        assert np.array_equal(m, mask), (W, H, sw, sh)
        assert np.array_equal(g, gray), (W, H, sw, sh)
        for f in range(3):
            want, _ = orc.detect_faces(oc, cv2.resize(frames[f], (sw, sh), interpolation=cv2.INTER_AREA))
            assert rectset(got[f]) == rectset(want), (W, H, sw, sh, f)
        det.close()


def test_crop_faces_from_device_frames(ug, synth, cv2ref):
    """SURVEY 8(f) N2, second half: the face patches of face_tracker.cc:296-302 (`cv::Mat img_face(curr_frame, rect)`, copied
    by Face's copy constructor, structure.h:137-146) cut out of device-resident frames == numpy slices of the same frames,
    byte for byte; BGR and gray, an empty list in the middle, and the two-call size query of the C ABI."""
    import ctypes as C
    import torch
    casc = ug.Cascade(cv2ref.cascade_path("alt"))
    det = ug.Detector(casc, ug.default_params(), max_size=(640, 360), max_batch=2)
    for channels in (3, 1):
        frames = [synth.bgr_frame(640, 360, 5, 700 + i) if channels == 3 else synth.gray_frame(640, 360, 5, 710 + i) for i in range(3)]
        frames[1] = np.zeros_like(frames[1])                                            # no detections on this one
        host = np.stack(frames)
        dev = torch.from_numpy(host).cuda()
        lists = det.detect_batch_device(dev.data_ptr(), channels == 3, 3, 640, 360, cap=32)
        assert lists.total() > 0 and len(lists[1]) == 0
        crops = det.crop_faces(dev.data_ptr(), 3, 640, 360, channels, lists)
        for f in range(3):
            assert len(crops[f]) == len(lists[f])
            for (x, y, w, h), patch in zip(lists[f].tolist(), crops[f]):
                assert np.array_equal(patch, host[f][y:y + h, x:x + w])
        # a buffer that is too small: nothing copied, the size reported
        cap = lists.rects.shape[1]
        offs = np.zeros(3 * cap, np.uint64); total = C.c_size_t(); small = np.zeros(16, np.uint8)
        rc = ug.lib().ugd_crop_faces(det._h, C.c_void_p(dev.data_ptr()), 3, 640, 360, channels, lists.rects.ctypes.data_as(C.c_void_p),
                                     lists.counts.astype(np.int32).ctypes.data_as(C.c_void_p), cap, small.ctypes.data_as(C.c_void_p),
                                     C.c_size_t(16), offs.ctypes.data_as(C.c_void_p), C.byref(total))
        assert rc == ug.ERR_CAPACITY and total.value == sum(int(w) * int(h) * channels for l in lists for (_, _, w, h) in l.tolist())
    det.close()


def test_randomized_call_sequences(dets, synth, orc, cv2ref):
    """Forty calls in a random order over the four shared detectors: frame sizes that make the plans, the buffers and the TMA
    tensor maps grow, shrink and come back; gray and BGR; single frames, batches below and above max_batch (one launch group,
    the group pipeline), device-resident batches, the submit/wait form and the fused decimation -- every list against the oracle.
    State that survives a call (plan cache, buffer sets, constant bank, pinned staging) is what this is after."""
    import cv2
    import torch
    rng = np.random.default_rng(20260101)
    ocs = {n: orc.OracleCascade.load(cv2ref.cascade_path(n)) for n in CASCADES}
    sizes = [(320, 180), (333, 217), (480, 270), (640, 360), (200, 400), (96, 64)]
    for it in range(40):
        name = CASCADES[int(rng.integers(len(CASCADES)))]
        _, det = dets[name]
        oc = ocs[name]
        w, h = sizes[int(rng.integers(len(sizes)))]
        kind = ["gray1", "bgr1", "gray_batch", "bgr_batch", "device", "submit", "scaled"][int(rng.integers(7))]
        seed = 3000 + it
        if kind == "gray1":
            g = synth.gray_frame(w, h, 3, seed)
            assert rectset(det.detect_gray(g)) == rectset(orc.detect_multiscale(oc, g)[0]), (it, name, kind, w, h)
        elif kind == "bgr1":
            f = synth.bgr_frame(w, h, 3, seed)
            assert rectset(det.detect_bgr(f)) == rectset(orc.detect_faces(oc, f)[0]), (it, name, kind, w, h)
        elif kind in ("gray_batch", "device", "submit"):
            n = int(rng.integers(1, 10))                                                  # max_batch is 4
            fr = np.stack([synth.gray_frame(w, h, 3, seed * 16 + i) for i in range(n)])
            if kind == "gray_batch":
                got = det.detect_batch(fr)
            elif kind == "device":
                dev = torch.from_numpy(fr).cuda()
                got = det.detect_batch_device(dev.data_ptr(), False, n, w, h)
            else:
                det.submit_batch(fr)
                got = det.wait_batch()
            for i in range(n):
                assert rectset(got[i]) == rectset(orc.detect_multiscale(oc, fr[i])[0]), (it, name, kind, w, h, i)
        elif kind == "bgr_batch":
            n = int(rng.integers(1, 7))
            fr = np.stack([synth.bgr_frame(w, h, 3, seed * 16 + i) for i in range(n)])
            got = det.detect_batch(fr)
            for i in range(n):
                assert rectset(got[i]) == rectset(orc.detect_faces(oc, fr[i])[0]), (it, name, kind, w, h, i)
        else:
            n = int(rng.integers(1, 7))
            W2, H2 = (2 * w, 2 * h) if rng.integers(2) else (3 * w, 3 * h)                 # fused 2x2, or k_area
            fr = np.stack([synth.bgr_frame(W2, H2, 3, seed * 16 + i) for i in range(n)])
            got = det.detect_batch_scaled(fr, (w, h))
            for i in range(n):
                small = cv2.resize(fr[i], (w, h), interpolation=cv2.INTER_AREA)
                assert rectset(got[i]) == rectset(orc.detect_faces(oc, small)[0]), (it, name, kind, w, h, i)


def test_edge_cases(dets, synth, ug):
    casc, det = dets["alt"]
    assert len(det.detect_gray(np.zeros((360, 640), np.uint8))) == 0          # flat: all low variance
    assert len(det.detect_gray(np.zeros((25, 25), np.uint8))) == 0            # smaller than minSize: no level
    g = synth.gray_frame(333, 217, 3, 4)                                      # ragged sizes
    det.detect_gray(g)
    with pytest.raises(ug.UgdError):
        det.detect_gray(np.zeros((721, 1280), np.uint8))                       # larger than created for


def test_parameter_variants_against_cv2(ug, synth, cv2ref):
    """The runtime parameters the reference hard-codes or reads from the cfg (detector.cc:25-31, structure.cc:83-85):
    minNeighbors 0 (raw list, clipped) and 4, maxSize, a different scale factor, non-square minSize -- vs live cv2."""
    casc = ug.Cascade(cv2ref.cascade_path("alt"))
    cc = cv2ref.load(cv2ref.cascade_path("alt"))
    g = synth.gray_frame(640, 360, 6, 21)
    variants = [dict(min_neighbors=0), dict(min_neighbors=4), dict(max_size=(120, 120)), dict(scale_factor=1.25),
                dict(min_size=(40, 60)), dict(scale_factor=1.05, min_size=(24, 24), max_size=(200, 200))]
    for v in variants:
        det = ug.Detector(casc, ug.default_params(**v), max_size=(640, 360), max_batch=1)
        got = det.detect_gray(g, cap=8192)
        want = cv2ref.detect_multiscale_cv2(cc, g, v.get("scale_factor", 1.1), v.get("min_neighbors", 2),
                                            v.get("min_size", (30, 30)), v.get("max_size"))
        assert len(want) > 0, v
        assert sorted(map(tuple, got.tolist())) == sorted(map(tuple, want.tolist())), v     # multiset (minNeighbors 0 repeats rects)
        det.close()


def test_capacity_errors_are_reported(ug, synth, cv2ref):
    """A candidate list that does not fit max_raw_per_frame is an error (UGD_ERR_CAPACITY), never a silent truncation;
    an output list longer than the caller's cap reports the true count."""
    casc = ug.Cascade(cv2ref.cascade_path("alt"))
    g = synth.gray_frame(640, 360, 6, 21)
    det = ug.Detector(casc, ug.default_params(max_raw_per_frame=16), max_size=(640, 360), max_batch=1)
    with pytest.raises(ug.UgdError):
        det.detect_gray(g)
    det.close()


def test_golden_fixtures_on_gpu(ug, synth, cv2ref):
    """The committed cv2-generated fixtures (tests/golden), replayed through the C ABI."""
    import hashlib
    import json
    import os
    gold = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "golden_detections.json")))
    dets = {}
    for c in gold["cases"]:
        key = (c["cascade"], c["scale"], tuple(c["min_size"]))
        if key not in dets:
            casc = ug.Cascade(cv2ref.cascade_path(c["cascade"]))
            dets[key] = (casc, ug.Detector(casc, ug.default_params(scale_factor=c["scale"], min_size=tuple(c["min_size"])),
                                           max_size=(1280, 720), max_batch=1))
        det = dets[key][1]
        if c["kind"] == "gray":
            f = synth.gray_frame(c["w"], c["h"], c["nfaces"], c["seed"])
            assert hashlib.sha1(f.tobytes()).hexdigest() == c["frame_sha1"]
            got = det.detect_gray(f)
            raw, _ = det.raw()
            x1 = np.minimum(raw[:, 0] + raw[:, 2], c["w"]); y1 = np.minimum(raw[:, 1] + raw[:, 3], c["h"])
            raw[:, 2] = x1 - raw[:, 0]; raw[:, 3] = y1 - raw[:, 1]
            assert len(raw) == c["raw_count"]
            assert hashlib.sha1(np.array(rectset(raw), np.int32).tobytes()).hexdigest() == c["raw_sha1"]
        else:
            f = synth.bgr_frame(c["w"], c["h"], c["nfaces"], c["seed"])
            got = det.detect_bgr(f)
        assert rectset(got) == rectset(c["rects"]), c
    for _, d in dets.values():
        d.close()


def test_full_size_configs_against_cv2(ug, synth, cv2ref):
    """BASELINE configs at full size: default@1080p (config 3) and the sf 1.05 / min 24 deep pyramid
    (config 5 parameters, on a 1080p frame to keep cv2 to seconds) -- final rectangles vs live cv2."""
    for name, w, h, sf, ms, nf in (("default", 1920, 1080, 1.1, (30, 30), 12), ("alt", 1920, 1080, 1.05, (24, 24), 12)):
        casc = ug.Cascade(cv2ref.cascade_path(name))
        det = ug.Detector(casc, ug.default_params(scale_factor=sf, min_size=ms), max_size=(w, h), max_batch=2)
        cc = cv2ref.load(cv2ref.cascade_path(name))
        frames = synth.gray_batch(w, h, 2, seed0=40, nfaces=nf)
        got = det.detect_batch(frames, cap=256)
        for i in range(2):
            assert rectset(got[i]) == rectset(cv2ref.detect_multiscale_cv2(cc, frames[i], sf, 2, ms))
        ctr = det.counters()
        assert ctr["windows_evaluated"] > 0 and ctr["raw"] > 0
        det.close()


def test_4k_deep_pyramid_against_cv2(ug, synth, cv2ref):
    """BASELINE config 5 at full size: one 3840x2160 frame, frontalface_alt, scale 1.05, min size 24
    (92 levels, 29 M grid windows) -- final rectangles vs live cv2, and the level/window counts of the plan."""
    casc = ug.Cascade(cv2ref.cascade_path("alt"))
    det = ug.Detector(casc, ug.default_params(scale_factor=1.05, min_size=(24, 24)), max_size=(3840, 2160), max_batch=1)
    cc = cv2ref.load(cv2ref.cascade_path("alt"))
    frame = synth.gray_frame(3840, 2160, 24, 5)
    got = det.detect_gray(frame, cap=1024)
    want = cv2ref.detect_multiscale_cv2(cc, frame, 1.05, 2, (24, 24))
    assert len(want) > 0 and rectset(got) == rectset(want)
    ctr = det.counters()
    assert ctr["levels"] == 92 and ctr["windows_evaluated"] <= ctr["grid_windows"]
    det.close()


def test_properties_at_bench_size(ug, synth, cv2ref):
    """Size-independent properties on a 720p batch: determinism (same batch twice), batch-order
    independence (frames permuted), idempotent grouping (output is canonically sorted), rects inside
    the frame, counters consistent."""
    casc = ug.Cascade(cv2ref.cascade_path("alt"))
    det = ug.Detector(casc, ug.default_params(), max_size=(1280, 720), max_batch=8)
    frames = synth.gray_batch(1280, 720, 8, seed0=70)
    a = det.detect_batch(frames)
    ctr = det.counters()
    b = det.detect_batch(frames)
    perm = np.array([3, 0, 7, 1, 6, 2, 5, 4])
    c = det.detect_batch(frames[perm])
    for i in range(8):
        assert np.array_equal(a[i], b[i])
        assert np.array_equal(c[i], a[perm[i]])
        r = a[i]
        assert (r[:, 0] >= 0).all() and (r[:, 1] >= 0).all() and (r[:, 0] + r[:, 2] <= 1280).all() and (r[:, 1] + r[:, 3] <= 720).all()
        assert [tuple(x) for x in r[:, [1, 0, 3, 2]].tolist()] == sorted(tuple(x) for x in r[:, [1, 0, 3, 2]].tolist())
    assert ctr["frames"] == 8 and ctr["levels"] == 33 and ctr["grid_windows"] == 8 * 1185515 or ctr["grid_windows"] > 0
    assert ctr["windows_evaluated"] <= ctr["grid_windows"] and ctr["grouped"] == sum(len(x) for x in a)
    det.close()


# ---- any Haar cascade cv2 can load (tilted features, non-square windows) -------------------

def test_tilted_integral_bit_exact(ug, synth):
    """k_tilted against the third output of cv2.integral3, every level, including a wrapping 720p frame."""
    import cv2
    c = ug.Cascade(cv2.data.haarcascades + "haarcascade_smile.xml")
    assert c.info.has_tilted == 1
    det = ug.Detector(c, ug.default_params(), max_size=(1280, 720), max_batch=2)
    bright = np.full((720, 1280), 255, np.uint8)
    bright[::7, ::5] = 3
    for g in (synth.gray_frame(640, 360, 6, 5), bright, synth.gray_frame(333, 217, 3, 1)):
        h, w = g.shape
        det.detect_gray(g)
        sc, lw, lh, _ = det.plan_info(w, h)
        for i in range(len(sc)):
            img, S, SQ = det.level(i, int(lw[i]), int(lh[i]))
            T = det.level_tilted(i, int(lw[i]), int(lh[i]))
            s2, q2, t2 = cv2.integral3(img, sdepth=cv2.CV_32S, sqdepth=cv2.CV_32S)
            assert np.array_equal(T, t2), f"{w}x{h} level {i} tilted integral"
            assert np.array_equal(S, s2) and np.array_equal(SQ, q2)
    det.close()
    # a cascade without tilted features never builds the plane
    c2 = ug.Cascade(cv2.data.haarcascades + "haarcascade_eye.xml")
    d2 = ug.Detector(c2, ug.default_params(), max_size=(320, 180), max_batch=1)
    d2.detect_gray(synth.gray_frame(320, 180, 2, 0))
    with pytest.raises(ug.UgdError) as e:
        d2.level_tilted(0, 10, 10)
    assert e.value.code == ug.ERR_UNSUPPORTED
    d2.close()


@pytest.mark.parametrize("path", all_cv2_haar_cascades(), ids=lambda p: os.path.basename(p)[12:-4])
def test_every_cv2_haar_cascade(ug, synth, orc, tmp_path, path):
    """All 17 Haar files cv2 ships -- 9 with tilted features, windows from 14x28 to 64x16 -- decision for
    decision against the oracle (itself pinned to cv2 on the same files): the full cascade on 640x360 and a
    4-stage truncation on 320x180 that accepts thousands of windows (final-stage sums bitwise)."""
    import cv2
    cases = ((path, synth.gray_frame(640, 360, 6, 3)),
             (truncate_cascade(path, 4, tmp_path / "t4.xml"), synth.gray_frame(320, 180, 3, 7)))
    for p, g in cases:
        h, w = g.shape
        c = ug.Cascade(p)
        oc = orc.OracleCascade.load(p)
        det = ug.Detector(c, ug.default_params(max_raw_per_frame=1 << 17), max_size=(w, h), max_batch=1)
        det.set_stage_map(True)
        got = det.detect_gray(g)
        raw, sums = det.raw(cap=1 << 17)
        smap = det.stage_map()
        ctr = det.counters()
        oraw, osums, octr, omap = orc.detect_raw(oc, g, want_map=True, cap=1 << 17)
        assert len(smap) == len(omap)
        bad = np.nonzero(smap != omap)[0]
        assert len(bad) == 0, f"{p}: {len(bad)} window decisions differ, first at {bad[:5]}: {smap[bad[:5]]} vs {omap[bad[:5]]}"
        for k in ("windows_evaluated", "grid_windows", "lowvar"):
            assert ctr[k] == octr[k], (p, k)
        if c.info.max_tree_nodes == 1:          # trees: the engine counts every node of a tree, the oracle the visited ones
            assert ctr["weak_evals"] == octr["weak_evals"], p
        a = sorted(zip(map(tuple, raw.tolist()), sums.tolist()))
        b = sorted(zip(map(tuple, oraw.tolist()), osums.tolist()))
        assert a == b, p
        want, _ = orc.detect_multiscale(oc, g)
        assert rectset(got) == rectset(want), p
        cc = cv2.CascadeClassifier(p)
        assert rectset(got) == rectset(cc.detectMultiScale(g, scaleFactor=1.1, minNeighbors=2, minSize=(30, 30)))
        det.close()


def test_window_grid_follows_cv2_row_stripes(ug, orc, tmp_path):
    """The stripe rule that drops the last grid row of some levels (oracle test of the same name): the engine's
    grid, enumerated with a pass-everything cascade, equals cv2's on sizes where rows are really lost."""
    import cv2
    rng = np.random.default_rng(11)
    p = pass_everything_cascade(tmp_path / "all.xml")
    c = ug.Cascade(p)
    cc = cv2.CascadeClassifier(p)
    for (w, h) in [(320, 180), (201, 202), (417, 233), (640, 360)]:
        g = rng.integers(0, 256, (h, w), dtype=np.uint8)
        det = ug.Detector(c, ug.default_params(min_neighbors=0, max_raw_per_frame=1 << 18), max_size=(w, h), max_batch=1)
        det.detect_gray(g, cap=16)
        raw, _ = det.raw(cap=1 << 18)
        want = np.asarray(cc.detectMultiScale(g, scaleFactor=1.1, minNeighbors=0, minSize=(30, 30))).reshape(-1, 4)
        clipped = raw.copy()
        clipped[:, 2] = np.minimum(raw[:, 0] + raw[:, 2], w) - raw[:, 0]
        clipped[:, 3] = np.minimum(raw[:, 1] + raw[:, 3], h) - raw[:, 1]
        assert rectset(clipped) == rectset(want), (w, h)
        det.close()
