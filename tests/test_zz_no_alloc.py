"""ugd_integral_noalloc (include/ugdet.h): the integral pass with its scratch in a static __device__ array, for callers
that cannot call cudaMalloc.  No counterpart in the reference.

Checked through the C ABI against the existing entry point for the same planes -- ugd_get_level after ugd_detect_gray
with minSize = the cascade window, so that pyramid level 0 is the frame itself -- bit for bit (integer work), on two
calls in a row that share the scratch: a 1280x720 frame, then a smaller, odd-sized one read through a padded stride.
This file is named to be collected last."""
import ctypes as C
import os

import numpy as np
import pytest

from conftest import ROOT

pytestmark = pytest.mark.gpu


def _engine_planes(ug, cv2ref, gray):
    """S and SQ of the frame from the existing entry points (detector-owned scratch)."""
    h, w = gray.shape
    casc = ug.Cascade(cv2ref.cascade_path("alt"))
    win = (casc.info.win_w, casc.info.win_h)
    det = ug.Detector(casc, ug.default_params(min_size=win, max_raw_per_frame=1 << 16), max_size=(w, h), max_batch=1)
    det.detect_gray(gray)
    sc, lw, lh, _ = det.plan_info(w, h)
    assert sc[0] == 1.0 and (int(lw[0]), int(lh[0])) == (w, h)
    img, S, SQ = det.level(0, w, h)
    det.close()
    assert np.array_equal(img, gray)
    return S, SQ


def _numpy_planes(gray):
    v = gray.astype(np.uint64)
    S = np.zeros((gray.shape[0] + 1, gray.shape[1] + 1), np.uint64)
    SQ = np.zeros_like(S)
    S[1:, 1:] = v.cumsum(0).cumsum(1)
    SQ[1:, 1:] = (v * v).cumsum(0).cumsum(1)
    return (S & 0xFFFFFFFF).astype(np.uint32).view(np.int32), (SQ & 0xFFFFFFFF).astype(np.uint32).view(np.int32)


def _call(lib, img_t, w, h, stride, S_t, SQ_t, stream=0):
    lib.ugd_integral_noalloc.argtypes = [C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_size_t, C.c_void_p, C.c_void_p,
                                         C.c_void_p]
    lib.ugd_integral_noalloc.restype = C.c_int
    return lib.ugd_integral_noalloc(0, C.c_void_p(img_t.data_ptr()), w, h, C.c_size_t(stride),
                                    C.c_void_p(S_t.data_ptr()), C.c_void_p(SQ_t.data_ptr()), C.c_void_p(stream))


def test_noalloc_integral_matches_engine_two_calls_in_a_row(ug, cv2ref, synth):
    import torch
    lib = ug.lib()
    a = synth.gray_frame(1280, 720, 6, 4242)
    b = np.ascontiguousarray(synth.gray_frame(1280, 720, 3, 4243)[100:579, 200:841])      # 641 x 479
    assert b.shape == (479, 641)
    want = [_engine_planes(ug, cv2ref, a), _engine_planes(ug, cv2ref, b)]
    for g, (S, SQ) in zip((a, b), want):                                                 # the engine vs plain numpy
        nS, nSQ = _numpy_planes(g)
        assert np.array_equal(S, nS) and np.array_equal(SQ, nSQ)

    # inputs and outputs are the caller's: allocated here, before the calls
    da = torch.from_numpy(a).cuda()
    stride_b = 704                                                                       # padded rows, filled with junk
    db = torch.full((479, stride_b), 255, dtype=torch.uint8, device="cuda")
    db[:, :641] = torch.from_numpy(b).cuda()
    outs = [(torch.full((721, 1281), -7, dtype=torch.int32, device="cuda"), torch.full((721, 1281), -7, dtype=torch.int32, device="cuda")),
            (torch.full((480, 642), -7, dtype=torch.int32, device="cuda"), torch.full((480, 642), -7, dtype=torch.int32, device="cuda"))]
    torch.cuda.synchronize()
    st = torch.cuda.current_stream().cuda_stream
    assert _call(lib, da, 1280, 720, 1280, *outs[0], stream=st) == 0                     # two calls in a row, same stream:
    assert _call(lib, db, 641, 479, stride_b, *outs[1], stream=st) == 0                  # the second reuses the scratch
    torch.cuda.synchronize()
    for (S_t, SQ_t), (S, SQ) in zip(outs, want):
        assert np.array_equal(S_t.cpu().numpy(), S)
        assert np.array_equal(SQ_t.cpu().numpy(), SQ)

    # and the first image again after the smaller one (stale strip sums of the other geometry in the scratch)
    S_t, SQ_t = outs[0]
    S_t.fill_(-7); SQ_t.fill_(-7)
    assert _call(lib, da, 1280, 720, 1280, S_t, SQ_t, stream=st) == 0
    torch.cuda.synchronize()
    assert np.array_equal(S_t.cpu().numpy(), want[0][0]) and np.array_equal(SQ_t.cpu().numpy(), want[0][1])


def test_noalloc_integral_limits(ug):
    import torch
    lib = ug.lib()
    hmax, wmax = 4096, 4096
    hdr = open(os.path.join(ROOT, "include", "ugdet.h")).read()
    assert "#define UGD_NOALLOC_MAX_W 4096" in hdr and "#define UGD_NOALLOC_MAX_H 4096" in hdr
    img = torch.zeros(64, dtype=torch.uint8, device="cuda")
    out = torch.zeros(4096, dtype=torch.int32, device="cuda")
    assert _call(lib, img, wmax + 1, 1, wmax + 1, out, out) == ug.ERR_CAPACITY           # beyond the scratch: rejected
    assert _call(lib, img, 1, hmax + 1, 1, out, out) == ug.ERR_CAPACITY
    assert _call(lib, img, 0, 4, 4, out, out) == ug.ERR_ARG
    assert _call(lib, img, 8, 4, 4, out, out) == ug.ERR_ARG                              # stride < w
    assert lib.ugd_integral_noalloc(0, None, 8, 8, C.c_size_t(8), None, None, None) == ug.ERR_ARG

    # the largest accepted size, with values that wrap the int32 planes
    rng = np.random.default_rng(7)
    g = rng.integers(200, 256, (hmax, wmax), dtype=np.uint8)
    dg = torch.from_numpy(g).cuda()
    S_t = torch.empty((hmax + 1, wmax + 1), dtype=torch.int32, device="cuda")
    SQ_t = torch.empty_like(S_t)
    torch.cuda.synchronize()
    assert _call(lib, dg, wmax, hmax, wmax, S_t, SQ_t, stream=torch.cuda.current_stream().cuda_stream) == 0
    torch.cuda.synchronize()
    nS, nSQ = _numpy_planes(g)
    assert np.array_equal(S_t.cpu().numpy(), nS) and np.array_equal(SQ_t.cpu().numpy(), nSQ)
