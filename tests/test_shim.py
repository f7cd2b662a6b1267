"""The drop-in header include/ugdet_face_detector.hpp: compiled against stand-in cv / Configuration types
(tests/cxx/cv_stub.hpp) and driven by a C++ caller written like the reference's call sites
(face_tracker.cc:43, :295).  CPU: it compiles, links libugdet.so and fails loudly without a device.
GPU: rect lists equal the oracle's detectFaces, strided cv::Mat included, and detectFaces appends."""
import os
import subprocess

import numpy as np
import pytest

from conftest import rectset

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def shim_exe(ug, tmp_path_factory):
    ug.lib()                                                    # builds libugdet.so if needed
    out = str(tmp_path_factory.mktemp("shim") / "shim_main")
    libdir = os.path.dirname(ug.lib_path())
    cmd = ["g++", "-std=c++17", "-Wall", "-Wextra", "-Werror", "-I" + os.path.join(ROOT, "include"),
           "-I" + os.path.join(ROOT, "tests", "cxx"), os.path.join(ROOT, "tests", "cxx", "shim_main.cc"), "-o", out,
           "-L" + libdir, "-lugdet", "-Wl,-rpath," + libdir]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    return out


def test_header_is_plain_c(tmp_path):
    """include/ugdet.h must stay bindable from C (cgo / JNI / ctypes style FFI): compile it as C99."""
    src = tmp_path / "c.c"
    src.write_text('#include "ugdet.h"\nint main(void) { ugd_params p; ugd_rect r; (void)p; (void)r; return UGD_OK; }\n')
    r = subprocess.run(["gcc", "-std=c99", "-Wall", "-Werror", "-pedantic", "-I" + os.path.join(ROOT, "include"), "-c", str(src),
                        "-o", str(tmp_path / "c.o")], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr


def test_shim_compiles_and_fails_loudly_without_a_device(shim_exe, cv2ref, synth, tmp_path):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a device is present: the GPU test covers the run")
    f = tmp_path / "f.bgr"
    synth.bgr_frame(64, 48, 1, 0).tofile(f)
    r = subprocess.run([shim_exe, cv2ref.cascade_path("alt2"), str(f), "64", "48", str(64 * 3)], capture_output=True, text=True)
    assert r.returncode == 1 and "ugd_create" in r.stderr       # no CPU fallback behind the shim


@pytest.mark.gpu
def test_shim_matches_detect_faces(shim_exe, cv2ref, synth, orc, tmp_path):
    oc = orc.OracleCascade.load(cv2ref.cascade_path("alt2"))
    nonempty = 0
    for seed, pad in ((0, 0), (3, 7)):                          # pad: row step larger than w*3, like a cv::Mat ROI
        bgr = synth.bgr_frame(640, 360, 6, seed)
        step = 640 * 3 + pad
        buf = np.zeros((360, step), np.uint8)
        buf[:, :640 * 3] = bgr.reshape(360, -1)
        f = tmp_path / f"f{seed}.bgr"
        buf.tofile(f)
        r = subprocess.run([shim_exe, cv2ref.cascade_path("alt2"), str(f), "640", "360", str(step)], capture_output=True, text=True)
        assert r.returncode == 0, r.stderr
        lines = r.stdout.split("\n")
        n = int(lines[0])
        got = [tuple(map(int, l.split())) for l in lines[1:1 + n]]
        want, _ = orc.detect_faces(oc, bgr)
        assert rectset(got) == rectset(want)
        nonempty += n > 0
    assert nonempty > 0
