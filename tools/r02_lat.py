"""Where the time of a one-frame call goes: per-kernel device times (profiling mode) next to the wall-clock time of the call."""
import sys, time, statistics
import numpy as np
sys.path.insert(0, '.')
import __graft_entry__ as ge
import importlib
ug = ge.load_package()
synth = importlib.import_module(ge.PKG_NAME + ".synth")
from oracle import cv2_ref
for (w, h) in ((1280, 720), (640, 360)):
    det = ug.Detector(ug.Cascade(cv2_ref.cascade_path('alt')), ug.default_params(), max_size=(w, h), max_batch=1)
    g = synth.gray_frame(w, h, 8, 7)
    for _ in range(10): det.detect_gray(g)
    ts = []
    for _ in range(100):
        t0 = time.perf_counter(); det.detect_gray(g); ts.append((time.perf_counter() - t0) * 1e3)
    det.set_profiling(True)
    det.detect_gray(g); det.detect_gray(g)
    kt = det.kernel_times()
    det.set_profiling(False)
    print(w, h, 'wall median %.3f ms min %.3f' % (statistics.median(ts), min(ts)), {k: round(v[0], 4) for k, v in kt.items() if v[1]}, 'sum %.3f' % sum(v[0] for v in kt.values()))
    det.close()
