"""Small batches: time per call of detect_batch_device for n frames (device-resident, one launch group)."""
import sys, time, statistics, os
import numpy as np
sys.path.insert(0, '.')
import __graft_entry__ as ge
import importlib, torch
ug = ge.load_package()
synth = importlib.import_module(ge.PKG_NAME + ".synth")
from oracle import cv2_ref
w, h = 1280, 720
det = ug.Detector(ug.Cascade(cv2_ref.cascade_path('alt')), ug.default_params(), max_size=(w, h), max_batch=64)
fr = torch.from_numpy(synth.gray_batch(w, h, 64, seed0=0)).cuda()
out = []
for n in (1, 2, 4, 8, 16, 32, 64):
    for _ in range(5): det.detect_batch_device(fr.data_ptr(), False, n, w, h)
    ts = []
    for _ in range(30):
        t0 = time.perf_counter(); det.detect_batch_device(fr.data_ptr(), False, n, w, h); ts.append((time.perf_counter() - t0) * 1e3)
    out.append('%d: %.3f' % (n, statistics.median(ts)))
print('UGD_CHUNK_BELOW=%s' % os.environ.get('UGD_CHUNK_BELOW', 'default'), ' ms per call by frames:', '  '.join(out))
