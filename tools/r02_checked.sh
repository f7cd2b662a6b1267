#!/bin/bash
# the scan kernel compiled with -DUGD_CHECKED=1 (device-side bounds checks on every tile / SQ-plane read, trap on a miss)
# under the decision-level parity tests: all sizes, all cascade shapes, TMA and fallback staging
cd /root/repo; mkdir -p gpurun_out
L=/root/repo/ugproj-detector_b200/build/libugdet_checked.so
ls -la $L
UGD_LIB=$L timeout 1500 python -m pytest tests -m gpu -x -q --deselect tests/test_shim.py > gpurun_out/r02_checked_build_pytest.log 2>&1; echo "checked build pytest rc=$?"
tail -3 gpurun_out/r02_checked_build_pytest.log
UGD_LIB=$L UGD_NO_TMA=1 timeout 900 python -m pytest tests -m gpu -x -q -k "decisions or stage_map_720p or tilted" > gpurun_out/r02_checked_build_notma_pytest.log 2>&1; echo "checked build, fallback staging: pytest rc=$?"
tail -3 gpurun_out/r02_checked_build_notma_pytest.log

# self-test of the checked build: a deliberately wild tile read (UGD_SCAN_KMIN=12345 arms it, checked builds only) must trap
UGD_LIB=$L UGD_SCAN_KMIN=12345 timeout 120 python - <<'PY' 2>&1 | tail -2
import sys, numpy as np
sys.path.insert(0, '.')
import __graft_entry__ as ge
ug = ge.load_package()
from oracle import cv2_ref
det = ug.Detector(ug.Cascade(cv2_ref.cascade_path('alt')), ug.default_params(), max_size=(320, 180), max_batch=1)
try:
    det.detect_gray(np.full((180, 320), 90, np.uint8) + (np.arange(320) % 50).astype(np.uint8)[None, :])
    print('SELF-TEST FAILED: the wild read was not caught')
except ug.UgdError as e:
    print('self-test ok: the wild read trapped ->', str(e)[:100])
PY
